// cbp_sv_kernels.cuh — single bound vector: the whole round loop of
// /root/reference/src/CbcBoundPropagation.cpp:107-195 as ONE persistent
// cooperative kernel.  Per round the matrix is streamed once: tiles of val/col
// are brought into shared memory by the TMA unit (cp.async.bulk + mbarrier,
// double buffered), the column bounds are gathered per nonzero, row
// activities are formed, and candidate bounds go out as atomic min/max on the
// *other* bound buffer (Jacobi).  Two grid-wide barriers per round (three when
// the matrix has rows longer than a tile); the host is not involved until the
// loop has stopped.
#pragma once
#include <cooperative_groups.h>
#include <math_constants.h>

#include "cbp_device.cuh"

namespace cbp {
namespace cg = cooperative_groups;

#ifndef CBP_SV_P2_TRIPS
#define CBP_SV_P2_TRIPS 1
#endif
#ifndef CBP_SV_P2_PACKED
#define CBP_SV_P2_PACKED 1  // second pass of the short rows: several rows per trip, lanes grouped by row (0: two rows per trip, one after the other)
#endif

constexpr int kP2Trips = CBP_SV_P2_TRIPS;  // trips of the packed second pass whose loads are in flight together

struct SvStatus {
  int rounds;
  int stopReason;
  int infeasRow;  // min original row index over violators (INT_MAX if none)
  int infeasCol;  // min crossing column (INT_MAX if none)
  int finalBuf;   // which of X[0], X[1] holds the answer
  int nChangedTotal;
  int timeUp;
  int chg[2];  // changed-column count of the running round (slot = round & 1)
  int sliceCtr[2];  // next short-row slice to hand out in the running round (slot = round & 1)
  int pad;
  unsigned long long t0;  // globaltimer at kernel start
  // row-sharded runs over peer memory: cross-GPU barriers passed when the kernel left (the next
  // run starts counting there), and a flag that a rank gave up waiting
  unsigned long long epochEnd;
  int p2pError;
  int abortRun;  // written by CTA 0 only: leave the kernel (a barrier timed out)
  unsigned long long goEpoch;  // written by CTA 0 only: the last cross-GPU barrier all ranks have reached
  unsigned long long ctaArrive;  // CTAs of this rank that have reached a cross-GPU barrier in this run
  // per round parity: every rank's first violated row (whole-matrix index; slot = rank, pushed by its
  // owner), and rank 0's "time is up"
  int inRowFlag[2][kMaxPeers];
  int inTime[2];
  // every rank's event-round limit (pushed by its owner after the start barrier): all ranks use the smallest
  int evMax[kMaxPeers];
  // where this rank's proposals go on every rank q: q's inbox slot for this rank (per round parity),
  // and q's status block
  Bnd *pushP[2][kMaxPeers];
  unsigned *pushD[2][kMaxPeers];
  struct SvStatus *peerSt[kMaxPeers];
  // constants of a row-sharded run, filled by the host: the exchange routines are separate functions
  // and fetch them here instead of taking them as arguments
  Bnd *inP[2];
  unsigned *inD[2];
  size_t strideP, strideD;
  int world, rank, nWords, pad3;
  Scal scal;
  // diagnostic: globaltimer seen by CTA 0 after the row phase, the long-row phase, the settle
  // phase and the decision of each of the first kStampRounds rounds (ns since t0)
  unsigned stamps[kStampRounds][4];
  unsigned stampsX[kStampRounds][4];  // CTA 0 only: after A1, after its own share of the rows (before the barrier)
  int roundChanged[kRoundCap];
  // row-sharded runs: arrivals of all ranks at the cross-GPU barrier.  Monotone over the life of the
  // block and never reset: a peer that is ahead may already have announced itself for the next run.
  unsigned long long arrive;
};

struct SvParams {
  // matrix (device)
  const double *val;
  const int *col;
  const Tile *tiles;
  const int *rp;
  const double *rowLB, *rowUB;
  const int *rowOrig;
  const LongRow *longRows;
  const signed char *colType;
  RowAct *longPart;  // scratch: one partial per long-row chunk
  int nTiles, firstLongTile, nLong, n;
  // short rows (<= kShortLen nonzeros) in sliced-ELL form: slice s holds rows
  // 32s .. 32s+31 (kept order), entry k of lane l at sellOff[s] + 32k + l
  const double *sellVal;
  const int *sellCol;
  const int *sellOff;  // nSlices + 1
  const int *sliceOrder;  // slices by decreasing length: the order in which they are handed out
  int nSlices, mShort;
  // column-major copy of the kept rows (which rows hold column j: cscRow[cscStart[j] .. cscStart[j+1]), kept-order
  // row indices) and the rows' packed CSR extents: event-driven rounds walk from the changed columns to their rows
  const int *cscStart;
  const int *cscRow;
  const int *rowNz, *rowCnt;
  int firstLongRow;      // kept rows from here on are longer than a tile (always scanned in full)
  int eventMaxChanged;   // a round that changed at most this many columns is followed by an event-driven round (0: never)
  // state
  Bnd *X[2];
  unsigned *dirty[2];
  int nWords;
  SvStatus *st;
  Scal s;
  int maxRounds;  // <=0: no limit
  // rows sharded over `world` GPUs.  inP[k] / inD[k] (k = round & 1) hold one slot per rank: n proposals
  // (strideP apart) and their mark words (strideD apart).  Accepted candidates go into this rank's OWN
  // slot; after the rows of a round every rank copies its marked proposals into its slot on every other
  // rank over NVLink; after the round's barrier every rank settles its own full copy of the bounds from
  // its (now local) slots.
  Bnd *inP[2];
  unsigned *inD[2];
  size_t strideP, strideD;
  // this rank's own slot and row flag per round parity, spelled out by the host: a plain parameter
  // can be re-read where it is needed, an address computed in the kernel would sit in registers
  // across the row loops
  Bnd *myP[2];
  unsigned *myD[2];
  int *myRowFlag[2];
  int world, rank, rowOffset;
  unsigned long long epoch0;  // cross-GPU barriers this group has passed in earlier runs
  int raw;        // 1: do the row phase of ONE round and stop (no R9, no bookkeeping): row-sharded mode
  unsigned long long budgetNs;  // time limit measured from kernel start
};

struct SvSmem {
  alignas(128) double val[2][kTileNnz];
  alignas(128) int col[2][kTileNnz];
  alignas(16) double2 lohi[kTileNnz];
  int rp[kTileRows + 1];
  alignas(8) uint64_t bar[2];
  RowAct red[kWarps];
  RowAct stash[kThreads];  // a long row's chunk partials, fetched one per thread
};

// publishes accepted candidates into the round's target buffer (row-sharded: the proposal buffer)
struct SvSink {
  Bnd *dst;
  unsigned *dirty;
  int *rowFlag;  // min over violated rows
  __device__ __forceinline__ void mark(int j) { red_or_b32(&dirty[j >> 5], 1u << (j & 31)); }
  __device__ __forceinline__ void ub(int j, double v)
  {
    atomic_min_f64(&dst[j].ub, v);
    mark(j);
  }
  __device__ __forceinline__ void lb(int j, double v)
  {
    atomic_max_f64(&dst[j].lb, v);
    mark(j);
  }
  __device__ __forceinline__ void flag_row(int row) { red_min_s32(rowFlag, row); }
};

__device__ __forceinline__ void sv_issue(const SvParams &p, SvSmem &sm, int t, int stage)
{
  const Tile T = p.tiles[t];
  const uint32_t cnt = (static_cast<uint32_t>(T.nzCount) + 3u) & ~3u;  // 16-byte multiples
  mbar_expect_tx(&sm.bar[stage], cnt * 12u);
  bulk_g2s(sm.val[stage], p.val + T.nzBegin, cnt * 8u, &sm.bar[stage]);
  bulk_g2s(sm.col[stage], p.col + T.nzBegin, cnt * 4u, &sm.bar[stage]);
}

// gather: one thread per staged nonzero; writes (lo, hi) = the bounds that
// enter minAct / maxAct for that nonzero (R3).  All loads of a thread are
// issued before the first use, so 8 gathers per thread are in flight.
__device__ __forceinline__ void sv_gather(const SvSmem &sm, double2 *lohi, int stage, int nz,
  const Bnd *__restrict__ src)
{
  constexpr int kPer = kTileNnz / kThreads;
  Bnd b[kPer];
#pragma unroll
  for (int i = 0; i < kPer; ++i) {
    const int k = threadIdx.x + i * kThreads;
    if (k < nz)
      b[i] = ldcg_bnd(&src[col_of(sm.col[stage][k])]);
  }
#pragma unroll
  for (int i = 0; i < kPer; ++i) {
    const int k = threadIdx.x + i * kThreads;
    if (k < nz) {
      const double a = sm.val[stage][k];
      lohi[k] = a > 0.0 ? make_double2(b[i].lb, b[i].ub) : make_double2(b[i].ub, b[i].lb);
    }
  }
}

__device__ __forceinline__ RowAct ldcg_rowact(const RowAct *q)
{
  const double2 a = __ldcg(reinterpret_cast<const double2 *>(q));
  const double2 b = __ldcg(reinterpret_cast<const double2 *>(q) + 1);
  RowAct r;
  r.minAct = a.x;
  r.maxAct = a.y;
  r.maxRange = b.x;
  const long long pk = __double_as_longlong(b.y);
  r.infMin = static_cast<int>(pk & 0xffffffffll);
  r.infMax = static_cast<int>(pk >> 32);
  return r;
}

// One pass over tiles [tBegin, tEnd) for this CTA.  phaseL=false: round phase A
// (short / medium rows completely, long-row chunks: partial sums only).
// phaseL=true: long-row chunks again, now with the whole row's activity known.
template <bool phaseL>
__device__ __forceinline__ void sv_stream(const SvParams &p, SvSmem &sm, uint32_t (&parity)[2],
  int tBegin, int tEnd, const Bnd *__restrict__ src, SvSink &sink)
{
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int t = tBegin + blockIdx.x;
  if (t < tEnd && tid == 0)
    sv_issue(p, sm, t, 0);
  for (int stage = 0; t < tEnd; t += gridDim.x, stage ^= 1) {
    const int tn = t + gridDim.x;
    if (tid == 0 && tn < tEnd)
      sv_issue(p, sm, tn, stage ^ 1);
    const Tile T = p.tiles[t];
    if (T.kind != kLong)
      for (int i = tid; i <= T.rowCount; i += kThreads)
        sm.rp[i] = p.rp[T.rpOff + i];
    mbar_wait(&sm.bar[stage], parity[stage]);
    parity[stage] ^= 1u;
    sv_gather(sm, sm.lohi, stage, T.nzCount, src);
    __syncthreads();
    const double *val = sm.val[stage];
    const int *col = sm.col[stage];
    if (T.kind == kMedium) {
      for (int r = warp; r < T.rowCount; r += kWarps) {
        const int kb = sm.rp[r], ke = sm.rp[r + 1];
        RowAct ra = { 0.0, 0.0, 0.0, 0, 0 };
        for (int k = kb + lane; k < ke; k += 32) {
          const double2 lh = sm.lohi[k];
          accumulate(ra, val[k], col_is_int(col[k]), lh.x, lh.y, p.s.inf);
        }
        warp_reduce(ra);
        const int row = T.rowBegin + r;
        const RowPlan pl = plan_row(ra, p.rowLB[row], p.rowUB[row], p.s);
        if (pl.infeasible && lane == 0)
          sink.flag_row(p.rowOrig[row] + p.rowOffset);
        if (pl.sideU || pl.sideL) {
          for (int k = kb + lane; k < ke; k += 32) {
            const double2 lh = sm.lohi[k];
            tighten_nz(sink, pl, val[k], col_of(col[k]), col_is_int(col[k]), lh.x, lh.y, p.s);
          }
        }
      }
    } else {  // one chunk of a long row
      const LongRow lr = p.longRows[T.longIdx];
      if (!phaseL) {
        RowAct ra = { 0.0, 0.0, 0.0, 0, 0 };
        for (int k = tid; k < T.nzCount; k += kThreads) {
          const double2 lh = sm.lohi[k];
          accumulate(ra, val[k], col_is_int(col[k]), lh.x, lh.y, p.s.inf);
        }
        warp_reduce(ra);
        if (lane == 0)
          sm.red[warp] = ra;
        __syncthreads();
        if (tid == 0) {
          RowAct tot = sm.red[0];
          for (int w = 1; w < kWarps; ++w)
            combine(tot, sm.red[w]);
          p.longPart[lr.firstPart + T.chunkIdx] = tot;
        }
      } else {
        // the row's partials, combined in chunk order by every thread alike;
        // they are fetched in parallel (one per thread) and summed from shared memory
        RowAct tot = { 0.0, 0.0, 0.0, 0, 0 };
        for (int c0 = 0; c0 < lr.nChunks; c0 += kThreads) {
          const int nc = min(kThreads, lr.nChunks - c0);
          if (tid < nc)
            sm.stash[tid] = ldcg_rowact(&p.longPart[lr.firstPart + c0 + tid]);
          __syncthreads();
          if (warp == 0)  // one warp walks them (all 256 threads would only queue up at the shared-memory port)
            for (int c = 0; c < nc; ++c)
              combine(tot, sm.stash[c]);
          __syncthreads();
        }
        if (tid == 0)
          sm.red[0] = tot;
        __syncthreads();
        tot = sm.red[0];
        const RowPlan pl = plan_row(tot, p.rowLB[lr.row], p.rowUB[lr.row], p.s);
        if (pl.infeasible && tid == 0 && T.chunkIdx == 0)
          sink.flag_row(p.rowOrig[lr.row] + p.rowOffset);
        if (pl.sideU || pl.sideL) {
          for (int k = tid; k < T.nzCount; k += kThreads) {
            const double2 lh = sm.lohi[k];
            tighten_nz(sink, pl, val[k], col_of(col[k]), col_is_int(col[k]), lh.x, lh.y, p.s);
          }
        }
      }
    }
    __syncthreads();  // everyone is done with this stage and with lohi
  }
}

// Short rows: one thread per row over the sliced-ELL copy.  The 32 lanes of a
// warp read 32 consecutive entries per step (coalesced), every lane walks ITS
// row in CSR order with the sums in registers — the order of the CPU
// restatement, so these rows agree with it bit for bit.  No shared memory, no
// block barrier; four gathers per lane are in flight.
// Slices are handed out through an atomic counter (a warp asks for its next
// slice before it starts the current one), so a round ends when the work
// ends, not when the unluckiest statically assigned warp does.
template <int kChunk>
__device__ __forceinline__ void sv_short_rows(const SvParams &p, const Bnd *__restrict__ src, SvSink &sink,
  int *ctr)
{
  const int lane = threadIdx.x & 31;
  int next = 0;
  if (lane == 0)
    next = atomicAdd(ctr, 1);
  next = __shfl_sync(0xffffffffu, next, 0);
  while (next < p.nSlices) {
    const int s = __ldg(&p.sliceOrder[next]);  // longest first, so the round's tail is made of the shortest
    if (lane == 0)
      next = atomicAdd(ctr, 1);
    const int off = __ldg(&p.sellOff[s]);
    const int len = (__ldg(&p.sellOff[s + 1]) - off) >> 5;
    const int row = (s << 5) + lane;
    const bool valid = row < p.mShort;
    const double rl = valid ? __ldg(&p.rowLB[row]) : 0.0, ru = valid ? __ldg(&p.rowUB[row]) : 0.0;
    const double *__restrict__ v = p.sellVal + off + lane;
    const int *__restrict__ c = p.sellCol + off + lane;
    RowAct ra = { 0.0, 0.0, 0.0, 0, 0 };
    for (int k = 0; k < len; k += kChunk) {
      double a[kChunk];
      int cj[kChunk];
      Bnd x[kChunk];
#pragma unroll
      for (int u = 0; u < kChunk; ++u) {
        a[u] = 0.0;
        cj[u] = 0;
        if (k + u < len) {  // uniform
          cj[u] = __ldg(&c[(k + u) << 5]);
          a[u] = __ldg(&v[(k + u) << 5]);
        }
      }
      unsigned isInt = 0u;  // the columns' integrality flags, so that the indices can die once the gathers are out
#pragma unroll
      for (int u = 0; u < kChunk; ++u) {  // padding points at column 0: a harmless, shared gather
        x[u] = ldcg_bnd(&src[col_of(cj[u])]);
        isInt |= col_is_int(cj[u]) ? 1u << u : 0u;
      }
#pragma unroll
      for (int u = 0; u < kChunk; ++u)  // padding has a == 0 and is skipped
        if (a[u] != 0.0)
          accumulate_fast(ra, a[u], (isInt >> u) & 1u, x[u], p.s.inf);
    }
    RowPlan pl = plan_row(ra, rl, ru, p.s);
    if (!valid) {
      pl.infeasible = false;
      pl.sideU = false;
      pl.sideL = false;
    }
    if (pl.infeasible)
      sink.flag_row(p.rowOrig[row] + p.rowOffset);
    // Second pass, one row at a time and nonzero-parallel: the lanes of the
    // warp take the (<= 32) nonzeros of each row that may tighten something;
    // that row's plan is broadcast by shuffle.  Rows that cannot tighten cost
    // nothing here, however many of them share the warp.
    unsigned todo = __ballot_sync(0xffffffffu, pl.sideU || pl.sideL);
    const int packed = min(pl.infMin, 2) | (min(pl.infMax, 2) << 2) | (pl.sideU ? 16 : 0) | (pl.sideL ? 32 : 0);
#if CBP_SV_P2_PACKED
    // Rows are short: a slice of length `len` gets 32 / len of its rows into one trip, the lanes cut into groups of
    // `len` (lane = group * len + entry), every group working for "its" row — the group-th of the rows still to do —
    // with that row's plan fetched by shuffle.  One copy of the tightening code serves them all.
    if (todo) {
      const int rowsPerTrip = 32 / len;  // (len <= kShortLen = 32, and a slice that wants anything has len >= 1)
      const int grp = lane / len, ent = lane - grp * len;
      while (todo) {
        // kP2Trips trips at a time: their loads are all issued before the first use
        int own[kP2Trips];
        bool have[kP2Trips];
        double a[kP2Trips];
        int cc[kP2Trips];
        Bnd x[kP2Trips];
#pragma unroll
        for (int t = 0; t < kP2Trips; ++t) {
          const unsigned ownU = grp < rowsPerTrip ? __fns(todo, 0, grp + 1) : 0xffffffffu;
          have[t] = ownU < 32u;
          own[t] = have[t] ? static_cast<int>(ownU) : 0;
          for (int r = 0; r < rowsPerTrip && todo; ++r)
            todo &= todo - 1;
          a[t] = 0.0;
          cc[t] = 0;
          if (have[t]) {
            const int at = off + (ent << 5) + own[t];
            cc[t] = __ldg(&p.sellCol[at]);
            a[t] = __ldg(&p.sellVal[at]);
          }
        }
#pragma unroll
        for (int t = 0; t < kP2Trips; ++t)
          x[t] = ldcg_bnd(&src[col_of(cc[t])]);
#pragma unroll
        for (int t = 0; t < kP2Trips; ++t) {
          RowPlan q;
          q.slackU = __shfl_sync(0xffffffffu, pl.slackU, own[t]);
          q.slackL = __shfl_sync(0xffffffffu, pl.slackL, own[t]);
          const int pk = __shfl_sync(0xffffffffu, packed, own[t]);
          q.infMin = pk & 3;
          q.infMax = (pk >> 2) & 3;
          q.sideU = (pk & 16) != 0;
          q.sideL = (pk & 32) != 0;
          q.infeasible = false;
          if (have[t] && a[t] != 0.0)
            tighten_nz(sink, q, a[t], col_of(cc[t]), col_is_int(cc[t]), a[t] > 0.0 ? x[t].lb : x[t].ub,
              a[t] > 0.0 ? x[t].ub : x[t].lb, p.s);
        }
      }
    }
#else
    while (todo) {
      // up to kSvPass2Rows rows per trip, all their loads issued before the first use
      int owner[kSvPass2Rows];
      double a4[kSvPass2Rows];
      int c4[kSvPass2Rows];
      Bnd x4[kSvPass2Rows];
#pragma unroll
      for (int t = 0; t < kSvPass2Rows; ++t) {
        owner[t] = todo ? __ffs(todo) - 1 : -1;
        todo &= todo - 1;
        a4[t] = 0.0;
        c4[t] = 0;
        if (owner[t] >= 0 && lane < len) {
          const int at = off + (lane << 5) + owner[t];
          c4[t] = __ldg(&p.sellCol[at]);
          a4[t] = __ldg(&p.sellVal[at]);
        }
      }
#pragma unroll
      for (int t = 0; t < kSvPass2Rows; ++t)
        x4[t] = ldcg_bnd(&src[col_of(c4[t])]);
#pragma unroll
      for (int t = 0; t < kSvPass2Rows; ++t) {
        if (owner[t] < 0)
          break;  // uniform
        RowPlan q;
        q.slackU = __shfl_sync(0xffffffffu, pl.slackU, owner[t]);
        q.slackL = __shfl_sync(0xffffffffu, pl.slackL, owner[t]);
        const int pk = __shfl_sync(0xffffffffu, packed, owner[t]);
        q.infMin = pk & 3;
        q.infMax = (pk >> 2) & 3;
        q.sideU = (pk & 16) != 0;
        q.sideL = (pk & 32) != 0;
        q.infeasible = false;
        if (a4[t] != 0.0)
          tighten_nz(sink, q, a4[t], col_of(c4[t]), col_is_int(c4[t]), a4[t] > 0.0 ? x4[t].lb : x4[t].ub,
            a4[t] > 0.0 ? x4[t].ub : x4[t].lb, p.s);
      }
    }
#endif
    next = __shfl_sync(0xffffffffu, next, 0);
  }
}

// One short row (<= kShortLen nonzeros) by one thread from the packed CSR copy: the same sums in the same order as
// the sliced-ELL pass (entries in CSR order, accumulate_fast), so a row gives the same plan whichever pass visits it.
__device__ __forceinline__ RowPlan sv_event_short_plan(const SvParams &p, const Bnd *__restrict__ src, SvSink &sink, int row)
{
  const int kb = __ldg(&p.rowNz[row]), cnt = __ldg(&p.rowCnt[row]);
  RowAct ra = { 0.0, 0.0, 0.0, 0, 0 };
  for (int k = 0; k < cnt; k += 8) {
    double a[8];
    int cj[8];
    Bnd x[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      a[u] = 0.0;
      cj[u] = 0;
      if (k + u < cnt) {
        cj[u] = __ldg(&p.col[kb + k + u]);
        a[u] = __ldg(&p.val[kb + k + u]);
      }
    }
#pragma unroll
    for (int u = 0; u < 8; ++u)
      x[u] = ldcg_bnd(&src[col_of(cj[u])]);
#pragma unroll
    for (int u = 0; u < 8; ++u)
      if (a[u] != 0.0)
        accumulate_fast(ra, a[u], col_is_int(cj[u]), x[u], p.s.inf);
  }
  const RowPlan pl = plan_row(ra, __ldg(&p.rowLB[row]), __ldg(&p.rowUB[row]), p.s);
  if (pl.infeasible)
    sink.flag_row(p.rowOrig[row] + p.rowOffset);
  return pl;
}

// The short rows of a warp's lanes that can tighten (want), nonzero-parallel: as many of them as fit into 32 lanes at
// a time — a row with cnt nonzeros gets cnt consecutive lanes — and each lane evaluates one nonzero of "its" row
// with that row's plan (fetched by shuffle).  One copy of the tightening code with all lanes busy, instead of every
// lane walking its own row through it.  Called by all 32 lanes together.
__device__ __forceinline__ void sv_tighten_packed(const SvParams &p, const Bnd *__restrict__ src, SvSink &sink, bool want,
  int row, const RowPlan &pl, int lane)
{
  unsigned todo = __ballot_sync(0xffffffffu, want);
  if (!todo)
    return;
  const int kb = want ? __ldg(&p.rowNz[row]) : 0, cnt = want ? __ldg(&p.rowCnt[row]) : 0;
  const int packed = min(pl.infMin, 2) | (min(pl.infMax, 2) << 2) | (pl.sideU ? 16 : 0) | (pl.sideL ? 32 : 0);
  while (todo) {
    int owner = -1, ent = 0, used = 0;
    while (todo) {
      const int o = __ffs(todo) - 1;
      const int c = __shfl_sync(0xffffffffu, cnt, o);
      if (used + c > 32 && used > 0)
        break;
      if (lane >= used && lane < used + c) {
        owner = o;
        ent = lane - used;
      }
      used += c;
      todo &= todo - 1;
    }
    const int from = owner >= 0 ? owner : 0;
    const int okb = __shfl_sync(0xffffffffu, kb, from);
    RowPlan q;
    q.slackU = __shfl_sync(0xffffffffu, pl.slackU, from);
    q.slackL = __shfl_sync(0xffffffffu, pl.slackL, from);
    const int pk = __shfl_sync(0xffffffffu, packed, from);
    q.infMin = pk & 3;
    q.infMax = (pk >> 2) & 3;
    q.sideU = (pk & 16) != 0;
    q.sideL = (pk & 32) != 0;
    q.infeasible = false;
    if (owner >= 0) {
      const int c = __ldg(&p.col[okb + ent]);
      const double a = __ldg(&p.val[okb + ent]);
      const Bnd x = ldcg_bnd(&src[col_of(c)]);
      if (a != 0.0)
        tighten_nz(sink, q, a, col_of(c), col_is_int(c), a > 0.0 ? x.lb : x.ub, a > 0.0 ? x.ub : x.lb, p.s);
    }
  }
}

// One medium row (kShortLen < nonzeros <= kTileNnz) by one warp: lanes stride over the row, butterfly — the
// order of the tile pass.
__device__ __forceinline__ void sv_event_medium_row(const SvParams &p, const Bnd *__restrict__ src, SvSink &sink, int row,
  int lane)
{
  const int kb = __ldg(&p.rowNz[row]), ke = kb + __ldg(&p.rowCnt[row]);
  RowAct ra = { 0.0, 0.0, 0.0, 0, 0 };
  for (int k = kb + lane; k < ke; k += 32) {
    const int c = __ldg(&p.col[k]);
    const double a = __ldg(&p.val[k]);
    const Bnd x = ldcg_bnd(&src[col_of(c)]);
    accumulate(ra, a, col_is_int(c), a > 0.0 ? x.lb : x.ub, a > 0.0 ? x.ub : x.lb, p.s.inf);
  }
  warp_reduce(ra);
  const RowPlan pl = plan_row(ra, __ldg(&p.rowLB[row]), __ldg(&p.rowUB[row]), p.s);
  if (pl.infeasible && lane == 0)
    sink.flag_row(p.rowOrig[row] + p.rowOffset);
  if (pl.sideU || pl.sideL)
    for (int k = kb + lane; k < ke; k += 32) {
      const int c = __ldg(&p.col[k]);
      const double a = __ldg(&p.val[k]);
      const Bnd x = ldcg_bnd(&src[col_of(c)]);
      tighten_nz(sink, pl, a, col_of(c), col_is_int(c), a > 0.0 ? x.lb : x.ub, a > 0.0 ? x.ub : x.lb, p.s);
    }
}

// Event-driven row phase: the previous round changed few columns, and a row none of whose columns moved would
// only repeat what it did last round (same bounds, same candidates, all of them already rejected or applied), so only
// the rows that hold a changed column are evaluated.  A warp takes a few words of the previous round's marks (as few
// as leaves every warp of the grid some: the phase is a chain of dependent loads per warp, not a throughput), brings
// the target buffer up to date for their marked columns (step A1), and then lays its marked columns end to end, 32
// at a time (lane = column: its range in the column-major copy), and those columns' rows end to end, 32 at a time
// (lane = row): short rows one per lane, medium rows one after the other by the whole warp.  So the number of trips
// follows the number of rows to visit, not the number of marked columns.  A row that holds two changed columns is
// evaluated twice; its candidates are max/min reductions, so that is harmless.  Rows longer than a tile are not
// handled here: the caller scans them in full.
__device__ __forceinline__ int sv_owner_of(int incl, int idx)
{
  // the first lane whose inclusive count exceeds idx (idx < the warp's total)
  int o = 0;
#pragma unroll
  for (int st = 16; st >= 1; st >>= 1)
    if (__shfl_sync(0xffffffffu, incl, o + st - 1) <= idx)
      o += st;
  return o;
}

__device__ __forceinline__ int sv_warp_incl(int v, int lane)
{
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, v, d);
    if (lane >= d)
      v += t;
  }
  return v;
}

__device__ __forceinline__ void sv_event_rows(const SvParams &p, const Bnd *__restrict__ src, Bnd *dst, unsigned *prev,
  SvSink &sink)
{
  const int lane = threadIdx.x & 31;
  const int gwarp = (blockIdx.x * kThreads + threadIdx.x) >> 5, nwarps = (gridDim.x * kThreads) >> 5;
  const int per = (p.nWords + nwarps - 1) / nwarps;
  const int wpw = per <= 1 ? 1 : (per >= 32 ? 32 : 1 << (32 - __clz(per - 1)));  // mark words per warp and trip
  for (int w0 = gwarp * wpw; w0 < p.nWords; w0 += nwarps * wpw) {
    const int wl = w0 + lane;
    const unsigned mine = (lane < wpw && wl < p.nWords) ? __ldcg(&prev[wl]) : 0u;
    if (mine) {
      prev[wl] = 0u;
      unsigned bits = mine;  // A1 for this lane's marked columns
      while (bits) {
        const int j = (wl << 5) + __ffs(bits) - 1;
        bits &= bits - 1;
        const Bnd b = ldcg_bnd(&src[j]);
        atomic_max_f64(&dst[j].lb, b.lb);
        atomic_min_f64(&dst[j].ub, b.ub);
      }
    }
    const int inclC = sv_warp_incl(__popc(mine), lane);
    const int nCols = __shfl_sync(0xffffffffu, inclC, 31);
    for (int cb = 0; cb < nCols; cb += 32) {  // (uniform)
      const int ci = cb + lane;
      const int o = sv_owner_of(inclC, min(ci, nCols - 1));
      const unsigned obits = __shfl_sync(0xffffffffu, mine, o);
      const int oexcl = __shfl_sync(0xffffffffu, inclC, o) - __popc(obits);
      int cs = 0, len = 0;
      if (ci < nCols) {
        const int j = ((w0 + o) << 5) + static_cast<int>(__fns(obits, 0, ci - oexcl + 1));
        cs = __ldg(&p.cscStart[j]);
        len = __ldg(&p.cscStart[j + 1]) - cs;
      }
      const int inclR = sv_warp_incl(len, lane);
      const int nRows = __shfl_sync(0xffffffffu, inclR, 31);
      for (int rb = 0; rb < nRows; rb += 32) {  // (uniform)
        const int ri = rb + lane;
        const int o2 = sv_owner_of(inclR, min(ri, nRows - 1));
        const int cs2 = __shfl_sync(0xffffffffu, cs, o2);
        const int ex2 = __shfl_sync(0xffffffffu, inclR, o2) - __shfl_sync(0xffffffffu, len, o2);
        const int row = ri < nRows ? __ldg(&p.cscRow[cs2 + (ri - ex2)]) : -1;
        RowPlan pl;
        pl.slackU = pl.slackL = 0.0;
        pl.infMin = pl.infMax = 0;
        pl.sideU = pl.sideL = pl.infeasible = false;
        const bool isShort = row >= 0 && row < p.mShort;
        if (isShort)
          pl = sv_event_short_plan(p, src, sink, row);
        sv_tighten_packed(p, src, sink, isShort && (pl.sideU || pl.sideL), row, pl, lane);
        unsigned med = __ballot_sync(0xffffffffu, row >= p.mShort && row < p.firstLongRow);
        while (med) {
          const int l = __ffs(med) - 1;
          med &= med - 1;
          sv_event_medium_row(p, src, sink, __shfl_sync(0xffffffffu, row, l), lane);
        }
      }
    }
  }
}

__device__ __forceinline__ void sv_stamp(const SvParams &p, int round, int which)
{
  if (blockIdx.x == 0 && threadIdx.x == 0 && round < kStampRounds) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    p.st->stamps[round][which] = static_cast<unsigned>(t - p.st->t0);
  }
}

// Cross-GPU barrier of a row-sharded run (no grid barrier inside: the CTAs count themselves in on a
// local word).  The last CTA of this rank to get here announces the rank on every rank's arrival
// counter; CTA 0 waits until all ranks have announced themselves `epoch` times and publishes its verdict
// on a local word, which the other CTAs wait for.  Only CTA 0 decides, so all CTAs of a rank leave the
// same way.  A rank that waits longer than its patience (a peer died) raises p2pError on every rank it
// can reach, and everybody leaves the kernel.  Threads that wrote to peer memory have fenced at system
// scope before they come here.
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *q)
{
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(q) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long ld_acquire_gpu(const unsigned long long *q)
{
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(q) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(unsigned long long *q, unsigned long long v)
{
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(q), "l"(v) : "memory");
}
__device__ __forceinline__ void red_release_sys_add(unsigned long long *q, unsigned long long v)
{
  asm volatile("red.release.sys.global.add.u64 [%0], %1;" ::"l"(q), "l"(v) : "memory");
}

// epoch: number of this barrier in the life of the group; nth: its number within this run (from 1).
// Not inlined (its registers must not weigh on the row loops), arguments by value.
__device__ __noinline__ bool sv_peer_barrier(SvStatus *st, unsigned long long epoch, unsigned long long nth,
  int world, unsigned long long patienceNs)
{
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned long long before = atomicAdd(&st->ctaArrive, 1ull);
    if (before + 1 == nth * gridDim.x) {  // the last CTA of this rank: everything this rank wrote is out
      __threadfence();
      for (int q = 0; q < world; ++q)
        red_release_sys_add(&st->peerSt[q]->arrive, 1ull);
    }
    if (blockIdx.x == 0) {
      const unsigned long long want = epoch * static_cast<unsigned long long>(world);
      unsigned long long t0, t;
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
      volatile int *err = &st->p2pError;
      bool ok = true;
      while (ld_acquire_sys(&st->arrive) < want) {
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
        if (*err || t - t0 > patienceNs) {
          for (int q = 0; q < world; ++q)
            atomicExch_system(&st->peerSt[q]->p2pError, 1);
          ok = false;
          break;
        }
      }
      if (ok)
        st_release_gpu(&st->goEpoch, epoch);
      else
        *reinterpret_cast<volatile int *>(&st->abortRun) = 1;
    } else {
      volatile int *ab = &st->abortRun;
      while (ld_acquire_gpu(&st->goEpoch) < epoch && !*ab) {
      }
    }
  }
  __syncthreads();
  return *reinterpret_cast<volatile int *>(&st->abortRun) == 0;
}

// Row-sharded, after the rows of a round: copy this rank's marked proposals (and the mark words, the
// row flag, rank 0's clock verdict) into this rank's slot on every other rank.  Plain stores over
// NVLink, nothing waits for an answer; the system fence at the end makes them visible before this
// thread reaches the barrier that announces them.  One thread per word of 32 columns.
// (Not inlined, constants fetched from the status block: its registers must not weigh on the row loops
// of the kernel.)
struct SvPeerArgs {
  Bnd *inP;       // this round's slots
  unsigned *inD;
  Bnd *oldP;      // this rank's own slot of the previous round
  unsigned *oldD;
  size_t strideP, strideD;
  int world, rank, nWords, k;
};

__device__ __forceinline__ SvPeerArgs sv_peer_args(const SvStatus *st, int k)
{
  SvPeerArgs a;
  a.strideP = st->strideP;
  a.strideD = st->strideD;
  a.world = __ldg(&st->world);
  a.rank = __ldg(&st->rank);
  a.nWords = __ldg(&st->nWords);
  a.k = k;
  a.inP = st->inP[k];  // (plain loads: written by the host before the launch, never by the kernel)
  a.inD = st->inD[k];
  a.oldP = st->inP[k ^ 1] + a.rank * a.strideP;
  a.oldD = st->inD[k ^ 1] + a.rank * a.strideD;
  return a;
}

__device__ __noinline__ void sv_peer_push(SvStatus *st, int k, int gtid, int gsize)
{
  const SvPeerArgs p = sv_peer_args(st, k);
  const Bnd *myP = p.inP + p.rank * p.strideP;
  const unsigned *myD = p.inD + p.rank * p.strideD;
  bool wrote = false;
  for (int w = gtid; w < p.nWords; w += gsize) {
    unsigned bits = __ldcg(&myD[w]);
    if (!bits)
      continue;
    wrote = true;
#pragma unroll
    for (int q = 0; q < kMaxPeers; ++q)
      if (q < p.world && q != p.rank)
        st->pushD[k][q][w] = bits;
    while (bits) {  // four columns per trip: their loads are issued together
      int j4[4];
      Bnd b4[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        j4[u] = bits ? (w << 5) + __ffs(bits) - 1 : -1;
        bits &= bits - 1;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (j4[u] >= 0)
          b4[u] = ldcg_bnd(&myP[j4[u]]);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (j4[u] < 0)
          break;
#pragma unroll
        for (int q = 0; q < kMaxPeers; ++q)
          if (q < p.world && q != p.rank)
            st->pushP[k][q][j4[u]] = b4[u];
      }
    }
  }
  if (gtid == 0) {
    const int rf = __ldcg(&st->inRowFlag[k][p.rank]);
    const int tf = p.rank == 0 ? __ldcg(&st->inTime[k]) : 0;
    for (int q = 0; q < p.world; ++q)
      if (q != p.rank) {
        if (rf != 0x7fffffff)
          st->peerSt[q]->inRowFlag[k][p.rank] = rf;
        if (tf)
          st->peerSt[q]->inTime[k] = 1;
      }
    wrote = true;
  }
  if (wrote)
    __threadfence_system();
}

// Row-sharded settle: every rank combines all ranks' proposals of this round for the marked columns,
// applies R9 to its own full copy of the bounds and counts the changed columns — the same work on the
// same data on every rank, so all ranks take the same decision without a second exchange.  One thread
// per word of 32 columns.  Also withdraws this rank's own proposals of the PREVIOUS round and clears the
// other ranks' marks of this one (nobody writes those slots again before the round after next, and
// nobody can get there without this rank reaching the next barrier first).
__device__ __noinline__ int sv_peer_settle(SvStatus *st, int k, const Bnd *__restrict__ src, Bnd *dst,
  unsigned *dirtyOut, int gtid, int gsize)
{
  const SvPeerArgs p = sv_peer_args(st, k);
  Scal sc;
  sc.inf = __ldg(&st->scal.inf);
  sc.tol = __ldg(&st->scal.tol);
  sc.feasTol = __ldg(&st->scal.feasTol);
  int mine = 0;
  Bnd *oldP = p.oldP;
  unsigned *oldD = p.oldD;
  for (int w = gtid; w < p.nWords; w += gsize) {
    unsigned bq[kMaxPeers];
    unsigned any = 0u;
#pragma unroll
    for (int q = 0; q < kMaxPeers; ++q) {
      bq[q] = q < p.world ? __ldcg(&p.inD[q * p.strideD + w]) : 0u;
      any |= bq[q];
    }
    const unsigned old = __ldcg(&oldD[w]);
    if (old) {
      oldD[w] = 0u;
      unsigned bits = old;
      while (bits) {
        const int j = (w << 5) + __ffs(bits) - 1;
        bits &= bits - 1;
        Bnd none;
        none.lb = -CUDART_INF;
        none.ub = CUDART_INF;
        oldP[j] = none;
      }
    }
    if (!any)
      continue;
#pragma unroll
    for (int q = 0; q < kMaxPeers; ++q)
      if (q < p.world && q != p.rank && bq[q])
        p.inD[q * p.strideD + w] = 0u;
    dirtyOut[w] = any;
    mine += __popc(any);
    unsigned bits = any;
    while (bits) {
      const int t = __ffs(bits) - 1;
      const int j = (w << 5) + t;
      bits &= bits - 1;
      Bnd b = ldcg_bnd(&src[j]);
      const double oldUb = b.ub;
      Bnd c[kMaxPeers];
#pragma unroll
      for (int q = 0; q < kMaxPeers; ++q)  // all slots' loads are issued before the first use
        if (q < p.world && ((bq[q] >> t) & 1u))
          c[q] = ldcg_bnd(&p.inP[q * p.strideP + j]);
#pragma unroll
      for (int q = 0; q < kMaxPeers; ++q)
        if (q < p.world && ((bq[q] >> t) & 1u)) {
          b.lb = fmax(b.lb, c[q].lb);
          b.ub = fmin(b.ub, c[q].ub);
        }
      if (settle(b.lb, b.ub, oldUb, sc))
        red_min_s32(&st->infeasCol, j);
      else
        dst[j] = b;
    }
  }
  // the flags of the round: every rank's first violated row, rank 0's clock
  if (gtid == 0) {
    int rf = 0x7fffffff;
    for (int q = 0; q < p.world; ++q)
      rf = min(rf, __ldcg(&st->inRowFlag[k][q]));
    if (rf != 0x7fffffff)
      st->infeasRow = rf;
    if (__ldcg(&st->inTime[k]))
      st->timeUp = 1;
  }
  return mine;
}

// kTiles=false is the variant for matrices whose rows are all short: no
// shared-memory staging, fewer registers, more resident warps.
template <bool kTiles, bool kP2P>
__global__ void __launch_bounds__(kThreads, kP2P ? kSvShardedBlocks : (kTiles ? 3 : kSvNoTileBlocks)) sv_fixpoint_kernel(const SvParams p)
{
  extern __shared__ __align__(128) unsigned char smem_raw[];
  SvSmem &sm = *reinterpret_cast<SvSmem *>(smem_raw);  // kTiles=false: launched without it, never touched
  __shared__ int s_cnt[kWarps];
  __shared__ int s_dec[5];
  cg::grid_group grid = cg::this_grid();
  const int tid = threadIdx.x;
  const int gtid = blockIdx.x * kThreads + tid, gsize = gridDim.x * kThreads;
  uint32_t parity[2] = { 0u, 0u };
  if (tid == 0) {
    if (kTiles) {
      mbar_init(&sm.bar[0], 1);
      mbar_init(&sm.bar[1], 1);
      mbar_fence_init();
    }
    if (blockIdx.x == 0) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
      p.st->t0 = t;
    }
  }
  __syncthreads();
  // R11 failed in the prepare kernel: infeasible before the first round
  if (__ldcg(&p.st->infeasCol) != 0x7fffffff) {
    if (gtid == 0) {
      p.st->rounds = 1;
      p.st->stopReason = 4;
      p.st->finalBuf = 0;
      p.st->epochEnd = p.epoch0;  // (sharded: every rank takes this exit, nobody waits for anybody)
    }
    return;
  }

  const int limit = p.maxRounds > 0 ? p.maxRounds : 0x7fffffff;
  // sharded runs pass one cross-GPU barrier at the start and one per round: barrier number
  // p.epoch0 + 1 at the start, p.epoch0 + 2 + round in a round
  // sharded: nobody reads a peer's block before that peer has reset it for this run (its kernel
  // follows its prepare kernels in stream order, so its arrival here says so).  Ranks reach this point
  // at host speed, hence the long patience.
  if (kP2P && !sv_peer_barrier(p.st, p.epoch0 + 1, 1ull, p.world, 30000000000ull)) {
    if (gtid == 0)
      p.st->epochEnd = p.epoch0 + 1;
    return;
  }
  if (kP2P && gtid == 0) {  // (after the start barrier: every rank has reset its block for this run)
    for (int q = 0; q < p.world; ++q)
      p.st->peerSt[q]->evMax[p.rank] = p.eventMaxChanged;
    __threadfence_system();
  }
  bool sparse = false;  // last round changed few columns: scan the marks word-wise
  bool event = false;   // ... so few that this round only visits the rows that hold one of them
  int round = 0;
  for (;; ++round) {
    const int cur = round & 1;
    const Bnd *src = p.X[cur];
    Bnd *dst = p.X[cur ^ 1];
#ifdef CBP_FINE_STAMPS
    if (blockIdx.x == 0 && tid == 0 && round < kStampRounds) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
      p.st->stampsX[round][2] = static_cast<unsigned>(t - p.st->t0);
    }
#endif
    // not sharded: candidates go straight into the target buffer; sharded: into this round's proposals
    SvSink sink = { kP2P ? p.myP[cur] : dst, kP2P ? p.myD[cur] : p.dirty[cur],
      kP2P ? p.myRowFlag[cur] : &p.st->infeasRow };

    // A1: bring the target buffer up to date at the columns the previous
    // round changed (it was that round's source, so it is stale there), and
    // clear that round's marks.  Atomics: rows of this round may hit the same
    // column at the same time.
    if (event) {
      // (A1 happens inside the event-driven row phase below, column by column)
    } else if (round > 0 && sparse) {
      // few columns changed last round: one mark word per thread, its few bits one after the other
      unsigned *prev = p.dirty[cur ^ 1];
      for (int w = gtid; w < p.nWords; w += gsize) {
        unsigned bits = __ldcg(&prev[w]);
        if (!bits)
          continue;
        prev[w] = 0u;
        while (bits) {
          const int j = (w << 5) + __ffs(bits) - 1;
          bits &= bits - 1;
          const Bnd b = ldcg_bnd(&src[j]);
          atomic_max_f64(&dst[j].lb, b.lb);
          atomic_min_f64(&dst[j].ub, b.ub);
        }
      }
    } else if (round > 0) {
      unsigned *prev = p.dirty[cur ^ 1];
      for (int j0 = gtid & ~31; j0 < p.n; j0 += gsize) {  // one column per thread, one word per warp
        const int j = j0 + (tid & 31);
        const unsigned bits = __ldcg(&prev[j0 >> 5]);
        if ((bits >> (tid & 31)) & 1u) {
          const Bnd b = ldcg_bnd(&src[j]);
          atomic_max_f64(&dst[j].lb, b.lb);
          atomic_min_f64(&dst[j].ub, b.ub);
        }
        __syncwarp();
        if ((tid & 31) == 0 && bits)
          prev[j0 >> 5] = 0u;
      }
    }
    if (blockIdx.x == 0 && tid == 0 && round < kStampRounds) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
      p.st->stampsX[round][0] = static_cast<unsigned>(t - p.st->t0);
    }
    // A2: stream the matrix: short rows from the sliced-ELL copy, the rest as TMA-staged tiles
    if (event)
      sv_event_rows(p, src, dst, p.dirty[cur ^ 1], sink);
    else
      sv_short_rows<kP2P ? kSvChunkSharded : kSvChunk>(p, src, sink, &p.st->sliceCtr[cur]);
    if (blockIdx.x == 0 && tid == 0 && round < kStampRounds) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
      p.st->stampsX[round][1] = static_cast<unsigned>(t - p.st->t0);
    }
    if (kTiles)  // (event rounds: the medium rows were visited from their columns; rows longer than a tile are scanned in full)
      sv_stream<false>(p, sm, parity, event ? p.firstLongTile : 0, p.nTiles, src, sink);
    if (kP2P) {
      if (kTiles && p.nLong > 0) {  // (long-row partials are rank-local: a row lives on one rank)
        grid.sync();
        sv_stream<true>(p, sm, parity, p.firstLongTile, p.nTiles, src, sink);
      }
      if (p.rank == 0 && gtid == 0) {  // rank 0's clock decides for everybody
        unsigned long long t;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
        if (t - p.st->t0 >= p.budgetNs)
          p.st->inTime[cur] = 1;
      }
      grid.sync();  // all proposals of this rank are in its own slot
      sv_peer_push(p.st, cur, gtid, gsize);
      if (!sv_peer_barrier(p.st, p.epoch0 + 2 + round, 2ull + round, p.world, 4000000000ull))
        break;
      sv_stamp(p, round, 0);
      sv_stamp(p, round, 1);
    } else {
      grid.sync();
      sv_stamp(p, round, 0);
      if (kTiles && p.nLong > 0) {
        sv_stream<true>(p, sm, parity, p.firstLongTile, p.nTiles, src, sink);
        grid.sync();
      }
      sv_stamp(p, round, 1);
    }
    if (p.raw) {  // the caller combines the proposals of all row blocks and settles them itself
      if (gtid == 0) {
        p.st->rounds = 1;
        p.st->stopReason = __ldcg(&p.st->infeasRow) != 0x7fffffff ? 4 : 2;
        p.st->finalBuf = cur ^ 1;
      }
      return;
    }
    // B: settle the changed columns (R9) and count them
    {
      int mine = 0;
      const unsigned *dw = p.dirty[cur];
      if (kP2P) {
        mine = sv_peer_settle(p.st, cur, src, dst, p.dirty[cur], gtid, gsize);
      } else if (sparse) {  // (the previous round's count is the predictor; either path is correct for any count)
        for (int w = gtid; w < p.nWords; w += gsize) {
          unsigned bits = __ldcg(&dw[w]);
          mine += __popc(bits);
          while (bits) {
            const int j = (w << 5) + __ffs(bits) - 1;
            bits &= bits - 1;
            Bnd b = ldcg_bnd(&dst[j]);
            const double oldUb = __ldcg(&src[j].ub);
            const double L0 = b.lb, U0 = b.ub;
            if (settle(b.lb, b.ub, oldUb, p.s))
              atomicMin(&p.st->infeasCol, j);
            else if (b.lb != L0 || b.ub != U0)
              dst[j] = b;
          }
        }
      } else
      for (int j0 = gtid & ~31; j0 < p.n; j0 += gsize) {  // one column per thread, one word per warp
        const int j = j0 + (tid & 31);
        const unsigned bits = __ldcg(&dw[j0 >> 5]);
        if ((bits >> (tid & 31)) & 1u) {
          ++mine;
          Bnd b = ldcg_bnd(&dst[j]);
          const double oldUb = __ldcg(&src[j].ub);
          const double L0 = b.lb, U0 = b.ub;
          if (settle(b.lb, b.ub, oldUb, p.s))
            atomicMin(&p.st->infeasCol, j);
          else if (b.lb != L0 || b.ub != U0)
            dst[j] = b;
        }
      }
#pragma unroll
      for (int m = 16; m >= 1; m >>= 1)
        mine += __shfl_xor_sync(0xffffffffu, mine, m);
      if ((tid & 31) == 0)
        s_cnt[tid >> 5] = mine;
      __syncthreads();
      if (tid == 0) {
        int tot = 0;
        for (int w = 0; w < kWarps; ++w)
          tot += s_cnt[w];
        if (tot)
          atomicAdd(&p.st->chg[cur], tot);
        if (blockIdx.x == 0) {
          p.st->chg[cur ^ 1] = 0;  // next round's slots; nobody touches them in this phase
          p.st->sliceCtr[cur ^ 1] = 0;
          unsigned long long t;
          asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
          if (!kP2P && t - p.st->t0 >= p.budgetNs)
            p.st->timeUp = 1;
        }
      }
    }
    grid.sync();  // (sharded too: every rank settled the same data, nothing more to exchange)
    sv_stamp(p, round, 2);
    // every thread takes the same decision from the same global words (fetched once per CTA: the
    // whole grid asking one L2 line for them costs microseconds)
    if (tid == 0) {
      s_dec[0] = __ldcg(&p.st->infeasRow);
      s_dec[1] = __ldcg(&p.st->infeasCol);
      s_dec[2] = __ldcg(&p.st->chg[cur]);
      s_dec[3] = __ldcg(&p.st->timeUp);
      int ev = p.eventMaxChanged;
      if (kP2P)  // (pushed before this round's cross-GPU barrier, the same values on every rank)
        for (int q = 0; q < p.world; ++q)
          ev = min(ev, __ldcg(&p.st->evMax[q]));
      s_dec[4] = ev;
    }
    __syncthreads();
    const int badRow = s_dec[0], badCol = s_dec[1];
    const int changed = s_dec[2];
    const int timeUp = s_dec[3];
    int stop = 0;
    int fin = cur;
    if (badRow != 0x7fffffff || badCol != 0x7fffffff)
      stop = 4;  // InfeasibleDetected: the failing round is not applied
    else if (changed == 0)
      stop = 1;  // ReachedFixpoint
    else if (round + 1 >= limit) {
      stop = 2;  // HitMaxRounds: the last round is applied
      fin = cur ^ 1;
    } else if (timeUp) {
      stop = 3;  // HitTimeLimit, checked before the next round would start
      fin = cur ^ 1;
    }
    sv_stamp(p, round, 3);
    if (gtid == 0) {
      if (stop != 4) {  // a failing round is not applied and not reported
        if (round < kRoundCap)
          p.st->roundChanged[round] = changed;
        p.st->nChangedTotal += changed;
      }
    }
    sparse = changed < (p.nWords >> 2);
    event = changed <= s_dec[4];
    if (stop) {
      if (gtid == 0) {
        p.st->rounds = round + 1;
        p.st->stopReason = stop;
        p.st->finalBuf = fin;
      }
      break;
    }
  }
  if (kP2P && gtid == 0)
    p.st->epochEnd = p.epoch0 + 2 + round;
}

// ---- small helper kernels ---------------------------------------------------

// row-sharded runs: empty proposal buffers (no proposal = (-inf, +inf))
__global__ void sv_no_proposals_kernel(Bnd *P, size_t count)
{
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < count) {
    Bnd none;
    none.lb = -CUDART_INF;
    none.ub = CUDART_INF;
    P[i] = none;
  }
}

// interleave lb/ub into both state buffers, apply R11, mark what it changed
__global__ void sv_prepare_kernel(const double *__restrict__ lb, const double *__restrict__ ub,
  const signed char *__restrict__ colType, int n, Bnd *X0, Bnd *X1, unsigned *dirty0, SvStatus *st,
  Scal s, int doPrepass)
{
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n)
    return;
  double l = lb[j] + 0.0, u = ub[j] + 0.0;  // (no -0.0 in the bound arrays: the integer-ordered atomics rely on it)
  const int r = doPrepass ? prepass_col(l, u, colType[j] != 0, s) : 0;
  if (r == 2)
    atomicMin(&st->infeasCol, j);
  else if (r == 1)
    atomicOr(&dirty0[j >> 5], 1u << (j & 31));
  Bnd b;
  b.lb = l;
  b.ub = u;
  X0[j] = b;
  X1[j] = b;
}

__global__ void sv_finish_kernel(const Bnd *__restrict__ X, int n, double *lb, double *ub)
{
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n)
    return;
  const Bnd b = X[j];
  lb[j] = b.lb;
  ub[j] = b.ub;
}

// Phase 1 (singleton rows): one thread per row that has exactly one nonzero.
// Candidates are rowUB/a and rowLB/a; acceptance as in R7/R8; Jacobi.
__global__ void sv_singleton_kernel(const int *__restrict__ singRows, int nSing,
  const int *__restrict__ rowNz, const double *__restrict__ val, const int *__restrict__ col,
  const double *__restrict__ rowLB, const double *__restrict__ rowUB,
  const signed char *__restrict__ colType, const Bnd *__restrict__ src, Bnd *dst, unsigned *dirty,
  Scal s)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nSing)
    return;
  const int row = singRows[i];
  const int k = rowNz[row];
  const double a = val[k];
  const int j = col_of(col[k]);
  const Bnd b = src[j];
  const bool isInt = col_is_int(col[k]);
  SvSink sink = { dst, dirty };
  const double ru = rowUB[row], rl = rowLB[row];
  if (ru < s.inf) {
    if (a > 0.0)
      offer_ub(sink, j, ru / a, b.ub, b.lb, isInt, s);
    else
      offer_lb(sink, j, ru / a, b.lb, b.ub, isInt, s);
  }
  if (rl > -s.inf) {
    if (a > 0.0)
      offer_lb(sink, j, rl / a, b.lb, b.ub, isInt, s);
    else
      offer_ub(sink, j, rl / a, b.ub, b.lb, isInt, s);
  }
}

// settle + count for the singleton phase; res[0]=changed, res[1]=fixed, res[2]=crossing
__global__ void sv_singleton_settle_kernel(const Bnd *__restrict__ src, Bnd *dst, unsigned *dirty,
  int nWords, int *res, Scal s)
{
  const int w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= nWords)
    return;
  unsigned bits = dirty[w];
  if (!bits)
    return;
  dirty[w] = 0u;
  while (bits) {
    const int j = (w << 5) + __ffs(bits) - 1;
    bits &= bits - 1;
    Bnd b = dst[j];
    if (settle(b.lb, b.ub, src[j].ub, s)) {
      atomicAdd(&res[2], 1);
      continue;
    }
    dst[j] = b;
    atomicAdd(&res[0], 1);
    if (b.lb == b.ub)
      atomicAdd(&res[1], 1);
  }
}

// plain activity kernel (diagnostic API): one warp per kept row, lanes stride
__global__ void row_activity_kernel(const int *__restrict__ rowNz, const int *__restrict__ rowCnt,
  const int *__restrict__ rowOrig, int mKept, const double *__restrict__ val,
  const int *__restrict__ col, const double *__restrict__ lb, const double *__restrict__ ub,
  double INF, double *minAct, double *maxAct, int *infMin, int *infMax)
{
  const int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (r >= mKept)
    return;
  const int kb = rowNz[r], ke = kb + rowCnt[r];
  RowAct ra = { 0.0, 0.0, 0.0, 0, 0 };
  if (rowCnt[r] <= kShortLen) {  // sequential order, as the oracle sums
    if (lane == 0)
      for (int k = kb; k < ke; ++k) {
        const double a = val[k];
        const int j = col_of(col[k]);
        accumulate(ra, a, true, a > 0.0 ? lb[j] : ub[j], a > 0.0 ? ub[j] : lb[j], INF);
      }
  } else {
    for (int k = kb + lane; k < ke; k += 32) {
      const double a = val[k];
      const int j = col_of(col[k]);
      accumulate(ra, a, true, a > 0.0 ? lb[j] : ub[j], a > 0.0 ? ub[j] : lb[j], INF);
    }
    warp_reduce(ra);
  }
  if (lane == 0) {
    const int o = rowOrig[r];
    minAct[o] = ra.minAct;
    maxAct[o] = ra.maxAct;
    infMin[o] = ra.infMin;
    infMax[o] = ra.infMax;
  }
}

}  // namespace cbp
