// cbp_batch_kernels.cuh — B bound vectors over one matrix.
//
// Layout in HBM ("slice-major"): the vectors are cut into slices of W (32, 16, 8 or 4, chosen
// when the batch is loaded); inside a slice the bounds are [column][vector]:
//     cell(j, b) = ((b / W) * n + j) * W + b % W,      each cell a Bnd = {lb, ub} (16 B).
// The W lanes that own a slice read the W vectors of one column as one contiguous W x 16-byte
// piece (512 B for W = 32), consecutive columns are adjacent, and everything that runs at one
// time touches one or two slices — a compact range for L2 and the TLB.  The marks (dirty /
// ever) are one 32-bit word per (32-vector group, column), group-major.
// A warp owns rows x vectors; the matrix entries are shared by its lanes while every lane
// accumulates its own vector's activities sequentially in CSR order — bit for bit the order of
// the CPU restatement.  Two bound buffers give Jacobi rounds; accepted candidates are atomic
// min/max on the other buffer plus a mark per (column, vector).
//
// Two forms of the bounds.  Classic: both buffers hold every cell (dense batches, bases that need R11).
// Overlay (base + diffs batches): baseX[n] holds the shared base bounds, dmask[G][n] says which
// vectors of a group have a cell of their own at a column, and only those cells of the two buffers are
// valid; a round reads "base unless masked", collects its candidates in a list, creates the cells the
// list touches for the first time, applies the list, and settles / carries on the marks as in the
// classic form.  The overlay kernels carry several vector groups per warp: what they read per
// nonzero comes from one shared-memory entry staged by one lane.
#pragma once
#include "cbp_device.cuh"

namespace cbp {

// one accepted candidate of round 0: column (bit 31 = upper bound), vector, value
struct __align__(16) BtLogEntry {
  unsigned var;
  int b;
  double v;
};

struct BtParams {
  const double *val;
  const int *col;
  const int *rowNz;   // packed offset of each kept row
  const int *rowCnt;  // nonzeros of each kept row
  const int *rowPos;  // rows longer than kBtLongLen: positive coefficients come first, this many
  const double *rowLB, *rowUB;
  const int *rowOrig;
  const signed char *colType;
  const int *shortRowsBt, *longRowsBt;  // kept-row indices by batch class
  int nShortBt, nLongBt;
  int mKept, n, Bp, G;  // Bp = padded vector count, G = Bp / 32
  int W;                // vectors per slice
  Bnd *X[2];
  unsigned *dirty[2];   // [G][n]
  unsigned *ever;       // [G][n] union of the marks of applied rounds
  unsigned *active;     // [G]
  int *vecInfRow, *vecInfCol;  // [Bp] min violating row / crossing column (INT_MAX none)
  int *vecChanged;             // [Bp] changed columns in the running round
  int *vecRounds, *vecTotal, *vecStop, *vecFinal;  // [Bp]
  int *anyActive;
  // round 0 of a base+deltas batch: accepted candidates go to this list instead of the second
  // bound buffer, which then is only built for the vector groups that need it
  struct BtLogEntry *log;
  unsigned *logCount;
  unsigned logCap;
  int *groupFlag;  // [G] this group has candidates or R11 changes
  // overlay form of a base+diffs batch: a vector's bounds are the shared base bounds except at the columns
  // its diffs or accepted candidates have touched; those cells alone are valid in X[0] / X[1]
  const Bnd *baseX;       // [n]
  const unsigned *dmask;  // [G][n] which vectors of the group differ from the base at column j
  Scal s;
};

// first cell (column 0) of vector b; column j is W cells further per column
__device__ __forceinline__ size_t bt_base(const BtParams &p, int b)
{
  return (static_cast<size_t>(b / p.W) * p.n) * p.W + (b % p.W);
}
// the same for the kernels that only exist for W = 32 (no runtime division)
__device__ __forceinline__ size_t bt_base32(const BtParams &p, int b)
{
  return (static_cast<size_t>(b >> 5) * p.n) * 32 + (b & 31);
}
__device__ __forceinline__ size_t bt_cell(const BtParams &p, int j, int b)
{
  return bt_base(p, b) + static_cast<size_t>(j) * p.W;
}
__device__ __forceinline__ size_t bt_word(const BtParams &p, int j, int g)
{
  return static_cast<size_t>(g) * p.n + j;
}

struct BtSink {
  Bnd *dst;         // already at this lane's vector, column 0
  unsigned *dirty;  // already at this warp's group, column 0
  int W;
  unsigned bit;
  __device__ __forceinline__ void ub(int j, double v)
  {
    atomic_min_f64(&dst[static_cast<size_t>(j) * W].ub, v);
    atomicOr(&dirty[j], bit);
  }
  __device__ __forceinline__ void lb(int j, double v)
  {
    atomic_max_f64(&dst[static_cast<size_t>(j) * W].lb, v);
    atomicOr(&dirty[j], bit);
  }
};

// round-0 variant: the candidate is appended to the list (entries past the capacity are counted but
// dropped; the host then falls back to the two-buffer round)
struct BtLogSink {
  BtLogEntry *log;
  unsigned *count;
  unsigned cap;
  int *flag;  // this warp's group flag
  int b;
  __device__ __forceinline__ void put(unsigned var, double v)
  {
    // one atomic per group of lanes that arrive here together (they usually do: same row, same
    // candidate, 32 similar vectors)
    const unsigned peers = __activemask();
    const int leader = __ffs(peers) - 1;
    const int laneId = threadIdx.x & 31;
    unsigned base = 0;
    if (laneId == leader)
      base = atomicAdd(count, static_cast<unsigned>(__popc(peers)));
    base = __shfl_sync(peers, base, leader);
    const unsigned at = base + __popc(peers & ((1u << laneId) - 1u));
    if (at < cap) {
      BtLogEntry e;
      e.var = var;
      e.b = b;
      e.v = v;
      log[at] = e;
    }
    *flag = 1;
  }
  __device__ __forceinline__ void ub(int j, double v) { put(static_cast<unsigned>(j) | 0x80000000u, v); }
  __device__ __forceinline__ void lb(int j, double v) { put(static_cast<unsigned>(j), v); }
};

template <bool kLog>
struct BtSinkFor {
  typedef BtSink type;
  static __device__ __forceinline__ BtSink make(const BtParams &p, int cur, int g, int b, int bitIdx)
  {
    BtSink s = { p.X[cur ^ 1] + bt_base(p, b), p.dirty[cur] + bt_word(p, 0, g), p.W, 1u << bitIdx };
    return s;
  }
};
template <>
struct BtSinkFor<true> {
  typedef BtLogSink type;
  static __device__ __forceinline__ BtLogSink make(const BtParams &p, int, int g, int b, int)
  {
    BtLogSink s = { p.log, p.logCount, p.logCap, p.groupFlag + g, b };
    return s;
  }
};

// One staged matrix entry: coefficient, column, integrality (16 bytes).
struct __align__(16) Stg {
  double a;
  int j;
  int isInt;
};

// The same for the first round of a base+diffs batch (32 bytes): coefficient, the column's BASE bounds,
// the group's diff mask for the column, and an upper bound (rounded up, single precision) of
// |a| * (ub - lb) of the base bounds.  The column index (integrality in bit 31) sits in a separate array,
// it is only needed by lanes whose own vector differs from the base there.
struct __align__(16) StgB {
  double a;
  double lo, hi;  // the base bounds that enter minAct / maxAct: (lb, ub) for a > 0, (ub, lb) for a < 0
  unsigned mask;
  float w;
};

// R3 for one (nonzero, vector).  kNoInf: no bound of this batch is infinite
// (checked when the batch is loaded; tightening never creates infinities), so
// the counters stay zero and the tests disappear.  kAllInt (walk): the matrix
// has no continuous column, so useful_range is the plain range.  kSign: +1 / -1 when the
// caller knows the sign of every coefficient in the range, 0 when mixed.
template <bool kNoInf, int kSign>
__device__ __forceinline__ void bt_acc(RowAct &ra, double a, bool isInt, const Bnd &x, double INF)
{
  const bool pos = kSign > 0 ? true : (kSign < 0 ? false : a > 0.0);
  const double lo = pos ? x.lb : x.ub, hi = pos ? x.ub : x.lb;
  if (kNoInf) {
    ra.minAct = fma(a, lo, ra.minAct);
    ra.maxAct = fma(a, hi, ra.maxAct);
    const double w = fabs(a) * useful_range(x.ub - x.lb, isInt);
    ra.maxRange = w > ra.maxRange ? w : ra.maxRange;
  } else {
    accumulate(ra, a, isInt, lo, hi, INF);
  }
}

// The row walk of one warp over nonzeros [kb, ke).  Per block of 32 nonzeros
// the lanes stage (a, j) in shared memory with one coalesced load each; then
// every lane walks the block in CSR order, reading (a, j) as a shared-memory
// broadcast and gathering the bounds of ITS vector for column j (the 32 lanes
// together read one 512-byte line of the [column][vector] array).  Eight
// gathers are in flight before the first is consumed.
template <bool kNoInf, int kSign, bool kAllInt>
__device__ __forceinline__ void bt_walk_sum(const BtParams &p, const char *__restrict__ src,
  unsigned strideB, int kb, int ke, int lane, Stg *stg, RowAct &ra)
{
  // the next block's (a, j) are fetched while the current block is being walked
  double nextA = 0.0;
  int nextC = 0;
  if (kb + lane < ke) {
    nextC = __ldg(&p.col[kb + lane]);
    nextA = __ldg(&p.val[kb + lane]);
  }
  for (int base = kb; base < ke; base += 32) {
    __syncwarp();
    {
      Stg e;
      e.a = nextA;
      e.j = col_of(nextC);
      e.isInt = col_is_int(nextC) ? 1 : 0;
      stg[lane] = e;
    }
    const int kn = base + 32 + lane;
    if (kn < ke) {
      nextC = __ldg(&p.col[kn]);
      nextA = __ldg(&p.val[kn]);
    }
    __syncwarp();
    const int cnt = min(32, ke - base);
    if (cnt == 32) {
#pragma unroll
      for (int i0 = 0; i0 < 32; i0 += 8) {
        Bnd x[8];
#pragma unroll
        for (int u = 0; u < 8; ++u)
          x[u] = ldcg_bnd(reinterpret_cast<const Bnd *>(
            src + static_cast<unsigned long long>(static_cast<unsigned>(stg[i0 + u].j)) * strideB));
#pragma unroll
        for (int u = 0; u < 8; ++u)
          bt_acc<kNoInf, kSign>(ra, stg[i0 + u].a, kAllInt ? true : stg[i0 + u].isInt != 0, x[u], p.s.inf);
      }
    } else {
      for (int i = 0; i < cnt; ++i) {
        const Bnd x = ldcg_bnd(reinterpret_cast<const Bnd *>(
          src + static_cast<unsigned long long>(static_cast<unsigned>(stg[i].j)) * strideB));
        bt_acc<kNoInf, kSign>(ra, stg[i].a, kAllInt ? true : stg[i].isInt != 0, x, p.s.inf);
      }
    }
  }
}

// First round of a base+diffs batch: the same walk, but a vector's bounds for column j are the shared
// base bounds unless the vector's diffs touch j (mask bit), and only then its own cell is read.  The
// lanes fetch base bounds and mask of "their" nonzero together with (a, j) into shared memory, one block
// ahead ((a, j) two blocks ahead: the base bounds need j), so the walk itself reads shared memory only:
// every (nonzero, vector) pair is still summed in CSR order, but the 16 bytes per pair no longer come
// through L2.
// kNoInf: the row-level shortcut of plan_row only needs SOME upper bound of max |a| * useful_range (a
// larger one lets more rows into the second walk, whose per-nonzero tests are exact, so results cannot
// change); here it is kept in single precision, rounded up, and for the base bounds it is computed once
// per nonzero by the staging lane.
template <bool kNoInf, int kSign, bool kAllInt>
__device__ __forceinline__ void bt_walk_sum_base(const BtParams &p, const char *__restrict__ src,
  unsigned strideB, const unsigned *__restrict__ dm, int kb, int ke, int lane, StgB *stg, int *stgJ, RowAct &ra)
{
  double a1 = 0.0, a2 = 0.0;
  int c1 = 0, c2 = 0;
  if (kb + lane < ke) {
    c1 = __ldg(&p.col[kb + lane]);
    a1 = __ldg(&p.val[kb + lane]);
  }
  if (kb + 32 + lane < ke) {
    c2 = __ldg(&p.col[kb + 32 + lane]);
    a2 = __ldg(&p.val[kb + 32 + lane]);
  }
  double2 x1 = __ldg(reinterpret_cast<const double2 *>(&p.baseX[col_of(c1)]));
  unsigned m1 = __ldg(&dm[col_of(c1)]);
  float mr = 0.f;  // kNoInf: running upper bound of the row's largest |a| * range
  for (int base = kb; base < ke; base += 32) {
    // does any vector of this group differ from the base at one of these 32 columns?  (hardly ever: then
    // the walk below needs no per-lane test at all)
    const bool diffs = __any_sync(0xffffffffu, m1 != 0u);
    __syncwarp();
    {
      StgB e;
      e.a = a1;
      e.lo = a1 > 0.0 ? x1.x : x1.y;
      e.hi = a1 > 0.0 ? x1.y : x1.x;
      e.mask = m1;
      e.w = __fmul_ru(__double2float_ru(fabs(a1)), __double2float_ru(__dsub_ru(x1.y, x1.x)));
      stg[lane] = e;
      stgJ[lane] = c1;
    }
    a1 = a2;
    c1 = c2;
    x1 = __ldg(reinterpret_cast<const double2 *>(&p.baseX[col_of(c1)]));  // (past the end: column 0, harmless)
    m1 = __ldg(&dm[col_of(c1)]);
    a2 = 0.0;
    c2 = 0;
    const int kn = base + 64 + lane;
    if (kn < ke) {
      c2 = __ldg(&p.col[kn]);
      a2 = __ldg(&p.val[kn]);
    }
    __syncwarp();
    const int cnt = min(32, ke - base);
    if (kNoInf && !diffs) {
      if (cnt == 32) {
#pragma unroll
        for (int i = 0; i < 32; ++i) {  // an entry = two 16-byte shared-memory broadcasts: {a, lo}, {hi, mask | w}
          const double2 h0 = reinterpret_cast<const double2 *>(&stg[i])[0];
          const double2 h1 = reinterpret_cast<const double2 *>(&stg[i])[1];
          const float w = __int_as_float(__double2hiint(h1.y));
          ra.minAct = fma(h0.x, h0.y, ra.minAct);
          ra.maxAct = fma(h0.x, h1.x, ra.maxAct);
          mr = fmaxf(mr, w);
        }
      } else {
        for (int i = 0; i < cnt; ++i) {
          const StgB e = stg[i];
          ra.minAct = fma(e.a, e.lo, ra.minAct);
          ra.maxAct = fma(e.a, e.hi, ra.maxAct);
          mr = fmaxf(mr, e.w);
        }
      }
      continue;
    }
    for (int i = 0; i < cnt; ++i) {
      const StgB e = stg[i];
      const bool pos = e.a > 0.0;
      Bnd x;
      x.lb = pos ? e.lo : e.hi;
      x.ub = pos ? e.hi : e.lo;
      float w = e.w;
      if ((e.mask >> lane) & 1u) {
        x = ldcg_bnd(reinterpret_cast<const Bnd *>(
          src + static_cast<unsigned long long>(static_cast<unsigned>(col_of(stgJ[i]))) * strideB));
        w = __fmul_ru(__double2float_ru(fabs(e.a)), __double2float_ru(__dsub_ru(x.ub, x.lb)));
      }
      if (kNoInf) {
        ra.minAct = fma(e.a, pos ? x.lb : x.ub, ra.minAct);
        ra.maxAct = fma(e.a, pos ? x.ub : x.lb, ra.maxAct);
        mr = fmaxf(mr, w);
      } else {
        bt_acc<false, kSign>(ra, e.a, kAllInt ? true : col_is_int(stgJ[i]), x, p.s.inf);
      }
    }
  }
  if (kNoInf)
    ra.maxRange = fmax(ra.maxRange, static_cast<double>(mr));
}

// second walk, only for warps where some lane's row can tighten something
template <bool kBase, class Sink>
__device__ __forceinline__ void bt_walk_tighten(const BtParams &p, const char *__restrict__ src,
  unsigned strideB, const unsigned *__restrict__ dm, int kb, int ke, int lane, Stg *stg, const RowPlan &pl,
  Sink &sink)
{
  const bool want = pl.sideU || pl.sideL;
  for (int base = kb; base < ke; base += 32) {
    const int k = base + lane;
    __syncwarp();
    if (k < ke) {
      const int c = __ldg(&p.col[k]);
      Stg e;
      e.a = __ldg(&p.val[k]);
      e.j = col_of(c);
      e.isInt = col_is_int(c) ? 1 : 0;
      stg[lane] = e;
    }
    __syncwarp();
    const int cnt = min(32, ke - base);
    if (want) {
      for (int i = 0; i < cnt; ++i) {
        const Stg e = stg[i];
        Bnd x;
        if (kBase && !((__ldg(&dm[e.j]) >> lane) & 1u)) {
          const double2 v = __ldg(reinterpret_cast<const double2 *>(&p.baseX[e.j]));
          x.lb = v.x;
          x.ub = v.y;
        } else {
          x = ldcg_bnd(reinterpret_cast<const Bnd *>(
            src + static_cast<unsigned long long>(static_cast<unsigned>(e.j)) * strideB));
        }
        tighten_nz(sink, pl, e.a, e.j, e.isInt != 0, e.a > 0.0 ? x.lb : x.ub, e.a > 0.0 ? x.ub : x.lb, p.s);
      }
    }
  }
}

// One round, rows phase, rows up to kBtLongLen nonzeros: a warp owns (row,
// group of 32 vectors) and sums in CSR order.  Work is group-major so that the
// warps running at the same time share few vector groups.
template <bool kNoInf, bool kAllInt, bool kLog, bool kBase = false>
__global__ void __launch_bounds__(kThreads, 4) bt_rows_kernel(const BtParams p, int cur)
{
  __shared__ __align__(16) unsigned char stageRaw[kWarps * 32 * (kBase ? sizeof(StgB) : sizeof(Stg))];
  __shared__ int stageJ[kBase ? kWarps : 1][32];
  Stg (*stage)[32] = reinterpret_cast<Stg (*)[32]>(stageRaw);
  StgB (*stageB)[32] = reinterpret_cast<StgB (*)[32]>(stageRaw);
  const long long W = static_cast<long long>(blockIdx.x) * kWarps + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (W >= static_cast<long long>(p.G) * p.nShortBt)
    return;
  const int g = static_cast<int>(W / p.nShortBt), r = __ldg(&p.shortRowsBt[W % p.nShortBt]);
  const unsigned act = __ldg(&p.active[g]);
  if (act == 0u)
    return;
  const bool mine = (act >> lane) & 1u;
  const int b = g * 32 + lane;
  const char *__restrict__ src = reinterpret_cast<const char *>(p.X[cur] + bt_base(p, b));
  const unsigned strideB = static_cast<unsigned>(p.W) * static_cast<unsigned>(sizeof(Bnd));
  const int kb = __ldg(&p.rowNz[r]), ke = kb + __ldg(&p.rowCnt[r]);

  const unsigned *__restrict__ dm = kBase ? p.dmask + bt_word(p, 0, g) : nullptr;
  RowAct ra = { 0.0, 0.0, 0.0, 0, 0 };
  if (kBase)
    bt_walk_sum_base<kNoInf, 0, kAllInt>(p, src, strideB, dm, kb, ke, lane, stageB[warp], stageJ[kBase ? warp : 0], ra);
  else
    bt_walk_sum<kNoInf, 0, kAllInt>(p, src, strideB, kb, ke, lane, stage[warp], ra);
  RowPlan pl = plan_row(ra, __ldg(&p.rowLB[r]), __ldg(&p.rowUB[r]), p.s);
  if (!mine) {
    pl.infeasible = false;
    pl.sideU = false;
    pl.sideL = false;
  }
  if (pl.infeasible)
    atomicMin(&p.vecInfRow[b], __ldg(&p.rowOrig[r]));
  if (!__any_sync(0xffffffffu, pl.sideU || pl.sideL))
    return;
  typename BtSinkFor<kLog>::type sink = BtSinkFor<kLog>::make(p, cur, g, b, lane);
  bt_walk_tighten<kBase>(p, src, strideB, dm, kb, ke, lane, kBase ? reinterpret_cast<Stg *>(stageB[warp]) : stage[warp],
    pl, sink);
}

// Same rows, but for matrices with so many columns that the bounds of a 32-vector group
// (n x 512 B) no longer fit in L2: a warp then owns (32/W rows, W vectors) with W = 16, 8 or 4,
// lanes [rowSlot][vector].  Work is (group, W-slice)-major, so everything running at one time
// touches n x W x 16 bytes of bounds, which L2 holds across all the rows that use a column
// (each cell comes from HBM once per round instead of once per row).  Every lane walks its own
// row in CSR order; W lanes read W x 16 contiguous bytes.
template <int W, bool kNoInf, bool kAllInt>
__global__ void __launch_bounds__(kThreads, 4) bt_rows_narrow_kernel(const BtParams p, int cur)
{
  constexpr int kRows = 32 / W;      // rows per warp
  constexpr int kSlices = 32 / W;    // W-slices per 32-vector group
  const int rowBlocks = (p.nShortBt + kRows - 1) / kRows;
  const long long widx = static_cast<long long>(blockIdx.x) * kWarps + (threadIdx.x >> 5);
  if (widx >= static_cast<long long>(p.G) * kSlices * rowBlocks)
    return;
  const int lane = threadIdx.x & 31;
  const int rb = static_cast<int>(widx % rowBlocks);
  const int gs = static_cast<int>(widx / rowBlocks);
  const int g = gs / kSlices, slice = gs % kSlices;
  const int vl = slice * W + (lane % W);  // this lane's vector inside the 32-group
  const unsigned act = __ldg(&p.active[g]);
  if (((act >> (slice * W)) & ((W == 32) ? 0xffffffffu : ((1u << W) - 1u))) == 0u)
    return;
  const int ri = rb * kRows + lane / W;
  const bool haveRow = ri < p.nShortBt;
  const bool mine = haveRow && ((act >> vl) & 1u);
  const int r = haveRow ? __ldg(&p.shortRowsBt[ri]) : 0;
  const int b = g * 32 + vl;
  const char *__restrict__ src = reinterpret_cast<const char *>(p.X[cur] + bt_base(p, b));
  const unsigned strideB = static_cast<unsigned>(p.W) * static_cast<unsigned>(sizeof(Bnd));
  const int kb = haveRow ? __ldg(&p.rowNz[r]) : 0, cnt = haveRow ? __ldg(&p.rowCnt[r]) : 0;

  // Each row slot stages its next 8 entries in shared memory (its W lanes fetch them, the entries
  // after those are already on their way), then every lane walks the 8 entries with 8 gathers in
  // flight.  All slots of the warp step together up to the longest of their rows.
  __shared__ Stg stage[kWarps][kRows * 8];
  Stg *stg = stage[threadIdx.x >> 5] + (lane / W) * 8;
  const int maxCnt = __reduce_max_sync(0xffffffffu, cnt);
  const int e0 = lane % W, e1 = e0 + W;  // W = 4: two entries per lane; W >= 8: one (e1 unused)
  double na0 = 0.0, na1 = 0.0;
  int nc0 = 0, nc1 = 0;
  if (e0 < 8 && e0 < cnt) {
    nc0 = __ldg(&p.col[kb + e0]);
    na0 = __ldg(&p.val[kb + e0]);
  }
  if (W < 8 && e1 < cnt) {
    nc1 = __ldg(&p.col[kb + e1]);
    na1 = __ldg(&p.val[kb + e1]);
  }
  RowAct ra = { 0.0, 0.0, 0.0, 0, 0 };
  for (int base = 0; base < maxCnt; base += 8) {
    __syncwarp();
    if (e0 < 8) {
      Stg e;
      e.a = na0;
      e.j = col_of(nc0);
      e.isInt = col_is_int(nc0) ? 1 : 0;
      stg[e0] = e;
    }
    if (W < 8) {
      Stg e;
      e.a = na1;
      e.j = col_of(nc1);
      e.isInt = col_is_int(nc1) ? 1 : 0;
      stg[e1] = e;
    }
    na0 = na1 = 0.0;
    nc0 = nc1 = 0;
    if (e0 < 8 && base + 8 + e0 < cnt) {
      nc0 = __ldg(&p.col[kb + base + 8 + e0]);
      na0 = __ldg(&p.val[kb + base + 8 + e0]);
    }
    if (W < 8 && base + 8 + e1 < cnt) {
      nc1 = __ldg(&p.col[kb + base + 8 + e1]);
      na1 = __ldg(&p.val[kb + base + 8 + e1]);
    }
    __syncwarp();
    Bnd x[8];
#pragma unroll
    for (int u = 0; u < 8; ++u)  // entries past the row's end are (0, column 0): a harmless gather
      x[u] = ldcg_bnd(reinterpret_cast<const Bnd *>(
        src + static_cast<unsigned long long>(static_cast<unsigned>(stg[u].j)) * strideB));
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const double a = stg[u].a;
      if (a != 0.0)
        bt_acc<kNoInf, 0>(ra, a, kAllInt ? true : stg[u].isInt != 0, x[u], p.s.inf);
    }
  }
  RowPlan pl = plan_row(ra, haveRow ? __ldg(&p.rowLB[r]) : 0.0, haveRow ? __ldg(&p.rowUB[r]) : 0.0, p.s);
  if (!mine) {
    pl.infeasible = false;
    pl.sideU = false;
    pl.sideL = false;
  }
  if (pl.infeasible)
    atomicMin(&p.vecInfRow[b], __ldg(&p.rowOrig[r]));
  if (pl.sideU || pl.sideL) {
    BtSink sink = { p.X[cur ^ 1] + bt_base(p, b), p.dirty[cur] + bt_word(p, 0, g), p.W, 1u << vl };
    for (int k = 0; k < cnt; ++k) {
      const int c = __ldg(&p.col[kb + k]);
      const double a = __ldg(&p.val[kb + k]);
      const int j = col_of(c);
      const Bnd x = ldcg_bnd(reinterpret_cast<const Bnd *>(src + static_cast<unsigned long long>(static_cast<unsigned>(j)) * strideB));
      tighten_nz(sink, pl, a, j, col_is_int(c), a > 0.0 ? x.lb : x.ub, a > 0.0 ? x.ub : x.lb, p.s);
    }
  }
}

// Rows longer than kBtLongLen: a CTA owns (row, group); its 8 warps split the
// row into 8 contiguous pieces, the partial activities are combined in warp
// order through shared memory.  With a whole CTA per row the CTAs in flight
// cover only one or two vector groups, whose bounds (n x 512 B each) then stay
// in L2 across all the rows that use them.  Such rows are stored with their
// positive coefficients first (rowPos of them), so no per-nonzero sign test.
#ifndef CBP_LONG_WARPS
#define CBP_LONG_WARPS 8
#endif
constexpr int kLongWarps = CBP_LONG_WARPS;  // warps that share one long row

template <bool kNoInf, bool kAllInt, bool kLog, bool kBase = false>
__global__ void __launch_bounds__(kLongWarps * 32, 2048 / (kLongWarps * 32) / 2) bt_rows_long_kernel(const BtParams p, int cur)
{
  __shared__ RowAct red[kLongWarps][32];
  __shared__ __align__(16) unsigned char stageRaw[kLongWarps * 32 * (kBase ? sizeof(StgB) : sizeof(Stg))];
  __shared__ int stageJ[kBase ? kLongWarps : 1][32];
  Stg (*stage)[32] = reinterpret_cast<Stg (*)[32]>(stageRaw);
  StgB (*stageB)[32] = reinterpret_cast<StgB (*)[32]>(stageRaw);
  const int g = blockIdx.x / p.nLongBt, r = __ldg(&p.longRowsBt[blockIdx.x % p.nLongBt]);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned act = __ldg(&p.active[g]);
  if (act == 0u)
    return;
  const bool mine = (act >> lane) & 1u;
  const int b = g * 32 + lane;
  const char *__restrict__ src = reinterpret_cast<const char *>(p.X[cur] + bt_base(p, b));
  const unsigned strideB = static_cast<unsigned>(p.W) * static_cast<unsigned>(sizeof(Bnd));
  const int kb = __ldg(&p.rowNz[r]), cnt = __ldg(&p.rowCnt[r]);
  const int ksplit = kb + __ldg(&p.rowPos[r]);              // first negative coefficient
  const int per = ((cnt + kLongWarps - 1) / kLongWarps + 31) & ~31;  // 32-aligned pieces
  const int wb = min(kb + warp * per, kb + cnt), we = min(wb + per, kb + cnt);

  const unsigned *__restrict__ dm = kBase ? p.dmask + bt_word(p, 0, g) : nullptr;
  RowAct ra = { 0.0, 0.0, 0.0, 0, 0 };
  if (kBase) {
    int *sj = stageJ[kBase ? warp : 0];
    bt_walk_sum_base<kNoInf, 1, kAllInt>(p, src, strideB, dm, wb, min(we, max(wb, ksplit)), lane, stageB[warp], sj, ra);
    bt_walk_sum_base<kNoInf, -1, kAllInt>(p, src, strideB, dm, max(wb, min(we, ksplit)), we, lane, stageB[warp], sj, ra);
  } else {
    bt_walk_sum<kNoInf, 1, kAllInt>(p, src, strideB, wb, min(we, max(wb, ksplit)), lane, stage[warp], ra);
    bt_walk_sum<kNoInf, -1, kAllInt>(p, src, strideB, max(wb, min(we, ksplit)), we, lane, stage[warp], ra);
  }
  red[warp][lane] = ra;
  __syncthreads();
  RowAct tot = red[0][lane];
#pragma unroll
  for (int w = 1; w < kLongWarps; ++w)
    combine(tot, red[w][lane]);
  RowPlan pl = plan_row(tot, __ldg(&p.rowLB[r]), __ldg(&p.rowUB[r]), p.s);
  if (!mine) {
    pl.infeasible = false;
    pl.sideU = false;
    pl.sideL = false;
  }
  if (pl.infeasible && warp == 0)
    atomicMin(&p.vecInfRow[b], __ldg(&p.rowOrig[r]));
  if (!__syncthreads_or(pl.sideU || pl.sideL))
    return;
  typename BtSinkFor<kLog>::type sink = BtSinkFor<kLog>::make(p, cur, g, b, lane);
  bt_walk_tighten<kBase>(p, src, strideB, dm, wb, we, lane, kBase ? reinterpret_cast<Stg *>(stageB[warp]) : stage[warp],
    pl, sink);
}

// Overlay rounds, long rows, no infinite bound in the batch: the walk above with V vector
// groups per warp.  What limits that walk is not arithmetic but the shared-memory broadcast: an entry is
// two 16-byte reads that are handed to all 32 lanes (512 bytes of register writes each, 128 bytes per
// clock and SM), two DFMAs behind them.  With V groups per warp every lane carries V vectors' sums and the
// same two reads feed 2V DFMAs.
template <int kSign, int V>
__device__ __forceinline__ void bt_walk_sum_base_multi(const BtParams &p, const char *const (&src)[V],
  unsigned strideB, const unsigned *const (&dm)[V], int kb, int ke, int lane, double2 *stgA, double *stgH, int *stgJ,
  unsigned (*stgM)[32], double (&mn)[V], double (&mx)[V], float (&mr)[V], float &mrBase)
{
  // A staged entry is 24 bytes here: {a, lo} in stgA, {hi} in stgH (the bytes that are broadcast to the lanes are
  // what the walk costs).  The base bounds' |a| * (ub - lb) bound does not travel at all: the staging lanes take
  // its maximum over the block with one warp reduction (adding the base value of a column that a lane has a cell
  // of its own for only makes that lane's upper bound larger, which is allowed).
  double a1 = 0.0, a2 = 0.0;
  int c1 = 0, c2 = 0;
  if (kb + lane < ke) {
    c1 = __ldg(&p.col[kb + lane]);
    a1 = __ldg(&p.val[kb + lane]);
  }
  if (kb + 32 + lane < ke) {
    c2 = __ldg(&p.col[kb + 32 + lane]);
    a2 = __ldg(&p.val[kb + 32 + lane]);
  }
  double2 x1 = __ldg(reinterpret_cast<const double2 *>(&p.baseX[col_of(c1)]));
  unsigned m1[V];
#pragma unroll
  for (int v = 0; v < V; ++v)
    m1[v] = dm[v] ? __ldg(&dm[v][col_of(c1)]) : 0u;
  for (int base = kb; base < ke; base += 32) {
    unsigned anyM = 0u;
#pragma unroll
    for (int v = 0; v < V; ++v)
      anyM |= m1[v];
    const bool diffs = __any_sync(0xffffffffu, anyM != 0u);
    const float w1 = __fmul_ru(__double2float_ru(fabs(a1)), __double2float_ru(__dsub_ru(x1.y, x1.x)));
    // (non-negative floats order like their bit patterns)
    mrBase = fmaxf(mrBase, __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(w1))));
    __syncwarp();
    stgA[lane] = make_double2(a1, a1 > 0.0 ? x1.x : x1.y);
    stgH[lane] = a1 > 0.0 ? x1.y : x1.x;
    if (diffs) {
      stgJ[lane] = c1;
#pragma unroll
      for (int v = 0; v < V; ++v)
        stgM[v][lane] = m1[v];
    }
    a1 = a2;
    c1 = c2;
    x1 = __ldg(reinterpret_cast<const double2 *>(&p.baseX[col_of(c1)]));
#pragma unroll
    for (int v = 0; v < V; ++v)
      m1[v] = dm[v] ? __ldg(&dm[v][col_of(c1)]) : 0u;
    a2 = 0.0;
    c2 = 0;
    const int kn = base + 64 + lane;
    if (kn < ke) {
      c2 = __ldg(&p.col[kn]);
      a2 = __ldg(&p.val[kn]);
    }
    __syncwarp();
    const int cnt = min(32, ke - base);
    if (!diffs) {
#pragma unroll 8
      for (int i = 0; i < cnt; ++i) {
        const double2 h0 = stgA[i];  // {a, lo}
        const double hi = stgH[i];
#pragma unroll
        for (int v = 0; v < V; ++v) {
          mn[v] = fma(h0.x, h0.y, mn[v]);
          mx[v] = fma(h0.x, hi, mx[v]);
        }
      }
      continue;
    }
    for (int i = 0; i < cnt; ++i) {
      const double2 h0 = stgA[i];
      const double a = h0.x;
      const double hiB = stgH[i];
      const bool pos = a > 0.0;
#pragma unroll
      for (int v = 0; v < V; ++v) {
        double lo = h0.y, hi = hiB;
        if ((stgM[v][i] >> lane) & 1u) {
          const Bnd x = ldcg_bnd(reinterpret_cast<const Bnd *>(
            src[v] + static_cast<unsigned long long>(static_cast<unsigned>(col_of(stgJ[i]))) * strideB));
          lo = pos ? x.lb : x.ub;
          hi = pos ? x.ub : x.lb;
          mr[v] = fmaxf(mr[v], __fmul_ru(__double2float_ru(fabs(a)), __double2float_ru(__dsub_ru(x.ub, x.lb))));
        }
        mn[v] = fma(a, lo, mn[v]);
        mx[v] = fma(a, hi, mx[v]);
      }
    }
  }
}

#ifndef CBP_BASE_V
#define CBP_BASE_V 4
#endif
#ifndef CBP_BASE_WARPS
#define CBP_BASE_WARPS 2
#endif
#ifndef CBP_BASE_MINB
#define CBP_BASE_MINB 8
#endif
constexpr int kBaseGroups = CBP_BASE_V;     // vector groups per warp
constexpr int kBaseWarps = CBP_BASE_WARPS;  // warps that share one long row

template <int V>
__global__ void __launch_bounds__(kBaseWarps * 32, CBP_BASE_MINB) bt_rows_long_base_kernel(const BtParams p, int cur)
{
  constexpr int kLongWarps = kBaseWarps;  // (shadows the generic kernels' constant on purpose)
  __shared__ RowAct red[V][kLongWarps][32];
  __shared__ double2 stageA[kLongWarps][32];
  __shared__ double stageH[kLongWarps][32];
  __shared__ int stageJ[kLongWarps][32];
  __shared__ unsigned stageM[kLongWarps][V][32];
  const int q = blockIdx.x / p.nLongBt, r = __ldg(&p.longRowsBt[blockIdx.x % p.nLongBt]);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned act[V];
  unsigned anyAct = 0u;
  const char *src[V];
  const unsigned *dm[V];
  constexpr unsigned strideB = 32u * static_cast<unsigned>(sizeof(Bnd));  // (these kernels exist for W = 32 only)
#pragma unroll
  for (int v = 0; v < V; ++v) {
    const int g = q * V + v;
    act[v] = g < p.G ? __ldg(&p.active[g]) : 0u;
    anyAct |= act[v];
    const int gc = min(g, p.G - 1);
    src[v] = reinterpret_cast<const char *>(p.X[cur] + (static_cast<size_t>(gc) * p.n) * 32 + lane);
    dm[v] = g < p.G ? p.dmask + static_cast<size_t>(g) * p.n : nullptr;
  }
  if (anyAct == 0u)
    return;
  const int kb = __ldg(&p.rowNz[r]), cnt = __ldg(&p.rowCnt[r]);
  const int ksplit = kb + __ldg(&p.rowPos[r]);
  const int per = ((cnt + kLongWarps - 1) / kLongWarps + 31) & ~31;
  const int wb = min(kb + warp * per, kb + cnt), we = min(wb + per, kb + cnt);

  double mn[V], mx[V];
  float mr[V], mrBase = 0.f;
#pragma unroll
  for (int v = 0; v < V; ++v) {
    mn[v] = 0.0;
    mx[v] = 0.0;
    mr[v] = 0.f;
  }
  bt_walk_sum_base_multi<1, V>(p, src, strideB, dm, wb, min(we, max(wb, ksplit)), lane, stageA[warp], stageH[warp],
    stageJ[warp], stageM[warp], mn, mx, mr, mrBase);
  bt_walk_sum_base_multi<-1, V>(p, src, strideB, dm, max(wb, min(we, ksplit)), we, lane, stageA[warp], stageH[warp],
    stageJ[warp], stageM[warp], mn, mx, mr, mrBase);
#pragma unroll
  for (int v = 0; v < V; ++v) {
    RowAct ra = { mn[v], mx[v], static_cast<double>(fmaxf(mr[v], mrBase)), 0, 0 };
    red[v][warp][lane] = ra;
  }
  __syncthreads();
  const double rl = __ldg(&p.rowLB[r]), ru = __ldg(&p.rowUB[r]);
#pragma unroll
  for (int v = 0; v < V; ++v) {
    const int g = q * V + v;
    if (g >= p.G)
      break;
    RowAct tot = red[v][0][lane];
#pragma unroll
    for (int w = 1; w < kLongWarps; ++w)
      combine(tot, red[v][w][lane]);
    RowPlan pl = plan_row(tot, rl, ru, p.s);
    if (!((act[v] >> lane) & 1u)) {
      pl.infeasible = false;
      pl.sideU = false;
      pl.sideL = false;
    }
    const int b = g * 32 + lane;
    if (pl.infeasible && warp == 0)
      atomicMin(&p.vecInfRow[b], __ldg(&p.rowOrig[r]));
    // (every warp holds the same plans, so they all take the same branch; the walk itself only needs the warp)
    if (__any_sync(0xffffffffu, pl.sideU || pl.sideL)) {
      BtLogSink sink = BtSinkFor<true>::make(p, cur, g, b, lane);
      bt_walk_tighten<true>(p, src[v], strideB, dm[v], wb, we, lane, reinterpret_cast<Stg *>(stageA[warp]), pl, sink);
    }
  }
}

#ifndef CBP_SHORT_MINB
#define CBP_SHORT_MINB 3
#endif
// The same for rows up to kBtLongLen nonzeros: a warp owns (row, V vector groups), no shared partial sums.
// Work is ordered group-quad-major so the warps in flight share their masks and touched cells.
template <int V>
__global__ void __launch_bounds__(kThreads, CBP_SHORT_MINB) bt_rows_short_base_kernel(const BtParams p, int cur)
{
  __shared__ double2 stageA[kWarps][32];
  __shared__ double stageH[kWarps][32];
  __shared__ int stageJ[kWarps][32];
  __shared__ unsigned stageM[kWarps][V][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // grid: x strides over the rows, y = group quad (quad-major: the warps in flight share masks and cells)
  const int ri = blockIdx.x * kWarps + warp;
  if (ri >= p.nShortBt)
    return;
  const int q = blockIdx.y, r = __ldg(&p.shortRowsBt[ri]);
  unsigned act[V];
  unsigned anyAct = 0u;
  const char *src[V];
  const unsigned *dm[V];
  constexpr unsigned strideB = 32u * static_cast<unsigned>(sizeof(Bnd));  // (these kernels exist for W = 32 only)
#pragma unroll
  for (int v = 0; v < V; ++v) {
    const int g = q * V + v;
    act[v] = g < p.G ? __ldg(&p.active[g]) : 0u;
    anyAct |= act[v];
    const int gc = min(g, p.G - 1);
    src[v] = reinterpret_cast<const char *>(p.X[cur] + (static_cast<size_t>(gc) * p.n) * 32 + lane);
    dm[v] = g < p.G ? p.dmask + static_cast<size_t>(g) * p.n : nullptr;
  }
  if (anyAct == 0u)
    return;
  const int kb = __ldg(&p.rowNz[r]), ke = kb + __ldg(&p.rowCnt[r]);
  double mn[V], mx[V];
  float mr[V], mrBase = 0.f;
#pragma unroll
  for (int v = 0; v < V; ++v) {
    mn[v] = 0.0;
    mx[v] = 0.0;
    mr[v] = 0.f;
  }
  bt_walk_sum_base_multi<0, V>(p, src, strideB, dm, kb, ke, lane, stageA[warp], stageH[warp], stageJ[warp],
    stageM[warp], mn, mx, mr, mrBase);
  const double rl = __ldg(&p.rowLB[r]), ru = __ldg(&p.rowUB[r]);
#pragma unroll
  for (int v = 0; v < V; ++v) {
    const int g = q * V + v;
    if (g >= p.G)
      break;
    if (act[v] == 0u)
      continue;
    const RowAct tot = { mn[v], mx[v], static_cast<double>(fmaxf(mr[v], mrBase)), 0, 0 };
    RowPlan pl = plan_row(tot, rl, ru, p.s);
    if (!((act[v] >> lane) & 1u)) {
      pl.infeasible = false;
      pl.sideU = false;
      pl.sideL = false;
    }
    const int b = g * 32 + lane;
    if (pl.infeasible)
      atomicMin(&p.vecInfRow[b], __ldg(&p.rowOrig[r]));
    if (__any_sync(0xffffffffu, pl.sideU || pl.sideL)) {
      BtLogSink sink = BtSinkFor<true>::make(p, cur, g, b, lane);
      bt_walk_tighten<true>(p, src[v], strideB, dm[v], kb, ke, lane, reinterpret_cast<Stg *>(stageA[warp]), pl, sink);
    }
  }
}

// After the rows phase: settle every (column, vector) marked in this round (R9) and count the
// changed columns per vector.  A warp takes 32 consecutive columns of one vector group: lane i
// fetches column i's mark word, then the words that are not zero are handled one at a time with
// lane = vector, so the 32 cells of a column are read and written as one 512-byte piece and every
// lane keeps its own vector's count in a register (one atomic per lane at the end, instead of one
// per changed bound).
__global__ void __launch_bounds__(kThreads) bt_settle_kernel(const BtParams p, int cur)
{
  const int g = blockIdx.y;
  const int lane = threadIdx.x & 31;
  const int warpsPerGrid = gridDim.x * kWarps;
  const int b = g * 32 + lane;
  const size_t base = bt_base(p, b);
  const unsigned *marks = p.dirty[cur] + bt_word(p, 0, g);
  int mine = 0;
  for (int j0 = (blockIdx.x * kWarps + (threadIdx.x >> 5)) * 32; j0 < p.n; j0 += warpsPerGrid * 32) {
    const int j = j0 + lane;
    const unsigned word = j < p.n ? marks[j] : 0u;
    unsigned any = __ballot_sync(0xffffffffu, word != 0u);
    while (any) {
      const int sl = __ffs(any) - 1;
      any &= any - 1;
      const unsigned w = __shfl_sync(0xffffffffu, word, sl);
      if ((w >> lane) & 1u) {
        const int col = j0 + sl;
        const size_t at = base + static_cast<size_t>(col) * p.W;
        Bnd x = p.X[cur ^ 1][at];
        const double L0 = x.lb, U0 = x.ub;
        if (settle(x.lb, x.ub, p.X[cur][at].ub, p.s))
          atomicMin(&p.vecInfCol[b], col);
        else if (x.lb != L0 || x.ub != U0)
          p.X[cur ^ 1][at] = x;
        ++mine;
      }
    }
  }
  if (mine)
    atomicAdd(&p.vecChanged[b], mine);
}

// Per-vector bookkeeping of src/CbcBoundPropagation.cpp:129-189 after a round.
// guard (may be null): two words — if either is non-zero the kernel does nothing (the host launches it ahead of
// knowing whether the first round was quiet; the words are the candidate count and the diffs' R11 flag).
__global__ void bt_state_kernel(const BtParams p, int cur, int round, int limit, int timeUp,
  const int *__restrict__ guard = nullptr)
{
  if (guard && (guard[0] | guard[1]))
    return;
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  bool on = false;
  if (b < p.Bp) {
    on = (p.active[b >> 5] >> lane) & 1u;
    if (on) {
      const int chg = p.vecChanged[b];
      p.vecChanged[b] = 0;
      p.vecRounds[b] = round + 1;
      int stop = 0, fin = cur;
      if (p.vecInfRow[b] != 0x7fffffff || p.vecInfCol[b] != 0x7fffffff)
        stop = 4;
      else if (chg == 0)
        stop = 1;
      else {
        p.vecTotal[b] += chg;
        if (round + 1 >= limit) {
          stop = 2;
          fin = cur ^ 1;
        } else if (timeUp) {
          stop = 3;
          fin = cur ^ 1;
        }
      }
      if (stop) {
        p.vecStop[b] = stop;
        p.vecFinal[b] = fin;
        on = false;
      }
    }
  }
  const unsigned m = __ballot_sync(0xffffffffu, on);
  if (lane == 0 && b < p.Bp) {
    p.active[b >> 5] = m;
    if (m)
      atomicOr(p.anyActive, 1);
  }
}

// Start of the next round: copy this round's settled values into the buffer that becomes the
// atomics target (it was the source, so it is stale exactly at the marked entries), only for vectors
// that go on; fold the marks of applied rounds into `ever`; clear the marks.  Same walk as the
// settle kernel: lane i owns column i's word, lanes become vectors for the words that are not zero.
__global__ void __launch_bounds__(kThreads) bt_carry_kernel(const BtParams p, int cur)
{
  const int g = blockIdx.y;
  const int lane = threadIdx.x & 31;
  const int warpsPerGrid = gridDim.x * kWarps;
  const int b = g * 32 + lane;
  const size_t base = bt_base(p, b);
  unsigned *marks = p.dirty[cur] + bt_word(p, 0, g);
  unsigned *ever = p.ever + bt_word(p, 0, g);
  const unsigned act = p.active[g];
  const bool running = (act >> lane) & 1u;
  // a vector's round was applied iff it is still running or stopped with the new buffer as its answer
  const bool appliedMe = running || p.vecFinal[b] == (cur ^ 1);
  const unsigned appliedMask = __ballot_sync(0xffffffffu, appliedMe);
  for (int j0 = (blockIdx.x * kWarps + (threadIdx.x >> 5)) * 32; j0 < p.n; j0 += warpsPerGrid * 32) {
    const int j = j0 + lane;
    const unsigned word = j < p.n ? marks[j] : 0u;
    if (word) {
      marks[j] = 0u;
      if (word & appliedMask)
        ever[j] |= word & appliedMask;
    }
    unsigned any = __ballot_sync(0xffffffffu, (word & act) != 0u);
    while (any) {
      const int sl = __ffs(any) - 1;
      any &= any - 1;
      const unsigned w = __shfl_sync(0xffffffffu, word, sl);
      if (running && ((w >> lane) & 1u)) {
        const size_t at = base + static_cast<size_t>(j0 + sl) * p.W;
        p.X[cur][at] = p.X[cur ^ 1][at];
      }
    }
  }
}

// ---- loading ------------------------------------------------------------------

// does any base column need R11 (non-integral integer bounds, crossing bounds)?
__global__ void bt_base_check_kernel(const double *__restrict__ baseLb, const double *__restrict__ baseUb,
  const signed char *__restrict__ colType, int n, Scal s, int *flag)
{
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n)
    return;
  double l = baseLb[j], u = baseUb[j];
  if (prepass_col(l, u, colType[j] != 0, s) != 0)
    atomicOr(flag, 1);
}

// is any bound of the loaded batch (base or diffs) infinite?
__global__ void bt_has_inf_kernel(const double *__restrict__ baseLb, const double *__restrict__ baseUb, int n,
  const double *__restrict__ dBnd, long long nDelta, double INF, int *flag)
{
  bool bad = false;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    bad = bad || is_inf(baseLb[i], INF) || is_inf(baseUb[i], INF);
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < nDelta; i += stride)
    bad = bad || is_inf(dBnd[i], INF);
  if (bad)
    *flag = 1;
}

// every vector starts from the base bounds; both buffers are written
__global__ void bt_expand_kernel(const double *__restrict__ baseLb, const double *__restrict__ baseUb,
  int n, int W, size_t cells, Bnd *X0, Bnd *X1)
{
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= cells)
    return;
  const int j = static_cast<int>((i / W) % n);
  Bnd x;
  x.lb = __ldg(&baseLb[j]) + 0.0;  // (no -0.0 in the bound arrays)
  x.ub = __ldg(&baseUb[j]) + 0.0;
  X0[i] = x;
  if (X1)  // null: the second buffer is built later, only for the groups that need it
    X1[i] = x;
}

// interleaved copy of the base bounds for the overlay form
__global__ void bt_base_pack_kernel(const double *__restrict__ baseLb, const double *__restrict__ baseUb, int n,
  Bnd *baseX)
{
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) {
    Bnd x;
    x.lb = baseLb[j] + 0.0;  // (no -0.0 in the bound arrays)
    x.ub = baseUb[j] + 0.0;
    baseX[j] = x;
  }
}

// Leave the overlay form: every cell no diff or candidate has touched yet gets the base bounds in both
// buffers (touched cells are valid in both already), so the classic two-buffer kernels can take over.
__global__ void bt_overlay_expand_kernel(const BtParams p)
{
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= static_cast<size_t>(p.n) * p.Bp)
    return;
  const int b = static_cast<int>(i / (static_cast<size_t>(p.W) * p.n)) * p.W + static_cast<int>(i % p.W);
  const int j = static_cast<int>((i / p.W) % p.n);
  if ((p.dmask[bt_word(p, j, b >> 5)] >> (b & 31)) & 1u)
    return;
  const Bnd x = p.baseX[j];
  p.X[0][i] = x;
  p.X[1][i] = x;
}

// Overlay rounds, before the candidate list is applied: a cell that a candidate is about to tighten for
// the first time becomes a real cell — base bounds in both buffers, mask bit set.  (A launch of its own:
// the atomics of bt_apply_log_kernel must find every cell initialised.)
__global__ void bt_overlay_touch_kernel(const BtParams p, unsigned count)
{
  const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count)
    return;
  const BtLogEntry e = p.log[i];
  const int j = static_cast<int>(e.var & 0x7fffffffu);
  unsigned *w = const_cast<unsigned *>(&p.dmask[bt_word(p, j, e.b >> 5)]);
  const unsigned bit = 1u << (e.b & 31);
  if (__ldcg(w) & bit)
    return;
  if (atomicOr(w, bit) & bit)
    return;  // somebody else got there first and writes the cell
  const Bnd x = p.baseX[j];
  const size_t at = bt_cell(p, j, e.b);
  p.X[0][at] = x;
  p.X[1][at] = x;
}

// second buffer := first buffer, for the vector groups whose flag is set (all if flags == nullptr)
__global__ void bt_copy_groups_kernel(const BtParams p, const int *__restrict__ flags)
{
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= static_cast<size_t>(p.n) * p.Bp)
    return;
  if (flags) {
    const int b = static_cast<int>(i / (static_cast<size_t>(p.W) * p.n)) * p.W + static_cast<int>(i % p.W);
    if (!flags[b >> 5])
      return;
  }
  p.X[1][i] = p.X[0][i];
}

// a list round's candidates, applied to the target buffer exactly as the direct sink would
__global__ void bt_apply_log_kernel(const BtParams p, unsigned count, int cur)
{
  const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count)
    return;
  const BtLogEntry e = p.log[i];
  const int j = static_cast<int>(e.var & 0x7fffffffu);
  Bnd *cell = &p.X[cur ^ 1][bt_cell(p, j, e.b)];
  if (e.var & 0x80000000u)
    atomic_min_f64(&cell->ub, e.v);
  else
    atomic_max_f64(&cell->lb, e.v);
  atomicOr(&p.dirty[cur][bt_word(p, j, e.b >> 5)], 1u << (e.b & 31));
}

// device-pointer batches: is every diff's column inside the matrix?  (host-pointer batches are checked on the host)
__global__ void bt_delta_check_kernel(const unsigned *__restrict__ var, long long count, int n, int *bad)
{
  const long long k = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k < count && diff_col(var[k]) >= n)
    *bad = 1;
}

// node bound diffs, applied in order inside a vector (one thread per vector).
// With cellPrepass the touched cells also get R11 here (the fast path, taken
// when no base column needs R11, so untouched cells are already final).
__global__ void bt_delta_kernel(const long long *__restrict__ start, const unsigned *__restrict__ var,
  const double *__restrict__ bnd, int B, const BtParams p, int cellPrepass, int lazy)
{
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B)
    return;
  const long long k0 = start[b], k1 = start[b + 1];
  unsigned *dmaskW = const_cast<unsigned *>(p.dmask);  // lazy == 2: nothing has been expanded, cells are the base
  for (long long k = k0; k < k1; ++k) {                // bounds until this vector's diffs touch them
    const unsigned v = var[k];
    const int jv = diff_col(v);
    if (jv >= p.n)  // (rejected when the batch was loaded; never index past the arrays)
      continue;
    const size_t at = bt_cell(p, jv, b);
    Bnd x;
    if (lazy == 2) {
      unsigned *w = &dmaskW[bt_word(p, jv, b >> 5)];
      const bool had = (*reinterpret_cast<volatile unsigned *>(w) >> (b & 31)) & 1u;  // (own bit: set by this thread only)
      x = had ? p.X[0][at] : p.baseX[jv];
      if (!had)
        atomicOr(w, 1u << (b & 31));
    } else {
      x = p.X[0][at];
    }
    // (-0.0 never enters the bound arrays: the integer-ordered atomics rely on it)
    if (v & kDiffUpper)
      x.ub = bnd[k] + 0.0;
    else
      x.lb = bnd[k] + 0.0;
    p.X[0][at] = x;
    if (lazy != 1)  // (overlay form: a touched cell is valid in both buffers)
      p.X[1][at] = x;
  }
  if (!cellPrepass)
    return;
  for (long long k = k0; k < k1; ++k) {
    const int j = diff_col(var[k]);
    if (j >= p.n)
      continue;
    const size_t at = bt_cell(p, j, b);
    Bnd x = p.X[0][at];
    const int r = prepass_col(x.lb, x.ub, p.colType[j] != 0, p.s);
    if (r == 2)
      atomicMin(&p.vecInfCol[b], j);
    else if (r == 1) {
      atomicOr(&p.dirty[0][bt_word(p, j, b >> 5)], 1u << (b & 31));
      p.X[0][at] = x;
      if (lazy != 1)
        p.X[1][at] = x;
      if (lazy) {
        p.groupFlag[b >> 5] = 1;
        p.groupFlag[p.G + 1] = 1;  // "R11 changed some diff cell"
      }
    }
  }
}

// after a quiet run from the base: take the diffs' bits out of the mask again, so that the next run of
// this (or another) batch finds it empty without clearing all of it
__global__ void bt_delta_unmask_kernel(const long long *__restrict__ start, const unsigned *__restrict__ var, int B,
  const BtParams p)
{
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B)
    return;
  unsigned *dmaskW = const_cast<unsigned *>(p.dmask);
  for (long long k = start[b]; k < start[b + 1]; ++k) {
    const int j = diff_col(var[k]);
    if (j < p.n)
      atomicAnd(&dmaskW[bt_word(p, j, b >> 5)], ~(1u << (b & 31)));
  }
}

// dense input, B x n row-major per vector -> slice-major cells; 32x32 tiles
__global__ void bt_transpose_in_kernel(const double *__restrict__ lb, const double *__restrict__ ub,
  int B, const BtParams p, double padLb, double padUb)
{
  __shared__ double tl[32][33], tu[32][33];
  const int n = p.n;
  const int j0 = blockIdx.x * 32, b0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int b = b0 + i, j = j0 + threadIdx.x;
    const bool ok = b < B && j < n;
    tl[i][threadIdx.x] = (ok ? lb[static_cast<size_t>(b) * n + j] : padLb) + 0.0;  // (no -0.0 in the bound arrays)
    tu[i][threadIdx.x] = (ok ? ub[static_cast<size_t>(b) * n + j] : padUb) + 0.0;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int j = j0 + i, b = b0 + threadIdx.x;
    if (j < n) {
      Bnd x;
      x.lb = tl[threadIdx.x][i];
      x.ub = tu[threadIdx.x][i];
      p.X[0][bt_cell(p, j, b)] = x;
    }
  }
}

// R11 per (column, vector); copies X0 into X1; seeds round 0's marks
__global__ void bt_prepass_kernel(const BtParams p, int B)
{
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= static_cast<size_t>(p.n) * p.Bp)
    return;
  const int v = static_cast<int>(i % p.W);
  const int j = static_cast<int>((i / p.W) % p.n);
  const int b = static_cast<int>(i / (static_cast<size_t>(p.W) * p.n)) * p.W + v;
  Bnd x = p.X[0][i];
  if (b < B) {
    const int r = prepass_col(x.lb, x.ub, p.colType[j] != 0, p.s);
    if (r == 2)
      atomicMin(&p.vecInfCol[b], j);
    else if (r == 1) {
      atomicOr(&p.dirty[0][bt_word(p, j, b >> 5)], 1u << (b & 31));
      p.X[0][i] = x;
    }
  }
  p.X[1][i] = x;
}

// a vector whose input already crosses (R11) stops before its first round
__global__ void bt_prepass_state_kernel(const BtParams p)
{
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  bool on = false;
  if (b < p.Bp) {
    on = (p.active[b >> 5] >> lane) & 1u;
    if (on && p.vecInfCol[b] != 0x7fffffff) {
      p.vecStop[b] = 4;
      p.vecRounds[b] = 1;
      p.vecFinal[b] = 0;
      on = false;
    }
  }
  const unsigned m = __ballot_sync(0xffffffffu, on);
  if (lane == 0 && b < p.Bp)
    p.active[b >> 5] = m;
}

__global__ void bt_init_state_kernel(const BtParams p, int B)
{
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= p.Bp)
    return;
  p.vecInfRow[b] = 0x7fffffff;
  p.vecInfCol[b] = 0x7fffffff;
  p.vecChanged[b] = 0;
  p.vecRounds[b] = 0;
  p.vecTotal[b] = 0;
  p.vecStop[b] = 0;
  p.vecFinal[b] = 0;
  const unsigned m = __ballot_sync(0xffffffffu, b < B);
  if ((threadIdx.x & 31) == 0)
    p.active[b >> 5] = m;
}

// ---- fetching -------------------------------------------------------------------

// dense answer: each vector's final buffer -> B x n row-major
__global__ void bt_transpose_out_kernel(const BtParams p, int B, double *lb, double *ub)
{
  __shared__ double tl[32][33], tu[32][33];
  const int j0 = blockIdx.x * 32, b0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int j = j0 + i, b = b0 + threadIdx.x;
    if (j < p.n && b < B) {
      const Bnd x = p.X[p.vecFinal[b]][bt_cell(p, j, b)];
      tl[i][threadIdx.x] = x.lb;
      tu[i][threadIdx.x] = x.ub;
    }
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int b = b0 + i, j = j0 + threadIdx.x;
    if (b < B && j < p.n) {
      lb[static_cast<size_t>(b) * p.n + j] = tl[threadIdx.x][i];
      ub[static_cast<size_t>(b) * p.n + j] = tu[threadIdx.x][i];
    }
  }
}

// ---- sparse answer (updatedBounds() per vector) ------------------------------
// `ever` is almost empty, so the scan is organised to skip it fast: a warp owns
// (vector group, chunk of kUpdChunk columns); its lanes read the marks of 32
// columns at a time, one word each, and only words that are not zero are looked
// at bit by bit (lane = vector).  Pass 1 counts per (chunk, vector), a tiny
// prefix kernel turns the counts into offsets, pass 2 writes (col, lb, ub) in
// ascending column order.
constexpr int kUpdChunk = 1024;

template <bool kFill>
__global__ void bt_updates_kernel(const BtParams p, int B, int *chunkCnt, const long long *__restrict__ start,
  int *updCol, double *updLb, double *updUb)
{
  const int g = blockIdx.x, chunk = blockIdx.y * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  const int c0 = chunk * kUpdChunk, c1 = min(p.n, c0 + kUpdChunk);
  if (c0 >= p.n)
    return;
  const int b = g * 32 + lane;
  const bool live = b < B && p.vecStop[b] != 4;  // a vector proved infeasible lists nothing
  const unsigned *e = p.ever + bt_word(p, 0, g);
  const size_t slot = static_cast<size_t>(chunk) * p.Bp + b;
  long long at = 0;
  const Bnd *X = nullptr;
  if (kFill && live) {
    at = start[b] + chunkCnt[slot];
    X = p.X[p.vecFinal[b]] + bt_base(p, b);
  }
  int cnt = 0;
  for (int j0 = c0; j0 < c1; j0 += 32) {
    const int j = j0 + lane;
    const unsigned word = j < c1 ? e[j] : 0u;
    unsigned any = __ballot_sync(0xffffffffu, word != 0u);
    while (any) {
      const int sl = __ffs(any) - 1;
      any &= any - 1;
      const unsigned w = __shfl_sync(0xffffffffu, word, sl);
      if (live && ((w >> lane) & 1u)) {
        if (kFill) {
          const int col = j0 + sl;
          const Bnd x = X[static_cast<size_t>(col) * p.W];
          updCol[at] = col;
          updLb[at] = x.lb;
          updUb[at] = x.ub;
          ++at;
        }
        ++cnt;
      }
    }
  }
  if (!kFill)
    chunkCnt[slot] = cnt;
}

// per vector: exclusive prefix of its chunk counts (in place) and its total
__global__ void bt_updates_prefix_kernel(int *chunkCnt, int nChunks, int Bp, int B, long long *count)
{
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B)
    return;
  long long run = 0;
  for (int c = 0; c < nChunks; ++c) {
    const size_t slot = static_cast<size_t>(c) * Bp + b;
    const int v = chunkCnt[slot];
    chunkCnt[slot] = static_cast<int>(run);
    run += v;
  }
  count[b] = run;
}

}  // namespace cbp
