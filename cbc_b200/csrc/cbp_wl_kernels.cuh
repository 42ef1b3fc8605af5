// cbp_wl_kernels.cuh — node batches in worklist form: the whole round loop of a base + diffs batch
// (/root/reference/src/CbcBoundPropagation.cpp:107-195, once per vector) as ONE persistent cooperative
// kernel, with no host round trip between the first and the last round.
//
// What makes it cheap.  The batch is in overlay form (cbp_batch_kernels.cuh): a vector's bounds are the shared
// node bounds except at the cells its diffs or its accepted candidates have touched.  If the node bounds
// themselves are a propagation fixpoint (the kernel checks: no row of the matrix is violated or can tighten
// anything from them, no column needs R11 — true for a node Cbc has already propagated), then in a vector's
// first round only the rows that hold one of its diff columns can do anything, and in every later round only
// the rows that hold a column the previous round changed: a row none of whose columns moved has the same bounds
// as the last time it was evaluated (or the node's), hence the same candidates, all of them already rejected
// or applied.  So a round evaluates a LIST of (vector, row) items, generated from the changed cells through the
// column-major copy of the matrix, instead of every (row, vector) pair.  Per item the row is walked in CSR
// order from "node bounds unless this vector has a cell of its own" — for rows of at most 32 nonzeros by one
// thread, sequentially, bit for bit the sum of the CPU restatement; longer rows by one warp (lanes stride,
// butterfly: the order of the single-vector kernel's medium rows).
//
// Phases of a round (grid-wide barriers between them; lists and counters live in global memory):
//   P1 items  -> row activity, R4 (violated row), R5-R8 candidates appended to the candidate list
//   P2 list   -> cells that a candidate touches for the first time are created (node bounds, mask bit)
//   P3 list   -> candidates applied (atomic max/min on the other buffer); the first to mark a cell lists it
//   P4 cells  -> R9 per changed cell, changed-column counts per vector
//   P5 + P6   -> per-vector bookkeeping of R12; changed cells carried into the next round's source buffer,
//                their rows listed as the next round's items
// When P1 leaves the candidate list empty and no cell is marked, P2-P4 are skipped.
// If the node bounds are not a fixpoint, or a list overflows, the kernel says so and the host runs the batch
// through the full-walk kernels instead (same results, every pair walked).
#pragma once
#include <cooperative_groups.h>

#include "cbp_batch_kernels.cuh"

namespace cbp {

struct WlItem {
  int b, row;
};
struct WlCell {
  int b, j;
};

enum { kWlOk = 0, kWlBaseNotFixpoint = 1, kWlOverflow = 2 };

// A (vector, row) pair whose row is too long for one warp (more than kBtLongLen nonzeros) and has to be walked — for
// its sums (needSum) or, sums known, for the tightening with the plan kept here: done after the items by a whole CTA
// per entry, so that one 4,000-nonzero row walked by one warp is not the tail of the phase.
struct __align__(16) WlWalk {
  int b, row;
  int needSum;
  int packed;  // plan: infMin | infMax << 2 | sideU << 4 | sideL << 5
  double slackU, slackL;
};

// One slot of the (vector, row) pair table: a 32-byte sector, so that finding the pair, adding to its sums and stamping
// it touch one sector of HBM, not four.
struct __align__(32) WlPair {
  unsigned long long key;  // ((b << 32) | row) + 1, 0 = free
  double dMin, dMax;       // sum over the vector's own cells in the row of a * (own bound - node's bound)
  int stamp;               // last round (+1) the pair was listed for
  int pad;
};
// The node's sums of a row (written by phase 0), one sector as well.
struct __align__(32) WlRowBase {
  double mn, mx, range;
  int exact;  // every sum over this row is exact in any order
  int pad;
};

// One entry of the head of the update lists, as it travels to the host with the control block.
struct WlUpdHead {
  int col, pad;
  double lb, ub;
};

// control block (device); zeroed by wl_setup_kernel before every launch
struct WlCtl {
  unsigned nS[2], nL[2];  // items of the running / next round: rows <= kShortLen nonzeros, longer rows
  unsigned nLog, nK, nCells;
  unsigned nWalk;  // entries of the long-row walk queue of the running round
  int baseBad, timeUp;
  int inexact;     // a diff is fractional or leaves the node's bounds: long rows are walked, not updated from the node's sums
  int overflow;    // a list other than the candidate list ran out of room (written outside P1 only)
  int overflowP1;  // the candidate list did (written in P1 only): two words, so that each is read while nobody writes it
  int anyActive[2];
  int status, rounds;
  int updDone;       // the update lists of the vectors were written by the kernel itself (nUpd entries)
  unsigned long long nUpd;
  unsigned nUsed;    // table slots taken by this run (listed in WlParams::used; the next run empties them first)
  int usedOverflow;  // more than the list holds: the host clears the whole table before the next run
  unsigned long long pairs;   // (nonzero, vector) pairs walked
  unsigned long long itemsTotal;
  unsigned long long roundsTotal;  // rounds summed over the vectors (each adds its count when it stops)
  unsigned long long t0;
  // diagnostic: ns since t0 at the first phase boundaries (CBP_WL_STAMPS=1 prints them)
  unsigned nStamps;
  unsigned stamps[96];
};

struct WlParams {
  BtParams bt;
  const int *cscStart, *cscRow;
  const double *cscVal;
  // sliced-ELL copy of the rows of at most kShortLen nonzeros (slice = 32 rows, entry k of lane l at
  // sellOff[s] + 32 k + l): the base check reads them coalesced
  const double *sellVal;
  const int *sellCol, *sellOff;
  int nSlices, mShort;
  // Rows longer than kShortLen: one table slot per (vector, row) pair that has come up, found by hashing.  It keeps
  // an item from being listed twice in a round and, where sums are exact (see exactOk), the pair's activity as
  // "the node's sums + what the vector's own cells add".
  WlPair *tab;
  unsigned hMask;            // slots in use - 1 (a power of two)
  int hShift;                // 64 - log2(slots in use): the hash is the top bits of key x odd constant
  unsigned *used;            // slots this run has taken, so that the run after it starts from an empty table without
  unsigned usedCap;          // a memset of all of it
  WlRowBase *rowBase;        // [mKept]
  int exactOk;  // host: every column is an integer column (rows that see an infinite bound are not exact: phase 0 finds them)
  // the loaded batch
  const long long *dStart;
  const unsigned *dVar;
  const double *dBnd;
  int B;
  // lists
  WlItem *itemsS[2], *itemsL[2];
  WlCell *K, *cells;
  WlWalk *walk;
  unsigned walkCap;
  unsigned itemCap, kCap, cellCap;
  unsigned prevUsed;   // table slots the previous worklist run took (listed in `used`; emptied first)
  unsigned prevCells;  // cells the previous worklist run left behind (their mask / ever bits are taken out first)
  WlCtl *ctl;
  int *vecChanged2[2];   // [Bp] changed columns, by round parity
  unsigned *active2[2];  // [G] running vectors, by round parity
  int limit;             // round limit (INT_MAX: none)
  unsigned long long budgetNs;
  // updatedBounds() of every vector, written at the end of the run when the buffers are large enough: per vector
  // the changed columns in ascending order with their final bounds (updTmp*: the same entries in arrival order)
  long long *updStart;  // [B + 1]
  int *updCnt;          // [2 Bp] counts, cursors
  int *updColTmp, *updCol;
  double *updLbTmp, *updUbTmp, *updLb, *updUb;
  long long updCap;
  WlUpdHead *updHead;  // the first updHeadCap entries of the ordered lists once more, next to the control block
  long long updHeadCap;
};

namespace cg = cooperative_groups;

__device__ __forceinline__ void wl_stamp(WlCtl *ctl, unsigned gtid)
{
  if (gtid == 0 && ctl->nStamps < 96) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    ctl->stamps[ctl->nStamps++] = static_cast<unsigned>(t - ctl->t0);
  }
}

// bounds of vector (g, vl) at column j in the round's source buffer: the node's unless the vector has its own cell
__device__ __forceinline__ Bnd wl_bounds(const BtParams &p, const Bnd *__restrict__ srcCells, const unsigned *__restrict__ dm,
  int vl, int j)
{
  const unsigned m = __ldcg(&dm[j]);
  if ((m >> vl) & 1u)
    return ldcg_bnd(&srcCells[static_cast<size_t>(j) * 32]);
  const double2 v = __ldg(reinterpret_cast<const double2 *>(&p.baseX[j]));
  Bnd x;
  x.lb = v.x;
  x.ub = v.y;
  return x;
}

// candidates of a worklist round go to the list (entries past the capacity are counted and dropped: the run is
// then repeated by the full-walk kernels)
constexpr unsigned kWlStage = 1024;  // candidates a CTA collects in shared memory before it asks the global list for room

struct WlLogSink {
  BtLogEntry *log;
  unsigned *count;
  unsigned cap;
  int *overflow;
  int b;
  // The candidates of a phase are staged per CTA in shared memory and appended to the global list with ONE
  // reservation per CTA when the phase ends (every candidate asking the one global counter for its slot is what the
  // phase would otherwise spend its time on); what does not fit the stage goes to the global list directly.
  BtLogEntry *stage;
  unsigned *staged;
  __device__ __forceinline__ void put(unsigned var, double v)
  {
    if (stage) {
      const unsigned at = atomicAdd(staged, 1u);
      if (at < kWlStage) {
        BtLogEntry e;
        e.var = var;
        e.b = b;
        e.v = v;
        stage[at] = e;
        return;
      }
    }
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
    } else {
      *overflow = 1;
    }
  }
  __device__ __forceinline__ void ub(int j, double v) { put(static_cast<unsigned>(j) | 0x80000000u, v); }
  __device__ __forceinline__ void lb(int j, double v) { put(static_cast<unsigned>(j), v); }
};

// base check: an accepted candidate means the node bounds are not a fixpoint
struct WlFlagSink {
  int *flag;
  __device__ __forceinline__ void ub(int, double) { *flag = 1; }
  __device__ __forceinline__ void lb(int, double) { *flag = 1; }
};

// Readers: the bounds a row walk sees.  get() for one column; get_n() for N columns at once with every load of a
// stage issued before the first use (mask words and node bounds first, then the few own cells).
struct WlBaseReader {  // the node's bounds, nothing else (base check)
  const Bnd *baseX;
  __device__ __forceinline__ Bnd get(int j) const
  {
    const double2 v = __ldg(reinterpret_cast<const double2 *>(&baseX[j]));
    Bnd x;
    x.lb = v.x;
    x.ub = v.y;
    return x;
  }
  template <int N>
  __device__ __forceinline__ void get_n(const int (&j)[N], Bnd (&x)[N]) const
  {
#pragma unroll
    for (int u = 0; u < N; ++u)
      x[u] = get(j[u]);
  }
};
struct WlVecReader {  // one vector: the node's bounds unless it has a cell of its own
  const BtParams *p;
  const Bnd *srcCells;
  const unsigned *dm;
  int vl;
  __device__ __forceinline__ Bnd get(int j) const { return wl_bounds(*p, srcCells, dm, vl, j); }
  template <int N>
  __device__ __forceinline__ void get_n(const int (&j)[N], Bnd (&x)[N]) const
  {
    unsigned m[N];
#pragma unroll
    for (int u = 0; u < N; ++u) {
      m[u] = __ldcg(&dm[j[u]]);
      const double2 v = __ldg(reinterpret_cast<const double2 *>(&p->baseX[j[u]]));
      x[u].lb = v.x;
      x[u].ub = v.y;
    }
#pragma unroll
    for (int u = 0; u < N; ++u)
      if ((m[u] >> vl) & 1u)
        x[u] = ldcg_bnd(&srcCells[static_cast<size_t>(j[u]) * 32]);
  }
};

constexpr int kWlChunk = 8;  // nonzeros of a thread-walked row whose loads are in flight together

// One row of at most kShortLen nonzeros by one thread, CSR order (the order of the CPU restatement): its sums.
// kCheck: also whether every sum over the row is exact in any order (integer coefficients and bounds, integer
// columns, magnitudes far below 2^53).
template <bool kCheck, class Reader>
__device__ __forceinline__ RowAct wl_thread_sum(const BtParams &p, int row, const Reader &rd, bool *whole)
{
  const int kb = __ldg(&p.rowNz[row]), cnt = __ldg(&p.rowCnt[row]);
  RowAct ra = { 0.0, 0.0, 0.0, 0, 0 };
  double mag = 0.0;
  bool ok = true;
  for (int k = 0; k < cnt; k += kWlChunk) {
    double a[kWlChunk];
    int cj[kWlChunk], jj[kWlChunk];
    Bnd x[kWlChunk];
#pragma unroll
    for (int u = 0; u < kWlChunk; ++u) {
      a[u] = 0.0;
      cj[u] = 0;
      if (k + u < cnt) {
        cj[u] = __ldg(&p.col[kb + k + u]);
        a[u] = __ldg(&p.val[kb + k + u]);
      }
      jj[u] = col_of(cj[u]);
    }
    rd.get_n(jj, x);
#pragma unroll
    for (int u = 0; u < kWlChunk; ++u)
      if (a[u] != 0.0) {
        accumulate_fast(ra, a[u], col_is_int(cj[u]), x[u], p.s.inf);
        if (kCheck) {
          mag += fabs(a[u]) * fmax(fabs(x[u].lb), fabs(x[u].ub));
          ok = ok && col_is_int(cj[u]) && a[u] == floor(a[u]) && x[u].lb == floor(x[u].lb) && x[u].ub == floor(x[u].ub);
        }
      }
  }
  if (kCheck)
    *whole = ok && mag < 1.0e15 && ra.infMin + ra.infMax == 0;
  return ra;
}

template <class Reader, class Sink>
__device__ __forceinline__ void wl_thread_tighten(const BtParams &p, int row, const RowPlan &pl, const Reader &rd, Sink &sink)
{
  const int kb = __ldg(&p.rowNz[row]), cnt = __ldg(&p.rowCnt[row]);
  for (int k = 0; k < cnt; k += kWlChunk) {
    double a[kWlChunk];
    int cj[kWlChunk], jj[kWlChunk];
    Bnd x[kWlChunk];
#pragma unroll
    for (int u = 0; u < kWlChunk; ++u) {
      a[u] = 0.0;
      cj[u] = 0;
      if (k + u < cnt) {
        cj[u] = __ldg(&p.col[kb + k + u]);
        a[u] = __ldg(&p.val[kb + k + u]);
      }
      jj[u] = col_of(cj[u]);
    }
    rd.get_n(jj, x);
#pragma unroll
    for (int u = 0; u < kWlChunk; ++u)
      if (a[u] != 0.0)
        tighten_nz(sink, pl, a[u], jj[u], col_is_int(cj[u]), a[u] > 0.0 ? x[u].lb : x[u].ub,
          a[u] > 0.0 ? x[u].ub : x[u].lb, p.s);
  }
}

// Second pass of the short-row items of one warp, nonzero-parallel: every lane may hold one (vector, row) item whose
// row can tighten (want); the warp takes as many of those items as fit into 32 lanes at a time — an item with cnt
// nonzeros gets cnt consecutive lanes — and each lane evaluates one nonzero of "its" item with that item's plan (fetched
// by shuffle).  One copy of the tightening code, all lanes busy, instead of every lane walking its own row through it.
// Called by all 32 lanes together.
__device__ __forceinline__ void wl_tighten_packed(const BtParams &p, const Bnd *srcX, bool want, int b, int row, int cnt,
  const RowPlan &pl, BtLogEntry *log, unsigned *nLog, int *overflow, BtLogEntry *stage, unsigned *staged, int lane)
{
  unsigned todo = __ballot_sync(0xffffffffu, want);
  if (!todo)
    return;
  const int kb = want ? __ldg(&p.rowNz[row]) : 0;
  const int packed = min(pl.infMin, 2) | (min(pl.infMax, 2) << 2) | (pl.sideU ? 16 : 0) | (pl.sideL ? 32 : 0);
  while (todo) {
    // the items of this trip: lane ranges [start, start + cnt) handed out in the order of the set bits
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
    const int src_l = owner >= 0 ? owner : 0;
    const int ob = __shfl_sync(0xffffffffu, b, src_l);
    const int okb = __shfl_sync(0xffffffffu, kb, src_l);
    RowPlan q;
    q.slackU = __shfl_sync(0xffffffffu, pl.slackU, src_l);
    q.slackL = __shfl_sync(0xffffffffu, pl.slackL, src_l);
    const int pk = __shfl_sync(0xffffffffu, packed, src_l);
    q.infMin = pk & 3;
    q.infMax = (pk >> 2) & 3;
    q.sideU = (pk & 16) != 0;
    q.sideL = (pk & 32) != 0;
    q.infeasible = false;
    if (owner >= 0) {
      const int c = __ldg(&p.col[okb + ent]);
      const double a = __ldg(&p.val[okb + ent]);
      const Bnd x = wl_bounds(p, srcX + bt_base32(p, ob), p.dmask + bt_word(p, 0, ob >> 5), ob & 31, col_of(c));
      WlLogSink sink = { log, nLog, p.logCap, overflow, ob, stage, staged };
      if (a != 0.0)
        tighten_nz(sink, q, a, col_of(c), col_is_int(c), a > 0.0 ? x.lb : x.ub, a > 0.0 ? x.ub : x.lb, p.s);
    }
  }
}

// One longer row by one warp: lanes stride over the row (four nonzeros per lane in flight), butterfly.
template <class Reader>
__device__ __forceinline__ RowAct wl_row_warp_sum(const BtParams &p, int row, const Reader &rd, int lane)
{
  const int kb = __ldg(&p.rowNz[row]), ke = kb + __ldg(&p.rowCnt[row]);
  RowAct ra = { 0.0, 0.0, 0.0, 0, 0 };
  for (int k0 = kb; k0 < ke; k0 += 128) {
    double a[4];
    int c[4], jj[4];
    Bnd x[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int k = k0 + 32 * u + lane;
      a[u] = 0.0;
      c[u] = 0;
      if (k < ke) {
        c[u] = __ldg(&p.col[k]);
        a[u] = __ldg(&p.val[k]);
      }
      jj[u] = col_of(c[u]);
    }
    rd.get_n(jj, x);
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (a[u] != 0.0)
        accumulate(ra, a[u], col_is_int(c[u]), a[u] > 0.0 ? x[u].lb : x[u].ub, a[u] > 0.0 ? x[u].ub : x[u].lb, p.s.inf);
  }
  warp_reduce(ra);
  return ra;
}

// the same with the exactness check of the base pass: magnitude of the row's terms, and whether they are all integers
template <class Reader>
__device__ __forceinline__ RowAct wl_row_warp_sum_checked(const BtParams &p, int row, const Reader &rd, int lane, double *magOut,
  bool *wholeOut)
{
  const int kb = __ldg(&p.rowNz[row]), ke = kb + __ldg(&p.rowCnt[row]);
  RowAct ra = { 0.0, 0.0, 0.0, 0, 0 };
  double mag = 0.0;
  bool whole = true;
  for (int k0 = kb; k0 < ke; k0 += 128) {
    double a[4];
    int c[4], jj[4];
    Bnd x[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int k = k0 + 32 * u + lane;
      a[u] = 0.0;
      c[u] = 0;
      if (k < ke) {
        c[u] = __ldg(&p.col[k]);
        a[u] = __ldg(&p.val[k]);
      }
      jj[u] = col_of(c[u]);
    }
    rd.get_n(jj, x);
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (a[u] != 0.0) {
        accumulate(ra, a[u], col_is_int(c[u]), a[u] > 0.0 ? x[u].lb : x[u].ub, a[u] > 0.0 ? x[u].ub : x[u].lb, p.s.inf);
        mag += fabs(a[u]) * fmax(fabs(x[u].lb), fabs(x[u].ub));
        whole = whole && col_is_int(c[u]) && a[u] == floor(a[u]) && x[u].lb == floor(x[u].lb) && x[u].ub == floor(x[u].ub);
      }
  }
  warp_reduce(ra);
#pragma unroll
  for (int m = 16; m >= 1; m >>= 1)
    mag += __shfl_xor_sync(0xffffffffu, mag, m);
  *magOut = mag;
  *wholeOut = __all_sync(0xffffffffu, whole);
  return ra;
}

template <class Reader, class Sink>
__device__ __forceinline__ void wl_row_warp_tighten(const BtParams &p, int row, const RowPlan &pl, const Reader &rd,
  Sink &sink, int lane)
{
  const int kb = __ldg(&p.rowNz[row]), ke = kb + __ldg(&p.rowCnt[row]);
  for (int k0 = kb; k0 < ke; k0 += 128) {
    double a[4];
    int c[4], jj[4];
    Bnd x[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int k = k0 + 32 * u + lane;
      a[u] = 0.0;
      c[u] = 0;
      if (k < ke) {
        c[u] = __ldg(&p.col[k]);
        a[u] = __ldg(&p.val[k]);
      }
      jj[u] = col_of(c[u]);
    }
    rd.get_n(jj, x);
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (a[u] != 0.0)
        tighten_nz(sink, pl, a[u], jj[u], col_is_int(c[u]), a[u] > 0.0 ? x[u].lb : x[u].ub,
          a[u] > 0.0 ? x[u].ub : x[u].lb, p.s);
  }
}

// Cells a warp takes per trip when n cells are to be listed: few enough that every warp of the grid gets some (the
// rows of a warp's cells are dealt out to all 32 lanes anyway, and a trip's cost is its chain of round trips to L2,
// not its width), a power of two, at most 32.
__device__ __forceinline__ unsigned wl_cells_per_warp(unsigned n, unsigned nwarps)
{
  const unsigned c = (n + nwarps - 1u) / nwarps;
  if (c <= 1u)
    return 1u;
  if (c >= 32u)
    return 32u;
  return 1u << (32 - __clz(c - 1u));
}

// One cell of a warp's 32 in wl_emit_rows (shared memory): where the column's rows start, how many rows the cells of
// the lanes below have, whose cell it is, what the change adds to a bound.
struct __align__(32) WlEmitCell {
  int cs, excl, b, pad;
  double dLb, dUb;
};

// Cell (b, j) went from bounds `was` to bounds `now` (lanes with active == false have no cell): the rows of column j
// become items of vector b in the lists of parity `nxt` (round `forRound`).  Called by all 32 lanes of a warp
// together, each with its own cell; `se` = 32 records of shared memory that belong to the warp.  Rows of at most
// kShortLen nonzeros go to the short list as (b, row) — unless sums are exact (w.exactOk), when every row goes the way
// of the longer rows: the pair's table slot takes what the change adds to the row's two sums, and the pair is listed
// (by slot) unless it already is for that round.
// The 32 columns' rows are laid end to end and dealt out to the lanes eight at a time: the number of trips (each a
// chain of round trips to L2 — probe, add, stamp, reserve) follows the columns' average length, not the longest one.
__device__ __forceinline__ void wl_emit_rows(const WlParams &w, bool active, int b, int j, int nxt, int forRound,
  const Bnd &was, const Bnd &now, int lane, WlEmitCell *se)
{
  const int cs = active ? __ldg(&w.cscStart[j]) : 0, ce = active ? __ldg(&w.cscStart[j + 1]) : 0;
  int incl = ce - cs;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d)
      incl += t;
  }
  const int total = __shfl_sync(0xffffffffu, incl, 31);
  if (total == 0)
    return;
  __syncwarp();
  {
    WlEmitCell e;
    e.cs = cs;
    e.excl = incl - (ce - cs);
    e.b = b;
    e.pad = 0;
    e.dLb = now.lb - was.lb;
    e.dUb = now.ub - was.ub;
    se[lane] = e;
  }
  __syncwarp();
  for (int base = 0; base < total; base += 256) {
    int r[8], q[8];
    unsigned char own[8];
    unsigned slot[8];
    unsigned have = 0u;
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int idx = base + 32 * u + lane;
      r[u] = -1;
      q[u] = 0;
      own[u] = 0;
      if (idx < total) {
        int o = 0;  // the cell: the last one with excl <= idx
#pragma unroll
        for (int st = 16; st >= 1; st >>= 1)
          if (se[o + st].excl <= idx)
            o += st;
        own[u] = static_cast<unsigned char>(o);
        q[u] = se[o].cs + (idx - se[o].excl);
        have |= 1u << u;
      }
    }
#pragma unroll
    for (int u = 0; u < 8; ++u)
      if ((have >> u) & 1u)
        r[u] = __ldg(&w.cscRow[q[u]]);
    unsigned shortRows = 0u;  // rows for the short list (with exact sums every row goes through the table)
    if (!w.exactOk) {
      int c[8];
#pragma unroll
      for (int u = 0; u < 8; ++u)
        c[u] = ((have >> u) & 1u) ? __ldg(&w.bt.rowCnt[r[u]]) : 0x7fffffff;
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if (c[u] <= kShortLen)
          shortRows |= 1u << u;
    }
    const int nS = __popc(shortRows);
    // The pairs' table slots (linear probing): every probe step is one compare-and-swap per row, and the eight rows'
    // steps go out together before any answer is waited for — one after the other they would be a chain of round trips.
    unsigned pend = have & ~shortRows, fresh = 0u;
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      slot[u] = ~0u;
      if ((pend >> u) & 1u) {
        const unsigned long long key
          = ((static_cast<unsigned long long>(static_cast<unsigned>(se[own[u]].b)) << 32) | static_cast<unsigned>(r[u])) + 1ull;
        slot[u] = static_cast<unsigned>((key * 0x9E3779B97F4A7C15ull) >> w.hShift);
      }
    }
    unsigned tbl = pend;  // rows that go through the table
    for (int probe = 0; probe < 128 && pend; ++probe) {
      unsigned long long seen[8];
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if ((pend >> u) & 1u)
          seen[u] = atomicCAS(&w.tab[slot[u]].key, 0ull,
            ((static_cast<unsigned long long>(static_cast<unsigned>(se[own[u]].b)) << 32) | static_cast<unsigned>(r[u])) + 1ull);
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if ((pend >> u) & 1u) {
          const unsigned long long key
            = ((static_cast<unsigned long long>(static_cast<unsigned>(se[own[u]].b)) << 32) | static_cast<unsigned>(r[u])) + 1ull;
          if (seen[u] == 0ull || seen[u] == key) {
            pend &= ~(1u << u);
            if (seen[u] == 0ull)
              fresh |= 1u << u;
          } else {
            slot[u] = (slot[u] + 1u) & w.hMask;
          }
        }
    }
    if (pend) {  // the table is too full
      w.ctl->overflow = 1;
      tbl &= ~pend;
    }
    const int nNew = __popc(fresh);
    if (w.exactOk) {
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if ((tbl >> u) & 1u) {
          const double a = __ldg(&w.cscVal[q[u]]);
          const double dLb = se[own[u]].dLb, dUb = se[own[u]].dUb;
          const double dMin = a * (a > 0.0 ? dLb : dUb), dMax = a * (a > 0.0 ? dUb : dLb);
          if (dMin != 0.0)
            atomicAdd(&w.tab[slot[u]].dMin, dMin);
          if (dMax != 0.0)
            atomicAdd(&w.tab[slot[u]].dMax, dMax);
        }
    }
    unsigned listed = 0u;  // table rows not yet listed for that round
    {
      int stamp[8];
#pragma unroll
      for (int u = 0; u < 8; ++u)
        stamp[u] = ((tbl >> u) & 1u) ? atomicMax(&w.tab[slot[u]].stamp, forRound + 1) : 0x7fffffff;
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if (stamp[u] < forRound + 1)
          listed |= 1u << u;
    }
    const int nL = __popc(listed);
    // Room in the three lists — short items, table items, newly taken table slots — with one atomic per list and warp;
    // an entry's place follows its place among the warp's rows (row 32 u + lane of this trip), so the lists keep a
    // vector's rows together, which is what the next round's item phase reads fastest.
    int totS = 0, totL = 0, totU = 0;
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      totS += __popc(__ballot_sync(0xffffffffu, (shortRows >> u) & 1u));
      totL += __popc(__ballot_sync(0xffffffffu, (listed >> u) & 1u));
      totU += __popc(__ballot_sync(0xffffffffu, (fresh >> u) & 1u));
    }
    unsigned bS = 0u, bL = 0u, bU = 0u;
    if (lane == 0) {
      if (totS)
        bS = atomicAdd(&w.ctl->nS[nxt], static_cast<unsigned>(totS));
      if (totL)
        bL = atomicAdd(&w.ctl->nL[nxt], static_cast<unsigned>(totL));
      if (totU)
        bU = atomicAdd(&w.ctl->nUsed, static_cast<unsigned>(totU));
    }
    bS = __shfl_sync(0xffffffffu, bS, 0);
    bL = __shfl_sync(0xffffffffu, bL, 0);
    bU = __shfl_sync(0xffffffffu, bU, 0);
    const bool okS = bS + static_cast<unsigned>(totS) <= w.itemCap, okL = bL + static_cast<unsigned>(totL) <= w.itemCap,
               okU = bU + static_cast<unsigned>(totU) <= w.usedCap;
    if (lane == 0) {
      if (!okS || !okL)
        w.ctl->overflow = 1;
      if (!okU)
        w.ctl->usedOverflow = 1;
    }
    const unsigned below = (1u << lane) - 1u;
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const unsigned mS = __ballot_sync(0xffffffffu, (shortRows >> u) & 1u), mL = __ballot_sync(0xffffffffu, (listed >> u) & 1u),
                     mU = __ballot_sync(0xffffffffu, (fresh >> u) & 1u);
      WlItem it;
      it.row = r[u];
      if ((shortRows >> u) & 1u) {
        if (okS) {
          it.b = se[own[u]].b;
          w.itemsS[nxt][bS + __popc(mS & below)] = it;
        }
      } else if ((listed >> u) & 1u) {
        if (okL) {
          it.b = static_cast<int>(slot[u]);
          w.itemsL[nxt][bL + __popc(mL & below)] = it;
        }
      }
      if (((fresh >> u) & 1u) && okU)
        w.used[bU + __popc(mU & below)] = slot[u];
      bS += __popc(mS);
      bL += __popc(mL);
      bU += __popc(mU);
    }
  }
}

// append to a list through one atomic per group of lanes that arrive together; returns the slot or ~0u on overflow
__device__ __forceinline__ unsigned wl_append(unsigned *count, unsigned cap, int *overflow)
{
  const unsigned peers = __activemask();
  const int leader = __ffs(peers) - 1;
  const int laneId = threadIdx.x & 31;
  unsigned base = 0;
  if (laneId == leader)
    base = atomicAdd(count, static_cast<unsigned>(__popc(peers)));
  base = __shfl_sync(peers, base, leader);
  const unsigned at = base + __popc(peers & ((1u << laneId) - 1u));
  if (at >= cap) {
    *overflow = 1;
    return ~0u;
  }
  return at;
}

#ifndef CBP_WL_BLOCKS
#define CBP_WL_BLOCKS 2
#endif

// a finished row of the base check: violated, or able to tighten something, means the node is no fixpoint
template <class Tighten>
__device__ __forceinline__ void wl_base_verdict(const BtParams &p, int row, const RowAct &ra, int *bad, Tighten &&walk)
{
  const RowPlan pl = plan_row(ra, __ldg(&p.rowLB[row]), __ldg(&p.rowUB[row]), p.s);
  if (pl.infeasible)
    *bad = 1;
  else if (pl.sideU || pl.sideL)
    walk(pl);
}

// ---- updatedBounds() of a worklist run, from the list of cells instead of a scan of the `ever` marks ----------------
// count per vector
__device__ __forceinline__ void wl_upd_count(const WlCell c, const BtParams &p, int *cnt)
{
  if (__ldcg(&p.vecStop[c.b]) == 4)
    return;  // a vector proved infeasible lists nothing
  if ((__ldcg(&p.ever[bt_word(p, c.j, c.b >> 5)]) >> (c.b & 31)) & 1u)
    atomicAdd(&cnt[c.b], 1);
}

// (col, lb, ub) of a changed cell into its vector's segment, in the order the threads get there
__device__ __forceinline__ void wl_upd_fill(const WlCell c, const BtParams &p, const long long *start, int *cursor, int *col,
  double *lb, double *ub)
{
  if (__ldcg(&p.vecStop[c.b]) == 4 || !((__ldcg(&p.ever[bt_word(p, c.j, c.b >> 5)]) >> (c.b & 31)) & 1u))
    return;
  const long long at = __ldcg(&start[c.b]) + atomicAdd(&cursor[c.b], 1);
  const Bnd x = ldcg_bnd(&p.X[__ldcg(&p.vecFinal[c.b])][bt_cell(p, c.j, c.b)]);
  col[at] = c.j;
  lb[at] = x.lb;
  ub[at] = x.ub;
}

// A vector's segment into ascending column order (the order updatedBounds() promises); a vector's columns are
// distinct.  Up to 32 entries: one warp, an entry per lane, every entry's place = the number of smaller columns (by
// shuffle).  Longer segments are a CTA's job (wl_upd_sort_cta).
__device__ __forceinline__ void wl_upd_emit(long long at, int c, double l, double u, int *colOut, double *lbOut, double *ubOut,
  WlUpdHead *head, long long headCap)
{
  colOut[at] = c;
  lbOut[at] = l;
  ubOut[at] = u;
  if (at < headCap) {
    WlUpdHead hd;
    hd.col = c;
    hd.pad = 0;
    hd.lb = l;
    hd.ub = u;
    head[at] = hd;
  }
}

constexpr int kWlSortWarp = 256;   // segments up to this length: a warp
constexpr int kWlSortSmem = 2048;  // ... up to this length: a CTA, bitonic in shared memory; longer: a CTA reads the marks in column order

// cols: kWlSortWarp ints of shared memory that belong to the warp
__device__ __forceinline__ void wl_upd_sort_warp(const long long *start, int b, int lane, int *cols, const int *colIn,
  const double *lbIn, const double *ubIn, int *colOut, double *lbOut, double *ubOut, WlUpdHead *head, long long headCap)
{
  const long long s0 = start[b];
  const int k = static_cast<int>(start[b + 1] - s0);
  if (k <= 0 || k > kWlSortWarp)
    return;
  if (k <= 32) {
    const int c = lane < k ? colIn[s0 + lane] : 0x7fffffff;
    int rank = 0;
#pragma unroll
    for (int f = 0; f < 32; ++f)
      rank += __shfl_sync(0xffffffffu, c, f) < c ? 1 : 0;
    if (lane < k)
      wl_upd_emit(s0 + rank, c, lbIn[s0 + lane], ubIn[s0 + lane], colOut, lbOut, ubOut, head, headCap);
    return;
  }
  __syncwarp();
  for (int e = lane; e < k; e += 32)
    cols[e] = colIn[s0 + e];
  __syncwarp();
  for (int e = lane; e < k; e += 32) {
    const int c = cols[e];
    int rank = 0;
    for (int f = 0; f < k; ++f)  // (all lanes read the same word: a broadcast)
      rank += cols[f] < c ? 1 : 0;
    wl_upd_emit(s0 + rank, c, lbIn[s0 + e], ubIn[s0 + e], colOut, lbOut, ubOut, head, headCap);
  }
  __syncwarp();
}

// keys: kWlSortSmem 64-bit words of shared memory; all kThreads threads of the CTA call this together
__device__ __forceinline__ void wl_upd_sort_cta(const BtParams &p, const long long *start, int b, int tid,
  unsigned long long *keys, const int *colIn, const double *lbIn, const double *ubIn, int *colOut, double *lbOut,
  double *ubOut, WlUpdHead *head, long long headCap)
{
  const long long s0 = start[b];
  const int k = static_cast<int>(start[b + 1] - s0);
  if (k <= kWlSortWarp)
    return;
  if (k <= kWlSortSmem) {
    int n2 = 64;
    while (n2 < k)
      n2 <<= 1;
    __syncthreads();
    for (int e = tid; e < n2; e += kThreads)  // (column << 32) | position: the order of the columns, the entry to fetch
      keys[e] = e < k ? (static_cast<unsigned long long>(static_cast<unsigned>(colIn[s0 + e])) << 32) | static_cast<unsigned>(e)
                      : ~0ull;
    __syncthreads();
    for (int span = 2; span <= n2; span <<= 1)
      for (int half = span >> 1; half >= 1; half >>= 1) {
        for (int t = tid; t < (n2 >> 1); t += kThreads) {
          const int lo = ((t / half) * (half << 1)) + (t % half), hi = lo + half;
          const bool up = (lo & span) == 0;
          const unsigned long long x = keys[lo], y = keys[hi];
          if ((x > y) == up) {
            keys[lo] = y;
            keys[hi] = x;
          }
        }
        __syncthreads();
      }
    for (int e = tid; e < k; e += kThreads) {
      const unsigned long long key = keys[e];
      const int from = static_cast<int>(key & 0xffffffffu);
      wl_upd_emit(s0 + e, static_cast<int>(key >> 32), lbIn[s0 + from], ubIn[s0 + from], colOut, lbOut, ubOut, head, headCap);
    }
    __syncthreads();
    return;
  }
  // Longer still: not ordered at all but read off the vector's `ever` marks column by column (the marks of a
  // 32-vector group are one word per column, consecutive columns adjacent): n / kThreads steps whatever the segment's
  // length, where counting places would take length^2 / kThreads.
  {
    unsigned *cnt = reinterpret_cast<unsigned *>(keys);  // [kWarps] warp totals, [kWarps] = the running place
    const unsigned *words = p.ever + bt_word(p, 0, b >> 5);
    const unsigned bit = 1u << (b & 31);
    const Bnd *fin = p.X[p.vecFinal[b]];
    const int lane = tid & 31, warp = tid >> 5;
    __syncthreads();
    if (tid == 0)
      cnt[kWarps] = 0u;
    __syncthreads();
    for (int j0 = 0; j0 < p.n; j0 += kThreads) {
      const int j = j0 + tid;
      const bool has = j < p.n && (words[j] & bit) != 0u;
      const unsigned m = __ballot_sync(0xffffffffu, has);
      if (lane == 0)
        cnt[warp] = __popc(m);
      __syncthreads();
      unsigned at = cnt[kWarps];
      for (int q = 0; q < warp; ++q)
        at += cnt[q];
      if (has) {
        const Bnd x = fin[bt_cell(p, j, b)];
        wl_upd_emit(s0 + at + __popc(m & ((1u << lane) - 1u)), j, x.lb, x.ub, colOut, lbOut, ubOut, head, headCap);
      }
      __syncthreads();
      if (tid == kThreads - 1)
        cnt[kWarps] = at + __popc(m);
      // (the next step's first barrier orders this write before the reads that follow it)
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(kThreads, CBP_WL_BLOCKS) bt_worklist_kernel(const WlParams w)
{
  __shared__ __align__(16) BtLogEntry s_stage[kWlStage];
  static_assert(sizeof(BtLogEntry) * kWlStage >= sizeof(unsigned long long) * kWlSortSmem, "P7 sorts in the staging buffer");
  __shared__ unsigned s_staged, s_base;
  __shared__ RowAct s_row[kWarps];
  __shared__ double s_mag[kWarps];
  __shared__ WlEmitCell s_emit[kWarps][32];
  cg::grid_group grid = cg::this_grid();
  const BtParams &p = w.bt;
  WlCtl *ctl = w.ctl;
  if (threadIdx.x == 0)
    s_staged = 0u;
  __syncthreads();
  const int tid = threadIdx.x, lane = tid & 31;
  const unsigned gtid = blockIdx.x * kThreads + tid, gsize = gridDim.x * kThreads;
  const unsigned gwarp = gtid >> 5, nwarps = gsize >> 5;
  unsigned long long myPairs = 0;

  // ---- phase 0: leftovers of the previous run out of the masks; is the node a propagation fixpoint?
  for (unsigned i = gtid; i < w.prevCells; i += gsize) {
    const WlCell c = w.cells[i];
    const size_t word = bt_word(p, c.j, c.b >> 5);
    const unsigned bit = 1u << (c.b & 31);
    atomicAnd(const_cast<unsigned *>(&p.dmask[word]), ~bit);
    atomicAnd(&p.ever[word], ~bit);
  }
  // (and the table slots it took are emptied: scattered stores that ride along with the matrix read of the base check)
  for (unsigned i = gtid; i < w.prevUsed; i += gsize) {
    uint4 *q = reinterpret_cast<uint4 *>(&w.tab[w.used[i]]);
    q[0] = make_uint4(0u, 0u, 0u, 0u);
    q[1] = make_uint4(0u, 0u, 0u, 0u);
  }
  if (gtid == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    ctl->t0 = t;
  }
  {
    const WlBaseReader rd = { p.baseX };
    WlFlagSink fs = { &ctl->baseBad };
    for (unsigned j = gtid; j < static_cast<unsigned>(p.n); j += gsize) {  // R11 on the node's own bounds
      const Bnd x = rd.get(j);
      double l = x.lb, u = x.ub;
      if (prepass_col(l, u, p.colType[j] != 0, p.s) != 0)
        ctl->baseBad = 1;
    }
    // the node's sums are kept for the rows whose pairs may later be evaluated as "these + the vector's own cells"
    for (unsigned sl = gwarp; sl < static_cast<unsigned>(w.nSlices); sl += nwarps) {  // short rows: lane = row of a slice
      const int off = __ldg(&w.sellOff[sl]), len = (__ldg(&w.sellOff[sl + 1]) - off) >> 5;
      const int r = (sl << 5) + lane;
      const bool valid = r < w.mShort;
      const double *__restrict__ sv = w.sellVal + off + lane;
      const int *__restrict__ sc = w.sellCol + off + lane;
      RowAct ra = { 0.0, 0.0, 0.0, 0, 0 };
      double mag = 0.0;
      bool whole = true;
      for (int k = 0; k < len; k += kWlChunk) {
        double a[kWlChunk];
        int cj[kWlChunk], jj[kWlChunk];
        Bnd x[kWlChunk];
#pragma unroll
        for (int u = 0; u < kWlChunk; ++u) {
          a[u] = 0.0;
          cj[u] = 0;
          if (k + u < len) {
            cj[u] = __ldg(&sc[(k + u) << 5]);
            a[u] = __ldg(&sv[(k + u) << 5]);
          }
          jj[u] = col_of(cj[u]);
        }
        rd.get_n(jj, x);
#pragma unroll
        for (int u = 0; u < kWlChunk; ++u)
          if (a[u] != 0.0) {  // (padding has a == 0)
            accumulate_fast(ra, a[u], col_is_int(cj[u]), x[u], p.s.inf);
            if (w.exactOk) {  // (only then does anybody ask whether the row's sums are exact)
              mag += fabs(a[u]) * fmax(fabs(x[u].lb), fabs(x[u].ub));
              whole = whole && col_is_int(cj[u]) && a[u] == floor(a[u]) && x[u].lb == floor(x[u].lb) && x[u].ub == floor(x[u].ub);
            }
          }
      }
      if (!valid)
        continue;
      if (w.exactOk) {
        WlRowBase rbw;
        rbw.mn = ra.minAct;
        rbw.mx = ra.maxAct;
        rbw.range = ra.maxRange;
        rbw.exact = (whole && mag < 1.0e15 && ra.infMin + ra.infMax == 0) ? 1 : 0;
        rbw.pad = 0;
        w.rowBase[r] = rbw;
      }
      wl_base_verdict(p, r, ra, &ctl->baseBad, [&](const RowPlan &pl) { wl_thread_tighten(p, r, pl, rd, fs); });
    }
    // (kept order: the rows of at most kShortLen nonzeros come first)
    for (unsigned r = w.mShort + gwarp; r < static_cast<unsigned>(p.mKept); r += nwarps) {  // rows of 33 .. kBtLongLen nonzeros: a warp each
      const int cnt = __ldg(&p.rowCnt[r]);
      if (cnt <= kShortLen || cnt > kBtLongLen)
        continue;
      bool whole = false;
      double mag = 0.0;
      const RowAct ra = wl_row_warp_sum_checked(p, r, rd, lane, &mag, &whole);
      whole = whole && mag < 1.0e15 && ra.infMin + ra.infMax == 0;
      if (lane == 0) {
        WlRowBase rbw;
        rbw.mn = ra.minAct;
        rbw.mx = ra.maxAct;
        rbw.range = ra.maxRange;
        rbw.exact = whole ? 1 : 0;
        rbw.pad = 0;
        w.rowBase[r] = rbw;
      }
      wl_base_verdict(p, r, ra, &ctl->baseBad, [&](const RowPlan &pl) { wl_row_warp_tighten(p, r, pl, rd, fs, lane); });
    }
    for (int i = blockIdx.x; i < p.nLongBt; i += gridDim.x) {  // longer rows: the whole CTA (a warp alone would be the phase's tail)
      const int r = __ldg(&p.longRowsBt[i]);
      const int kb = __ldg(&p.rowNz[r]), ke = kb + __ldg(&p.rowCnt[r]);
      RowAct ra = { 0.0, 0.0, 0.0, 0, 0 };
      double mag = 0.0;
      bool whole = true;
      for (int k = kb + tid; k < ke; k += kThreads) {
        const int c = __ldg(&p.col[k]);
        const double a = __ldg(&p.val[k]);
        const Bnd x = rd.get(col_of(c));
        accumulate(ra, a, col_is_int(c), a > 0.0 ? x.lb : x.ub, a > 0.0 ? x.ub : x.lb, p.s.inf);
        mag += fabs(a) * fmax(fabs(x.lb), fabs(x.ub));
        whole = whole && col_is_int(c) && a == floor(a) && x.lb == floor(x.lb) && x.ub == floor(x.ub);
      }
      warp_reduce(ra);
#pragma unroll
      for (int m = 16; m >= 1; m >>= 1)
        mag += __shfl_xor_sync(0xffffffffu, mag, m);
      whole = __all_sync(0xffffffffu, whole);
      __syncthreads();
      if (lane == 0) {
        s_row[tid >> 5] = ra;
        s_mag[tid >> 5] = whole ? mag : 1.0e300;
      }
      __syncthreads();
      RowAct tot = s_row[0];
      double magTot = s_mag[0];
      for (int q = 1; q < kWarps; ++q) {  // (every thread alike, warp order)
        combine(tot, s_row[q]);
        magTot += s_mag[q];
      }
      if (tid == 0) {
        WlRowBase rbw;
        rbw.mn = tot.minAct;
        rbw.mx = tot.maxAct;
        rbw.range = tot.maxRange;
        rbw.exact = (magTot < 1.0e15 && tot.infMin + tot.infMax == 0) ? 1 : 0;
        rbw.pad = 0;
        w.rowBase[r] = rbw;
      }
      wl_base_verdict(p, r, tot, &ctl->baseBad, [&](const RowPlan &pl) {
        for (int k = kb + tid; k < ke; k += kThreads) {
          const int c = __ldg(&p.col[k]);
          const double a = __ldg(&p.val[k]);
          const Bnd x = rd.get(col_of(c));
          tighten_nz(fs, pl, a, col_of(c), col_is_int(c), a > 0.0 ? x.lb : x.ub, a > 0.0 ? x.ub : x.lb, p.s);
        }
      });
    }
  }
  grid.sync();
  wl_stamp(ctl, gtid);
  if (*reinterpret_cast<volatile int *>(&ctl->baseBad)) {
    if (gtid == 0)
      ctl->status = kWlBaseNotFixpoint;
    return;
  }

  // ---- phase 0b: the diffs, one lane per diff over all vectors' diffs laid end to end (a vector with one diff — a
  // strong-branching arm — costs a lane, not a warp).  Applying a vector's diffs in order means: a bound ends at the
  // LAST diff that names it.  The lane of the FIRST diff of a column owns the cell: it folds the vector's later diffs
  // of that column into the node's bounds, gives the result R11, creates the cell (mask bit, both buffers) and lists
  // it; the column's rows become items of the first round once it is known which vectors already cross (below).
  {
    const long long nDiff = w.dStart[w.B];
    unsigned *dmaskW = const_cast<unsigned *>(p.dmask);
    for (long long kk = static_cast<long long>(gwarp) << 5; kk < nDiff; kk += static_cast<long long>(nwarps) << 5) {
      const long long k = kk + lane;
      if (k >= nDiff)
        continue;
      int lo = 0, hi = w.B;  // the vector: the last b with dStart[b] <= k
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(&w.dStart[mid]) <= k)
          lo = mid;
        else
          hi = mid;
      }
      const int b = lo;
      const long long k0 = __ldg(&w.dStart[b]), k1 = __ldg(&w.dStart[b + 1]);
      const unsigned v = __ldg(&w.dVar[k]);
      const int j = diff_col(v);
      if (j >= p.n)
        continue;
      bool first = true;
      for (long long q = k0; q < k && first; ++q)
        first = diff_col(__ldg(&w.dVar[q])) != j;
      if (!first)
        continue;
      const Bnd nb = p.baseX[j];
      Bnd x0 = nb;
      bool inexact = false;
      for (long long q = k; q < k1; ++q) {
        const unsigned vq = q == k ? v : __ldg(&w.dVar[q]);
        if (diff_col(vq) != j)
          continue;
        const double nv = __ldg(&w.dBnd[q]) + 0.0;
        // exact sums need every bound of the vector to be an integer inside the node's bounds
        inexact = inexact || !(nv == floor(nv) && nv >= nb.lb && nv <= nb.ub);
        if (vq & kDiffUpper)
          x0.ub = nv;
        else
          x0.lb = nv;
      }
      if (inexact)
        ctl->inexact = 1;
      Bnd x = x0;
      const int r = prepass_col(x.lb, x.ub, p.colType[j] != 0, p.s);  // R11 on the cell
      const size_t at = bt_cell(p, j, b);
      const unsigned bit = 1u << (b & 31);
      if (r != 1)
        x = x0;
      p.X[0][at] = x;
      p.X[1][at] = x;
      atomicOr(&dmaskW[bt_word(p, j, b >> 5)], bit);
      const unsigned ci = wl_append(&ctl->nCells, w.cellCap, &ctl->overflow);
      if (ci != ~0u) {
        WlCell c;
        c.b = b;
        c.j = j;
        w.cells[ci] = c;
      }
      if (r == 2) {
        atomicMin(&p.vecInfCol[b], j);  // the input already crosses: the vector stops before its first round
      } else if (r == 1) {
        const unsigned old = atomicOr(&p.dirty[0][bt_word(p, j, b >> 5)], bit);
        if (!(old & bit)) {
          const unsigned ki = wl_append(&ctl->nK, w.kCap, &ctl->overflow);
          if (ki != ~0u) {
            WlCell c;
            c.b = b;
            c.j = j;
            w.K[ki] = c;
          }
        }
      }
    }
    grid.sync();
    for (unsigned b = gtid; b < static_cast<unsigned>(w.B); b += gsize)
      if (__ldcg(&p.vecInfCol[b]) != 0x7fffffff) {
        p.vecStop[b] = 4;
        p.vecRounds[b] = 1;
        p.vecFinal[b] = 0;
        atomicAnd(&w.active2[0][b >> 5], ~(1u << (b & 31)));
        atomicAdd(&ctl->roundsTotal, 1ull);
      }
    const unsigned nNew = min(*reinterpret_cast<volatile unsigned *>(&ctl->nCells), w.cellCap);
    const unsigned cpwNew = wl_cells_per_warp(nNew, nwarps);
    for (unsigned i0 = gwarp * cpwNew; i0 < nNew; i0 += nwarps * cpwNew) {  // (whole warps: the listing is a warp's job)
      const unsigned i = lane < cpwNew ? i0 + lane : nNew;
      bool go = false;
      WlCell c;
      c.b = 0;
      c.j = 0;
      Bnd was, now;
      was.lb = was.ub = now.lb = now.ub = 0.0;
      if (i < nNew) {
        c = w.cells[i];
        if (__ldcg(&p.vecInfCol[c.b]) == 0x7fffffff) {
          go = true;
          was = p.baseX[c.j];
          now = ldcg_bnd(&p.X[0][bt_cell(p, c.j, c.b)]);
        }
      }
      wl_emit_rows(w, go, c.b, c.j, 0, 0, was, now, lane, s_emit[tid >> 5]);
    }
  }
  grid.sync();
  wl_stamp(ctl, gtid);

  // ---- the rounds
  int round = 0;
  int status = kWlOk;
  for (;; ++round) {
    const int cur = round & 1, nxt = cur ^ 1;
    if (*reinterpret_cast<volatile int *>(&ctl->overflow)) {  // (written in phase 0b, P2, P3, P6: none of them runs now)
      status = kWlOverflow;
      break;
    }
    const unsigned nS = *reinterpret_cast<volatile unsigned *>(&ctl->nS[cur]);
    const unsigned nL = *reinterpret_cast<volatile unsigned *>(&ctl->nL[cur]);
    const Bnd *srcX = p.X[cur];
    // P1a: short-row items: the sums one thread per item (CSR order), the tightening nonzero-parallel per warp
    for (unsigned i0 = gwarp << 5; i0 < nS; i0 += nwarps << 5) {
      const unsigned i = i0 + lane;
      const bool have = i < nS;
      WlItem it;
      it.b = 0;
      it.row = 0;
      if (have)
        it = w.itemsS[cur][i];
      const int cnt = have ? __ldg(&p.rowCnt[it.row]) : 0;
      RowPlan pl;
      pl.slackU = pl.slackL = 0.0;
      pl.infMin = pl.infMax = 0;
      pl.sideU = pl.sideL = pl.infeasible = false;
      if (have) {
        const WlVecReader rd = { &p, srcX + bt_base32(p, it.b), p.dmask + bt_word(p, 0, it.b >> 5), it.b & 31 };
        const RowAct ra = wl_thread_sum<false>(p, it.row, rd, nullptr);
        pl = plan_row(ra, __ldg(&p.rowLB[it.row]), __ldg(&p.rowUB[it.row]), p.s);
        if (pl.infeasible)
          atomicMin(&p.vecInfRow[it.b], __ldg(&p.rowOrig[it.row]));
        myPairs += cnt;
      }
      wl_tighten_packed(p, srcX, have && !pl.infeasible && (pl.sideU || pl.sideL), it.b, it.row, cnt, pl, p.log, &ctl->nLog,
        &ctl->overflowP1, s_stage, &s_staged, lane);
    }
    // P1b: pairs listed by table slot, one lane each; what needs a warp (a long row's walk) is done by the warp,
    // one pair after the other
    const bool exact = w.exactOk && !*reinterpret_cast<volatile int *>(&ctl->inexact);
    // (neighbours in the list are one vector's rows and tend to need the same amount of work: a warp takes items that
    // lie nwarps apart, so the expensive ones spread over all warps)
    for (unsigned t0 = 0; t0 * nwarps < nL; t0 += 32) {
      const unsigned i = (t0 + lane) * nwarps + gwarp;
      const bool have = i < nL;
      WlItem it;
      it.b = 0;
      it.row = 0;
      if (have)
        it = w.itemsL[cur][i];
      const unsigned slot = static_cast<unsigned>(it.b);
      // (the pair's slot: key | dMin in one 16-byte load, dMax | stamp in the other — one sector)
      ulonglong2 pk = make_ulonglong2(1ull, 0ull);
      double2 pm = make_double2(0.0, 0.0);
      if (have) {
        pk = __ldcg(reinterpret_cast<const ulonglong2 *>(&w.tab[slot]));
        pm = __ldcg(reinterpret_cast<const double2 *>(&w.tab[slot].dMax));
      }
      const int b = static_cast<int>((pk.x - 1ull) >> 32);
      const int row = it.row;
      const int cnt = have ? __ldg(&p.rowCnt[row]) : 0;
      const WlVecReader rd = { &p, srcX + bt_base32(p, b), p.dmask + bt_word(p, 0, b >> 5), b & 31 };
      RowAct ra = { 0.0, 0.0, 0.0, 0, 0 };
      bool needSum = have;
      bool have_plan = have;  // (false once the pair has gone to the long-row queue)
      if (have && exact) {
        const double2 r0 = __ldcg(reinterpret_cast<const double2 *>(&w.rowBase[row]));
        const double2 r1 = __ldcg(reinterpret_cast<const double2 *>(&w.rowBase[row].range));
        if (__double2hiint(r1.y) != 0 || __double2loint(r1.y) != 0) {  // (exact | pad as the second word pair: exact != 0)
          ra.minAct = r0.x + __longlong_as_double(static_cast<long long>(pk.y));
          ra.maxAct = r0.y + pm.x;
          ra.maxRange = r1.x;  // (the node's: an upper bound of the vector's, all R6's shortcut needs)
          needSum = false;
        }
      }
      if (needSum && cnt <= kShortLen) {
        ra = wl_thread_sum<false>(p, row, rd, nullptr);
        myPairs += cnt;
        needSum = false;
      }
      if (needSum && cnt > kBtLongLen) {  // too long for a warp: a CTA walks it after the items (sums, then the rest)
        const unsigned qi = wl_append(&ctl->nWalk, w.walkCap, &ctl->overflowP1);
        if (qi != ~0u) {
          WlWalk e;
          e.b = b;
          e.row = row;
          e.needSum = 1;
          e.packed = 0;
          e.slackU = e.slackL = 0.0;
          w.walk[qi] = e;
        }
        needSum = false;
        have_plan = false;
      }
      unsigned todo = __ballot_sync(0xffffffffu, needSum);
      while (todo) {
        const int l = __ffs(todo) - 1;
        todo &= todo - 1;
        const int rb = __shfl_sync(0xffffffffu, b, l), rr = __shfl_sync(0xffffffffu, row, l);
        const WlVecReader rd2 = { &p, srcX + bt_base32(p, rb), p.dmask + bt_word(p, 0, rb >> 5), rb & 31 };
        const RowAct r2 = wl_row_warp_sum(p, rr, rd2, lane);
        if (lane == l) {
          ra = r2;
          myPairs += cnt;
        }
      }
      RowPlan pl = plan_row(ra, have ? __ldg(&p.rowLB[row]) : 0.0, have ? __ldg(&p.rowUB[row]) : 0.0, p.s);
      if (!have_plan) {
        pl.infeasible = false;
        pl.sideU = false;
        pl.sideL = false;
      }
      if (pl.infeasible)
        atomicMin(&p.vecInfRow[b], __ldg(&p.rowOrig[row]));
      bool tight = !pl.infeasible && (pl.sideU || pl.sideL);
      if (tight && cnt > kBtLongLen) {  // the sums were exact and cheap, the tightening walk is not: queue it with its plan
        const unsigned qi = wl_append(&ctl->nWalk, w.walkCap, &ctl->overflowP1);
        if (qi != ~0u) {
          WlWalk e;
          e.b = b;
          e.row = row;
          e.needSum = 0;
          e.packed = min(pl.infMin, 2) | (min(pl.infMax, 2) << 2) | (pl.sideU ? 16 : 0) | (pl.sideL ? 32 : 0);
          e.slackU = pl.slackU;
          e.slackL = pl.slackL;
          w.walk[qi] = e;
        }
        tight = false;
      }
      {  // rows of at most kShortLen nonzeros: nonzero-parallel, several items per trip
        const bool shortTight = tight && cnt <= kShortLen;
        wl_tighten_packed(p, srcX, shortTight, b, row, cnt, pl, p.log, &ctl->nLog, &ctl->overflowP1, s_stage, &s_staged, lane);
        if (shortTight) {
          myPairs += cnt;
          tight = false;
        }
      }
      todo = __ballot_sync(0xffffffffu, tight);
      while (todo) {
        const int l = __ffs(todo) - 1;
        todo &= todo - 1;
        const int rb = __shfl_sync(0xffffffffu, b, l), rr = __shfl_sync(0xffffffffu, row, l);
        RowPlan q;
        q.slackU = __shfl_sync(0xffffffffu, pl.slackU, l);
        q.slackL = __shfl_sync(0xffffffffu, pl.slackL, l);
        q.infMin = __shfl_sync(0xffffffffu, pl.infMin, l);
        q.infMax = __shfl_sync(0xffffffffu, pl.infMax, l);
        q.sideU = __shfl_sync(0xffffffffu, pl.sideU ? 1 : 0, l) != 0;
        q.sideL = __shfl_sync(0xffffffffu, pl.sideL ? 1 : 0, l) != 0;
        q.infeasible = false;
        const WlVecReader rd2 = { &p, srcX + bt_base32(p, rb), p.dmask + bt_word(p, 0, rb >> 5), rb & 31 };
        WlLogSink sink = { p.log, &ctl->nLog, p.logCap, &ctl->overflowP1, rb, s_stage, &s_staged };
        wl_row_warp_tighten(p, rr, q, rd2, sink, lane);
        if (lane == l)
          myPairs += cnt;
      }
    }
    // the CTA's staged candidates go to the global list: one reservation
    __syncthreads();
    {
      const unsigned n = min(s_staged, kWlStage);
      if (tid == 0 && n)
        s_base = atomicAdd(&ctl->nLog, n);
      __syncthreads();
      for (unsigned i = tid; i < n; i += kThreads) {
        if (s_base + i < p.logCap)
          p.log[s_base + i] = s_stage[i];
        else
          ctl->overflowP1 = 1;
      }
      __syncthreads();
      if (tid == 0)
        s_staged = 0u;
    }
    if (gtid == 0) {
      ctl->itemsTotal += nS + nL;
      ctl->anyActive[nxt] = 0;
    }
    grid.sync();
    wl_stamp(ctl, gtid);
    if (*reinterpret_cast<volatile int *>(&ctl->overflowP1)) {  // (P1 is over: nobody writes this word now)
      status = kWlOverflow;
      break;
    }
    const unsigned nWalk = *reinterpret_cast<volatile unsigned *>(&ctl->nWalk);
    if (nWalk > 0) {
      // P1c: the long rows that have to be walked, a CTA per (vector, row) pair
      for (unsigned e = blockIdx.x; e < nWalk; e += gridDim.x) {
        const WlWalk q = w.walk[e];
        const WlVecReader rd = { &p, srcX + bt_base32(p, q.b), p.dmask + bt_word(p, 0, q.b >> 5), q.b & 31 };
        const int kb = __ldg(&p.rowNz[q.row]), ke = kb + __ldg(&p.rowCnt[q.row]);
        RowPlan pl;
        pl.slackU = q.slackU;
        pl.slackL = q.slackL;
        pl.infMin = q.packed & 3;
        pl.infMax = (q.packed >> 2) & 3;
        pl.sideU = (q.packed & 16) != 0;
        pl.sideL = (q.packed & 32) != 0;
        pl.infeasible = false;
        if (q.needSum) {
          RowAct ra = { 0.0, 0.0, 0.0, 0, 0 };
          for (int k0 = kb; k0 < ke; k0 += 2 * kThreads) {  // two nonzeros per thread in flight
            double a[2];
            int c[2], jj[2];
            Bnd x[2];
#pragma unroll
            for (int u = 0; u < 2; ++u) {
              const int k = k0 + u * kThreads + tid;
              a[u] = 0.0;
              c[u] = 0;
              if (k < ke) {
                c[u] = __ldg(&p.col[k]);
                a[u] = __ldg(&p.val[k]);
              }
              jj[u] = col_of(c[u]);
            }
            rd.get_n(jj, x);
#pragma unroll
            for (int u = 0; u < 2; ++u)
              if (a[u] != 0.0)
                accumulate(ra, a[u], col_is_int(c[u]), a[u] > 0.0 ? x[u].lb : x[u].ub, a[u] > 0.0 ? x[u].ub : x[u].lb, p.s.inf);
          }
          warp_reduce(ra);
          __syncthreads();
          if (lane == 0)
            s_row[tid >> 5] = ra;
          __syncthreads();
          RowAct tot = s_row[0];
          for (int q2 = 1; q2 < kWarps; ++q2)  // (every thread alike, warp order)
            combine(tot, s_row[q2]);
          pl = plan_row(tot, __ldg(&p.rowLB[q.row]), __ldg(&p.rowUB[q.row]), p.s);
          if (pl.infeasible && tid == 0)
            atomicMin(&p.vecInfRow[q.b], __ldg(&p.rowOrig[q.row]));
          if (tid == 0)
            myPairs += ke - kb;
        }
        if (!pl.infeasible && (pl.sideU || pl.sideL)) {
          WlLogSink sink = { p.log, &ctl->nLog, p.logCap, &ctl->overflowP1, q.b, s_stage, &s_staged };
          for (int k = kb + tid; k < ke; k += kThreads) {
            const int c = __ldg(&p.col[k]);
            const double a = __ldg(&p.val[k]);
            const Bnd x = rd.get(col_of(c));
            tighten_nz(sink, pl, a, col_of(c), col_is_int(c), a > 0.0 ? x.lb : x.ub, a > 0.0 ? x.ub : x.lb, p.s);
          }
          if (tid == 0)
            myPairs += ke - kb;
        }
        __syncthreads();
      }
      __syncthreads();
      {
        const unsigned n = min(s_staged, kWlStage);
        if (tid == 0 && n)
          s_base = atomicAdd(&ctl->nLog, n);
        __syncthreads();
        for (unsigned i = tid; i < n; i += kThreads) {
          if (s_base + i < p.logCap)
            p.log[s_base + i] = s_stage[i];
          else
            ctl->overflow = 1;
        }
        __syncthreads();
        if (tid == 0)
          s_staged = 0u;
      }
      grid.sync();
      wl_stamp(ctl, gtid);
      if (gtid == 0)
        ctl->nWalk = 0u;  // (every thread has read it: the barrier above lies in between)
      if (*reinterpret_cast<volatile int *>(&ctl->overflowP1) || *reinterpret_cast<volatile int *>(&ctl->overflow)) {
        status = kWlOverflow;
        break;
      }
    }
    const unsigned nLog = min(*reinterpret_cast<volatile unsigned *>(&ctl->nLog), p.logCap);
    if (gtid == 0) {  // (the consumed item lists; nobody touches these counters before P6)
      ctl->nS[cur] = 0u;
      ctl->nL[cur] = 0u;
    }
    if (nLog > 0 || *reinterpret_cast<volatile unsigned *>(&ctl->nK) > 0) {
      // P2: cells that candidates touch for the first time
      for (unsigned i = gtid; i < nLog; i += gsize) {
        const BtLogEntry e = p.log[i];
        const int j = static_cast<int>(e.var & 0x7fffffffu);
        unsigned *word = const_cast<unsigned *>(&p.dmask[bt_word(p, j, e.b >> 5)]);
        const unsigned bit = 1u << (e.b & 31);
        if (__ldcg(word) & bit)
          continue;
        if (atomicOr(word, bit) & bit)
          continue;  // somebody else got there first and writes the cell
        const Bnd x = p.baseX[j];
        const size_t at = bt_cell(p, j, e.b);
        p.X[0][at] = x;
        p.X[1][at] = x;
        const unsigned ci = wl_append(&ctl->nCells, w.cellCap, &ctl->overflow);
        if (ci != ~0u) {
          WlCell c;
          c.b = e.b;
          c.j = j;
          w.cells[ci] = c;
        }
      }
      grid.sync();
      wl_stamp(ctl, gtid);
      // P3: the candidates, applied as the direct sink would have; whoever marks a cell first lists it
      for (unsigned i = gtid; i < nLog; i += gsize) {
        const BtLogEntry e = p.log[i];
        const int j = static_cast<int>(e.var & 0x7fffffffu);
        Bnd *cell = &p.X[nxt][bt_cell(p, j, e.b)];
        if (e.var & 0x80000000u)
          atomic_min_f64(&cell->ub, e.v);
        else
          atomic_max_f64(&cell->lb, e.v);
        const unsigned bit = 1u << (e.b & 31);
        const unsigned old = atomicOr(&p.dirty[cur][bt_word(p, j, e.b >> 5)], bit);
        if (!(old & bit)) {
          const unsigned ki = wl_append(&ctl->nK, w.kCap, &ctl->overflow);
          if (ki != ~0u) {
            WlCell c;
            c.b = e.b;
            c.j = j;
            w.K[ki] = c;
          }
        }
      }
      grid.sync();
      wl_stamp(ctl, gtid);
      // P4: R9 per changed cell, counts per vector
      const unsigned nK4 = min(*reinterpret_cast<volatile unsigned *>(&ctl->nK), w.kCap);
      for (unsigned i = gtid; i < nK4; i += gsize) {
        const WlCell c = w.K[i];
        const size_t at = bt_cell(p, c.j, c.b);
        Bnd x = p.X[nxt][at];
        const double L0 = x.lb, U0 = x.ub;
        if (settle(x.lb, x.ub, p.X[cur][at].ub, p.s))
          atomicMin(&p.vecInfCol[c.b], c.j);
        else if (x.lb != L0 || x.ub != U0)
          p.X[nxt][at] = x;
        atomicAdd(&w.vecChanged2[cur][c.b], 1);
      }
      if (gtid == 0) {
        ctl->nLog = 0u;
        unsigned long long t;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
        if (t - ctl->t0 >= w.budgetNs)
          ctl->timeUp = 1;
      }
      grid.sync();
      wl_stamp(ctl, gtid);
    }
    // P5: per-vector bookkeeping of src/CbcBoundPropagation.cpp:129-189; P6: carry + next round's items
    const int timeUp = *reinterpret_cast<volatile int *>(&ctl->timeUp);
    for (unsigned b = gtid; b < static_cast<unsigned>(p.Bp); b += gsize) {
      bool on = (w.active2[cur][b >> 5] >> (b & 31)) & 1u;
      int stoppedRounds = 0;
      if (on) {
        const int chg = w.vecChanged2[cur][b];
        p.vecRounds[b] = round + 1;
        int stop = 0, fin = cur;
        if (p.vecInfRow[b] != 0x7fffffff || p.vecInfCol[b] != 0x7fffffff)
          stop = 4;
        else if (chg == 0)
          stop = 1;
        else {
          p.vecTotal[b] += chg;
          if (round + 1 >= w.limit) {
            stop = 2;
            fin = nxt;
          } else if (timeUp) {
            stop = 3;
            fin = nxt;
          }
        }
        if (stop) {
          p.vecStop[b] = stop;
          p.vecFinal[b] = fin;
          on = false;
          stoppedRounds = round + 1;
        }
      }
      stoppedRounds = __reduce_add_sync(0xffffffffu, stoppedRounds);
      if (lane == 0 && stoppedRounds)
        atomicAdd(&ctl->roundsTotal, static_cast<unsigned long long>(stoppedRounds));
      w.vecChanged2[nxt][b] = 0;
      const unsigned m = __ballot_sync(0xffffffffu, on);  // (Bp and the stride are multiples of 32: whole warps)
      if (lane == 0) {
        w.active2[nxt][b >> 5] = m;
        if (m)
          ctl->anyActive[nxt] = 1;
      }
    }
    const unsigned nK6 = min(*reinterpret_cast<volatile unsigned *>(&ctl->nK), w.kCap);
    const unsigned cpw6 = wl_cells_per_warp(nK6, nwarps);
    for (unsigned i0 = gwarp * cpw6; i0 < nK6; i0 += nwarps * cpw6) {  // (whole warps: the listing below is a warp's job)
      const unsigned i = lane < cpw6 ? i0 + lane : nK6;
      bool go = false;
      WlCell c;
      c.b = 0;
      c.j = 0;
      Bnd was, now;
      was.lb = was.ub = now.lb = now.ub = 0.0;
      if (i < nK6) {
        c = w.K[i];
        const size_t word = bt_word(p, c.j, c.b >> 5);
        const unsigned bit = 1u << (c.b & 31);
        atomicAnd(&p.dirty[cur][word], ~bit);
        const bool applied = p.vecInfRow[c.b] == 0x7fffffff && p.vecInfCol[c.b] == 0x7fffffff;
        if (applied) {
          if (!(atomicOr(&p.ever[word], bit) & bit))
            atomicAdd(&w.updCnt[c.b], 1);  // (the cell's first change: one more entry of the vector's update list)
          if (!(round + 1 >= w.limit || timeUp)) {
            const size_t at = bt_cell(p, c.j, c.b);
            was = p.X[cur][at];
            now = p.X[nxt][at];
            p.X[cur][at] = now;
            go = true;
          }
        }
      }
      wl_emit_rows(w, go, c.b, c.j, nxt, round + 1, was, now, lane, s_emit[tid >> 5]);
    }
    grid.sync();
    wl_stamp(ctl, gtid);
    if (gtid == 0)
      ctl->nK = 0u;
    if (!*reinterpret_cast<volatile int *>(&ctl->anyActive[nxt])) {
      ++round;
      break;
    }
    // (nK is read again only after the next P3 has appended to it: two barriers from here)
  }
  // pairs walked: one atomic per warp
#pragma unroll
  for (int m = 16; m >= 1; m >>= 1)
    myPairs += __shfl_xor_sync(0xffffffffu, myPairs, m);
  if (lane == 0 && myPairs)
    atomicAdd(&ctl->pairs, myPairs);
  // (every way out of the loop passes a grid barrier after the last write to these flags)
  if (status == kWlOk
    && (*reinterpret_cast<volatile int *>(&ctl->overflow) || *reinterpret_cast<volatile int *>(&ctl->overflowP1)))
    status = kWlOverflow;  // a list filled up in the last round
  if (gtid == 0) {
    ctl->status = status;
    ctl->rounds = round;
  }
  if (status != kWlOk || w.updCap <= 0)
    return;
  // ---- P7: the vectors' update lists, so that fetching them is a copy: offsets, fill, order
  // (the counts: P6 adds one when it sets a cell's `ever` bit; a vector proved infeasible lists nothing)
  const unsigned nAll = min(*reinterpret_cast<volatile unsigned *>(&ctl->nCells), w.cellCap);
  if (blockIdx.x == 0) {  // offsets: one CTA, 8,192 vectors at a time: a warp takes 1,024 consecutive ones, 32 per lane
    __shared__ long long s_tot[kWarps];  // in registers (all loads in flight at once), and scans them by shuffles
    const int wq = tid >> 5;
    long long done = 0;  // entries of the vectors before this piece
    for (int base = 0; base < w.B; base += kWarps * 1024) {
      const int b0 = base + wq * 1024;
      int v[32];
      int sum = 0;
#pragma unroll
      for (int u = 0; u < 32; ++u) {
        const int b = b0 + 32 * u + lane;
        v[u] = (b < w.B && p.vecStop[b] != 4) ? w.updCnt[b] : 0;
      }
#pragma unroll
      for (int u = 0; u < 32; ++u)
        sum += v[u];
#pragma unroll
      for (int m = 16; m >= 1; m >>= 1)
        sum += __shfl_xor_sync(0xffffffffu, sum, m);
      __syncthreads();
      if (lane == 0)
        s_tot[wq] = sum;
      __syncthreads();
      long long run = done;
      long long all = done;
      for (int q = 0; q < kWarps; ++q) {
        if (q < wq)
          run += s_tot[q];
        all += s_tot[q];
      }
#pragma unroll
      for (int u = 0; u < 32; ++u) {
        int incl = v[u];
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const int t = __shfl_up_sync(0xffffffffu, incl, d);
          if (lane >= d)
            incl += t;
        }
        const int b = b0 + 32 * u + lane;
        if (b < w.B)
          w.updStart[b] = run + (incl - v[u]);
        run += __shfl_sync(0xffffffffu, incl, 31);
      }
      done = all;
    }
    if (tid == 0) {
      w.updStart[w.B] = done;
      ctl->nUpd = static_cast<unsigned long long>(done);
    }
  }
  grid.sync();
  wl_stamp(ctl, gtid);
  const long long nUpd = static_cast<long long>(*reinterpret_cast<volatile unsigned long long *>(&ctl->nUpd));
  if (nUpd > w.updCap)
    return;  // (the host grows the buffers and takes the lists with the stand-alone kernels)
  for (unsigned i = gtid; i < nAll; i += gsize)
    wl_upd_fill(w.cells[i], p, w.updStart, w.updCnt + p.Bp, w.updColTmp, w.updLbTmp, w.updUbTmp);
  grid.sync();
  wl_stamp(ctl, gtid);
  // (what the fill wrote is read with plain loads: the barrier's fence has invalidated L1, and nobody writes it now)
  for (unsigned b = gwarp; b < static_cast<unsigned>(w.B); b += nwarps)
    wl_upd_sort_warp(w.updStart, b, lane, reinterpret_cast<int *>(s_emit[tid >> 5]), w.updColTmp, w.updLbTmp, w.updUbTmp,
      w.updCol, w.updLb, w.updUb, w.updHead, w.updHeadCap);
  {  // the long segments: a CTA each; every CTA looks through its share of the vectors at once
    const int perCta = (w.B + gridDim.x - 1) / gridDim.x;
    const int c0 = min(static_cast<int>(blockIdx.x) * perCta, w.B), c1 = min(c0 + perCta, w.B);
    for (int base = c0; base < c1; base += kThreads) {
      const int b = base + tid;
      const bool big = b < c1 && w.updStart[b + 1] - w.updStart[b] > kWlSortWarp;
      __syncthreads();
      if (tid == 0)
        s_base = 0u;
      __syncthreads();
      if (big)
        reinterpret_cast<int *>(s_emit)[atomicAdd(&s_base, 1u)] = b;  // (at most kThreads ints; the warps are done with s_emit)
      __syncthreads();
      const unsigned nBig = s_base;
      for (unsigned q = 0; q < nBig; ++q)
        wl_upd_sort_cta(p, w.updStart, reinterpret_cast<int *>(s_emit)[q], tid, reinterpret_cast<unsigned long long *>(s_stage),
          w.updColTmp, w.updLbTmp, w.updUbTmp, w.updCol, w.updLb, w.updUb, w.updHead, w.updHeadCap);
    }
  }
  if (gtid == 0)
    ctl->updDone = 1;
}

// ---- before and after the run ----------------------------------------------------------------------------------------
// Everything the run starts from, in one launch: the control block and the per-vector words zeroed, vectors 0..B-1
// running, the node's bounds packed as (lb, ub) pairs.
__global__ void wl_setup_kernel(const WlParams w, const double *__restrict__ baseLb, const double *__restrict__ baseUb,
  Bnd *__restrict__ baseX)
{
  const BtParams &p = w.bt;
  const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < sizeof(WlCtl) / sizeof(unsigned))
    reinterpret_cast<unsigned *>(w.ctl)[i] = 0u;
  if (i < static_cast<unsigned>(p.n)) {
    Bnd x;
    x.lb = baseLb[i] + 0.0;  // (no -0.0 in the bound arrays)
    x.ub = baseUb[i] + 0.0;
    baseX[i] = x;
  }
  if (i < static_cast<unsigned>(p.Bp)) {  // (Bp is a multiple of 32: whole warps)
    p.vecInfRow[i] = 0x7fffffff;
    p.vecInfCol[i] = 0x7fffffff;
    p.vecChanged[i] = 0;
    p.vecRounds[i] = 0;
    p.vecTotal[i] = 0;
    p.vecStop[i] = 0;
    p.vecFinal[i] = 0;
    w.vecChanged2[0][i] = 0;
    w.vecChanged2[1][i] = 0;
    w.updCnt[i] = 0;
    w.updCnt[p.Bp + i] = 0;
    const unsigned m = __ballot_sync(0xffffffffu, static_cast<int>(i) < w.B);
    if ((threadIdx.x & 31) == 0) {
      p.active[i >> 5] = m;
      w.active2[0][i >> 5] = m;
      w.active2[1][i >> 5] = 0u;
    }
  }
}

// ---- updatedBounds() of a worklist run, stand-alone (the run's own P7 did not have the room) ------------------------
__global__ void wl_upd_count_kernel(const WlCell *__restrict__ cells, unsigned nCells, const BtParams p, int *cnt)
{
  const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nCells)
    wl_upd_count(cells[i], p, cnt);
}

__global__ void wl_upd_fill_kernel(const WlCell *__restrict__ cells, unsigned nCells, const BtParams p,
  const long long *__restrict__ start, int *cursor, int *col, double *lb, double *ub)
{
  const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nCells)
    wl_upd_fill(cells[i], p, start, cursor, col, lb, ub);
}

// one CTA of kThreads threads per vector
__global__ void __launch_bounds__(kThreads) wl_upd_sort_kernel(const BtParams p, const long long *__restrict__ start, int B,
  const int *colIn, const double *lbIn, const double *ubIn, int *colOut, double *lbOut, double *ubOut)
{
  __shared__ unsigned long long keys[kWlSortSmem];
  const int b = blockIdx.x;
  if (b >= B)
    return;
  __shared__ int cols[kWlSortWarp];
  if ((threadIdx.x >> 5) == 0)
    wl_upd_sort_warp(start, b, threadIdx.x & 31, cols, colIn, lbIn, ubIn, colOut, lbOut, ubOut, nullptr, 0);
  wl_upd_sort_cta(p, start, b, threadIdx.x, keys, colIn, lbIn, ubIn, colOut, lbOut, ubOut, nullptr, 0);
}

}  // namespace cbp
