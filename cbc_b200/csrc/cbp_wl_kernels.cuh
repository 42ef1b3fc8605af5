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

// control block (device); zeroed by the host before every launch
struct WlCtl {
  unsigned nS[2], nL[2];  // items of the running / next round: rows <= kShortLen nonzeros, longer rows
  unsigned nLog, nK, nCells;
  int baseBad, timeUp;
  int overflow;    // a list other than the candidate list ran out of room (written outside P1 only)
  int overflowP1;  // the candidate list did (written in P1 only): two words, so that each is read while nobody writes it
  int anyActive[2];
  int status, rounds;
  unsigned long long pairs;   // (nonzero, vector) pairs walked
  unsigned long long itemsTotal;
  unsigned long long t0;
};

struct WlParams {
  BtParams bt;
  const int *cscStart, *cscRow;
  // the loaded batch
  const long long *dStart;
  const unsigned *dVar;
  const double *dBnd;
  int B;
  // lists
  WlItem *itemsS[2], *itemsL[2];
  WlCell *K, *cells;
  unsigned itemCap, kCap, cellCap;
  unsigned prevCells;  // cells the previous worklist run left behind (their mask / ever bits are taken out first)
  WlCtl *ctl;
  int *vecChanged2[2];   // [Bp] changed columns, by round parity
  unsigned *active2[2];  // [G] running vectors, by round parity
  int limit;             // round limit (INT_MAX: none)
  unsigned long long budgetNs;
};

namespace cg = cooperative_groups;

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
struct WlLogSink {
  BtLogEntry *log;
  unsigned *count;
  unsigned cap;
  int *overflow;
  int b;
  __device__ __forceinline__ void put(unsigned var, double v)
  {
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

// reads "the node's bounds, nothing else" (base check)
struct WlBaseReader {
  const Bnd *baseX;
  __device__ __forceinline__ Bnd operator()(int j) const
  {
    const double2 v = __ldg(reinterpret_cast<const double2 *>(&baseX[j]));
    Bnd x;
    x.lb = v.x;
    x.ub = v.y;
    return x;
  }
};
struct WlVecReader {
  const BtParams *p;
  const Bnd *srcCells;
  const unsigned *dm;
  int vl;
  __device__ __forceinline__ Bnd operator()(int j) const { return wl_bounds(*p, srcCells, dm, vl, j); }
};

// One row of at most kShortLen nonzeros by one thread, CSR order (the order of the CPU restatement).
// Returns true if the row is violated.
template <class Reader, class Sink>
__device__ __forceinline__ bool wl_row_thread(const BtParams &p, int row, const Reader &rd, Sink &sink)
{
  const int kb = __ldg(&p.rowNz[row]), cnt = __ldg(&p.rowCnt[row]);
  RowAct ra = { 0.0, 0.0, 0.0, 0, 0 };
  for (int k = 0; k < cnt; k += 4) {
    double a[4];
    int cj[4];
    Bnd x[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      a[u] = 0.0;
      cj[u] = 0;
      if (k + u < cnt) {
        cj[u] = __ldg(&p.col[kb + k + u]);
        a[u] = __ldg(&p.val[kb + k + u]);
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u)
      x[u] = rd(col_of(cj[u]));
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (a[u] != 0.0)
        accumulate_fast(ra, a[u], col_is_int(cj[u]), x[u], p.s.inf);
  }
  const RowPlan pl = plan_row(ra, __ldg(&p.rowLB[row]), __ldg(&p.rowUB[row]), p.s);
  if (pl.infeasible)
    return true;
  if (pl.sideU || pl.sideL)
    for (int k = 0; k < cnt; ++k) {
      const int c = __ldg(&p.col[kb + k]);
      const double a = __ldg(&p.val[kb + k]);
      const Bnd x = rd(col_of(c));
      tighten_nz(sink, pl, a, col_of(c), col_is_int(c), a > 0.0 ? x.lb : x.ub, a > 0.0 ? x.ub : x.lb, p.s);
    }
  return false;
}

// One longer row by one warp: lanes stride over the row, butterfly.  Every lane returns the verdict.
template <class Reader, class Sink>
__device__ __forceinline__ bool wl_row_warp(const BtParams &p, int row, const Reader &rd, Sink &sink, int lane)
{
  const int kb = __ldg(&p.rowNz[row]), ke = kb + __ldg(&p.rowCnt[row]);
  RowAct ra = { 0.0, 0.0, 0.0, 0, 0 };
  for (int k = kb + lane; k < ke; k += 32) {
    const int c = __ldg(&p.col[k]);
    const double a = __ldg(&p.val[k]);
    const Bnd x = rd(col_of(c));
    accumulate(ra, a, col_is_int(c), a > 0.0 ? x.lb : x.ub, a > 0.0 ? x.ub : x.lb, p.s.inf);
  }
  warp_reduce(ra);
  const RowPlan pl = plan_row(ra, __ldg(&p.rowLB[row]), __ldg(&p.rowUB[row]), p.s);
  if (pl.infeasible)
    return true;
  if (pl.sideU || pl.sideL)
    for (int k = kb + lane; k < ke; k += 32) {
      const int c = __ldg(&p.col[k]);
      const double a = __ldg(&p.val[k]);
      const Bnd x = rd(col_of(c));
      tighten_nz(sink, pl, a, col_of(c), col_is_int(c), a > 0.0 ? x.lb : x.ub, a > 0.0 ? x.ub : x.lb, p.s);
    }
  return false;
}

// The rows of column j become items of vector b in the lists of parity `nxt`.  One reservation per list and
// thread (a column's rows are few), so the counters see one atomic per thread, not one per item.
__device__ __forceinline__ void wl_emit_rows(const WlParams &w, int b, int j, int nxt)
{
  const int cs = __ldg(&w.cscStart[j]), ce = __ldg(&w.cscStart[j + 1]);
  int nShort = 0;
  for (int q = cs; q < ce; ++q)
    nShort += __ldg(&w.bt.rowCnt[__ldg(&w.cscRow[q])]) <= kShortLen ? 1 : 0;
  const int nLong = (ce - cs) - nShort;
  unsigned atS = nShort ? atomicAdd(&w.ctl->nS[nxt], static_cast<unsigned>(nShort)) : 0u;
  unsigned atL = nLong ? atomicAdd(&w.ctl->nL[nxt], static_cast<unsigned>(nLong)) : 0u;
  if ((nShort && atS + nShort > w.itemCap) || (nLong && atL + nLong > w.itemCap)) {
    w.ctl->overflow = 1;
    return;
  }
  for (int q = cs; q < ce; ++q) {
    const int row = __ldg(&w.cscRow[q]);
    WlItem it;
    it.b = b;
    it.row = row;
    if (__ldg(&w.bt.rowCnt[row]) <= kShortLen)
      w.itemsS[nxt][atS++] = it;
    else
      w.itemsL[nxt][atL++] = it;
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
#define CBP_WL_BLOCKS 3
#endif

__global__ void __launch_bounds__(kThreads, CBP_WL_BLOCKS) bt_worklist_kernel(const WlParams w)
{
  cg::grid_group grid = cg::this_grid();
  const BtParams &p = w.bt;
  WlCtl *ctl = w.ctl;
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
  if (gtid == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    ctl->t0 = t;
  }
  {
    const WlBaseReader rd = { p.baseX };
    WlFlagSink fs = { &ctl->baseBad };
    for (unsigned j = gtid; j < static_cast<unsigned>(p.n); j += gsize) {  // R11 on the node's own bounds
      const Bnd x = rd(j);
      double l = x.lb, u = x.ub;
      if (prepass_col(l, u, p.colType[j] != 0, p.s) != 0)
        ctl->baseBad = 1;
    }
    for (unsigned r = gtid; r < static_cast<unsigned>(p.mKept); r += gsize)
      if (__ldg(&p.rowCnt[r]) <= kShortLen && wl_row_thread(p, r, rd, fs))
        ctl->baseBad = 1;
    for (unsigned r = gwarp; r < static_cast<unsigned>(p.mKept); r += nwarps)
      if (__ldg(&p.rowCnt[r]) > kShortLen && wl_row_warp(p, r, rd, fs, lane))
        ctl->baseBad = 1;
  }
  grid.sync();
  if (*reinterpret_cast<volatile int *>(&ctl->baseBad)) {
    if (gtid == 0)
      ctl->status = kWlBaseNotFixpoint;
    return;
  }

  // ---- phase 0b: the diffs.  One thread per vector, diffs applied in order; the cells they touch are created
  // (mask bit, both buffers), get R11, and their columns' rows are the items of the first round.
  for (unsigned b = gtid; b < static_cast<unsigned>(w.B); b += gsize) {
    const long long k0 = w.dStart[b], k1 = w.dStart[b + 1];
    unsigned *dmaskW = const_cast<unsigned *>(p.dmask);
    const unsigned bit = 1u << (b & 31);
    for (long long k = k0; k < k1; ++k) {
      const unsigned v = w.dVar[k];
      const int j = diff_col(v);
      if (j >= p.n)
        continue;
      const size_t at = bt_cell(p, j, b);
      unsigned *word = &dmaskW[bt_word(p, j, b >> 5)];
      const bool had = (*reinterpret_cast<volatile unsigned *>(word) & bit) != 0u;  // (own bit: set by this thread only)
      Bnd x = had ? p.X[0][at] : p.baseX[j];
      if (!had) {
        atomicOr(word, bit);
        const unsigned ci = atomicAdd(&ctl->nCells, 1u);
        if (ci < w.cellCap) {
          WlCell c;
          c.b = b;
          c.j = j;
          w.cells[ci] = c;
        } else {
          ctl->overflow = 1;
        }
      }
      if (v & kDiffUpper)
        x.ub = w.dBnd[k] + 0.0;
      else
        x.lb = w.dBnd[k] + 0.0;
      p.X[0][at] = x;
      p.X[1][at] = x;
    }
    bool dead = false;
    for (long long k = k0; k < k1; ++k) {
      const int j = diff_col(w.dVar[k]);
      if (j >= p.n)
        continue;
      const size_t at = bt_cell(p, j, b);
      Bnd x = p.X[0][at];
      const int r = prepass_col(x.lb, x.ub, p.colType[j] != 0, p.s);
      if (r == 2) {
        atomicMin(&p.vecInfCol[b], j);
        dead = true;
      } else if (r == 1) {
        p.X[0][at] = x;
        p.X[1][at] = x;
        const unsigned old = atomicOr(&p.dirty[0][bt_word(p, j, b >> 5)], bit);
        if (!(old & bit)) {
          const unsigned ki = atomicAdd(&ctl->nK, 1u);
          if (ki < w.kCap) {
            WlCell c;
            c.b = b;
            c.j = j;
            w.K[ki] = c;
          } else {
            ctl->overflow = 1;
          }
        }
      }
    }
    if (dead) {  // the input already crosses (R11): stops before its first round
      p.vecStop[b] = 4;
      p.vecRounds[b] = 1;
      p.vecFinal[b] = 0;
      atomicAnd(&w.active2[0][b >> 5], ~bit);
    } else {
      for (long long k = k0; k < k1; ++k) {
        const int j = diff_col(w.dVar[k]);
        if (j < p.n)
          wl_emit_rows(w, b, j, 0);
      }
    }
  }
  grid.sync();

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
    // P1: the items
    for (unsigned i = gtid; i < nS; i += gsize) {
      const WlItem it = w.itemsS[cur][i];
      const WlVecReader rd = { &p, srcX + bt_base32(p, it.b), p.dmask + bt_word(p, 0, it.b >> 5), it.b & 31 };
      WlLogSink sink = { p.log, &ctl->nLog, p.logCap, &ctl->overflowP1, it.b };
      if (wl_row_thread(p, it.row, rd, sink))
        atomicMin(&p.vecInfRow[it.b], __ldg(&p.rowOrig[it.row]));
      myPairs += __ldg(&p.rowCnt[it.row]);
    }
    for (unsigned i = gwarp; i < nL; i += nwarps) {
      const WlItem it = w.itemsL[cur][i];
      const WlVecReader rd = { &p, srcX + bt_base32(p, it.b), p.dmask + bt_word(p, 0, it.b >> 5), it.b & 31 };
      WlLogSink sink = { p.log, &ctl->nLog, p.logCap, &ctl->overflowP1, it.b };
      if (wl_row_warp(p, it.row, rd, sink, lane) && lane == 0)
        atomicMin(&p.vecInfRow[it.b], __ldg(&p.rowOrig[it.row]));
      if (lane == 0)
        myPairs += __ldg(&p.rowCnt[it.row]);
    }
    if (gtid == 0) {
      ctl->itemsTotal += nS + nL;
      ctl->anyActive[nxt] = 0;
    }
    grid.sync();
    if (*reinterpret_cast<volatile int *>(&ctl->overflowP1)) {  // (P1 is over: nobody writes this word now)
      status = kWlOverflow;
      break;
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
    }
    // P5: per-vector bookkeeping of src/CbcBoundPropagation.cpp:129-189; P6: carry + next round's items
    const int timeUp = *reinterpret_cast<volatile int *>(&ctl->timeUp);
    for (unsigned b = gtid; b < static_cast<unsigned>(p.Bp); b += gsize) {
      bool on = (w.active2[cur][b >> 5] >> (b & 31)) & 1u;
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
        }
      }
      w.vecChanged2[nxt][b] = 0;
      const unsigned m = __ballot_sync(0xffffffffu, on);  // (Bp is a multiple of 32 and the stride of 32: whole warps)
      if (lane == 0) {
        w.active2[nxt][b >> 5] = m;
        if (m)
          ctl->anyActive[nxt] = 1;
      }
    }
    const unsigned nK6 = min(*reinterpret_cast<volatile unsigned *>(&ctl->nK), w.kCap);
    for (unsigned i = gtid; i < nK6; i += gsize) {
      const WlCell c = w.K[i];
      const size_t word = bt_word(p, c.j, c.b >> 5);
      const unsigned bit = 1u << (c.b & 31);
      atomicAnd(&p.dirty[cur][word], ~bit);
      const bool applied = p.vecInfRow[c.b] == 0x7fffffff && p.vecInfCol[c.b] == 0x7fffffff;
      if (!applied)
        continue;
      atomicOr(&p.ever[word], bit);
      if (round + 1 >= w.limit || timeUp)
        continue;
      const size_t at = bt_cell(p, c.j, c.b);
      p.X[cur][at] = p.X[nxt][at];
      wl_emit_rows(w, c.b, c.j, nxt);
    }
    grid.sync();
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
  if (gtid == 0) {
    ctl->status = status;
    ctl->rounds = round;
  }
}

}  // namespace cbp
