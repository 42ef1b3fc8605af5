// cbp_device.cuh — device-side building blocks shared by the single-vector
// and the batched kernels (sm_100a).  Compiled with -fmad=false: every
// multiply-add that the rules want fused is written as fma(); nothing else
// may contract, so the arithmetic is operation-for-operation the one frozen
// in DESIGN.md §3 (rules R1-R12).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace cbp {

constexpr int kThreads = 256;      // threads per CTA of the streaming kernels
constexpr int kWarps = kThreads / 32;
constexpr int kTileNnz = 1024;     // nonzeros staged per tile (12 KB of val+col per stage)
constexpr int kTileRows = 256;     // rows per tile (one per thread)
constexpr int kShortLen = 32;      // rows up to this length: one thread per row (sliced-ELL)
#ifndef CBP_SV_CHUNK
#define CBP_SV_CHUNK 8
#endif
#ifndef CBP_SV_NT_BLOCKS
#define CBP_SV_NT_BLOCKS 3
#endif
constexpr int kSvChunk = CBP_SV_CHUNK;        // sliced-ELL steps whose loads are in flight together
#ifndef CBP_SV_CHUNK_SHARDED
#define CBP_SV_CHUNK_SHARDED 8
#endif
constexpr int kSvChunkSharded = CBP_SV_CHUNK_SHARDED;  // the same for the row-sharded kernel (which has fewer registers to spare)
constexpr int kSvNoTileBlocks = CBP_SV_NT_BLOCKS;  // CTAs per SM of the short-rows-only variant
#ifndef CBP_SV_SHARDED_BLOCKS
#define CBP_SV_SHARDED_BLOCKS 3
#endif
constexpr int kSvShardedBlocks = CBP_SV_SHARDED_BLOCKS;  // CTAs per SM of the row-sharded variants
#ifndef CBP_SV_P2ROWS
#define CBP_SV_P2ROWS 2
#endif
constexpr int kSvPass2Rows = CBP_SV_P2ROWS;  // rows whose second-pass loads are in flight together
constexpr int kSellWindow = 4096;  // short rows are sorted by length inside windows of this many rows
constexpr int kAlign = 16;         // tile starts are multiples of 16 entries (128 B of val)
constexpr int kMaxPeers = 8;       // GPUs of one NVSwitch box in a row-sharded run
constexpr int kStampRounds = 64;   // rounds whose phase timestamps are kept (diagnostic)
constexpr int kRoundCap = 1024;    // per-round counts kept on the device
constexpr int kBtLongLen = 256;    // batch mode: longer rows get a whole CTA per (row, group)

constexpr double kContRelThr = 1.0e-3;  // R8
constexpr double kIntThr = 0.5;         // R8

enum TileKind { kShort = 0, kMedium = 1, kLong = 2 };

// One unit of streaming work: a run of whole rows (kShort / kMedium) or one
// chunk of a long row (kLong).  nzBegin indexes the packed val/col arrays.
struct Tile {
  int nzBegin;   // multiple of kAlign
  int nzCount;   // real entries
  int rowBegin;  // first row (kept order)
  int rowCount;
  int rpOff;     // offset of this tile's rowCount+1 local row pointers
  int kind;
  int longIdx;   // kLong: index into LongRow[]
  int chunkIdx;  // kLong: chunk number inside the row
};

struct LongRow {
  int row;        // kept-order row index
  int firstPart;  // index of chunk 0's partial
  int nChunks;
  int pad;
};

// Partial (or whole) row activity: R3 plus the largest |a|*(ub-lb).
struct __align__(16) RowAct {
  double minAct, maxAct, maxRange;
  int infMin, infMax;
};
static_assert(sizeof(RowAct) == 32, "RowAct is read back as two 16-byte words");

struct __align__(16) Bnd {  // one column's bounds
  double lb, ub;
};

// ---- scalars every kernel needs ----------------------------------------
struct Scal {
  double inf;      // solver infinity
  double tol;      // primal tolerance
  double feasTol;  // 100 * tol (R4, R9)
};

__device__ __forceinline__ bool is_inf(double x, double INF) { return !(fabs(x) < INF); }

// The packed column index carries the column's integrality in bit 31 (a static
// property, set at upload), so the tightening pass needs no colType lookup.
constexpr int kColMask = 0x7fffffff;
__device__ __forceinline__ int col_of(int packed) { return packed & kColMask; }
__device__ __forceinline__ bool col_is_int(int packed) { return packed < 0; }

// Node bound diffs arrive in CbcPartialNodeInfo's encoding (/root/reference/src/CbcPartialNodeInfo.cpp:139,196,
// 208,220): bit 31 = upper bound, bit 30 = "can go odd way" (a flag applyBounds() sets, not part of the column),
// bits 0-29 = column.
constexpr unsigned kDiffUpper = 0x80000000u;
constexpr unsigned kDiffColMask = 0x3fffffffu;
__host__ __device__ __forceinline__ int diff_col(unsigned v) { return static_cast<int>(v & kDiffColMask); }

// atomic max / min on doubles through the integer atomics (no -0.0 is ever
// stored: R7 adds +0.0).  For v >= 0 signed-int order equals double order and
// any negative double is a negative int; for v < 0 unsigned order is reversed
// double order among negatives and every non-negative double is smaller.
// Reductions written as PTX `red`: nothing comes back, the thread does not wait.  (Left to the
// compiler, an atomicMax whose result is unused becomes RED in some kernels and a returning ATOM in
// others; the latter holds the issuing warp's scoreboard until L2 answers.)
__device__ __forceinline__ void red_max_s64(void *addr, long long v)
{
  asm volatile("red.relaxed.gpu.global.max.s64 [%0], %1;" ::"l"(addr), "l"(v) : "memory");
}
__device__ __forceinline__ void red_min_s64(void *addr, long long v)
{
  asm volatile("red.relaxed.gpu.global.min.s64 [%0], %1;" ::"l"(addr), "l"(v) : "memory");
}
__device__ __forceinline__ void red_max_u64(void *addr, unsigned long long v)
{
  asm volatile("red.relaxed.gpu.global.max.u64 [%0], %1;" ::"l"(addr), "l"(v) : "memory");
}
__device__ __forceinline__ void red_min_u64(void *addr, unsigned long long v)
{
  asm volatile("red.relaxed.gpu.global.min.u64 [%0], %1;" ::"l"(addr), "l"(v) : "memory");
}
__device__ __forceinline__ void red_min_s32(int *addr, int v)
{
  asm volatile("red.relaxed.gpu.global.min.s32 [%0], %1;" ::"l"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ void red_or_b32(unsigned *addr, unsigned v)
{
  asm volatile("red.relaxed.gpu.global.or.b32 [%0], %1;" ::"l"(addr), "r"(v) : "memory");
}

// max / min on doubles through the ordering of their bit patterns: as signed integers for values >= 0,
// reversed as unsigned integers for negative ones
__device__ __forceinline__ void atomic_max_f64(double *addr, double v)
{
  if (v >= 0.0)
    red_max_s64(addr, __double_as_longlong(v));
  else
    red_min_u64(addr, static_cast<unsigned long long>(__double_as_longlong(v)));
}
__device__ __forceinline__ void atomic_min_f64(double *addr, double v)
{
  if (v >= 0.0)
    red_min_s64(addr, __double_as_longlong(v));
  else
    red_max_u64(addr, static_cast<unsigned long long>(__double_as_longlong(v)));
}

__device__ __forceinline__ Bnd ldcg_bnd(const Bnd *p)
{
  const double2 v = __ldcg(reinterpret_cast<const double2 *>(p));
  Bnd b;
  b.lb = v.x;
  b.ub = v.y;
  return b;
}

// ---- R4 / R5: what a finished row activity allows ------------------------
struct RowPlan {
  double slackU, slackL;
  int infMin, infMax;
  bool sideU, sideL, infeasible;
};

__device__ __forceinline__ RowPlan plan_row(const RowAct &a, double rl, double ru, const Scal &s)
{
  RowPlan p;
  const bool ruFin = ru < s.inf, rlFin = rl > -s.inf;
  p.infeasible = (ruFin && a.infMin == 0 && a.minAct > ru + s.feasTol)
    || (rlFin && a.infMax == 0 && a.maxAct < rl - s.feasTol);
  p.sideU = ruFin && a.infMin <= 1 && !(a.infMax == 0 && a.maxAct <= ru + s.tol);
  p.sideL = rlFin && a.infMax <= 1 && !(a.infMin == 0 && a.minAct >= rl - s.tol);
  p.slackU = ru - a.minAct;
  p.slackL = a.maxAct - rl;
  p.infMin = a.infMin;
  p.infMax = a.infMax;
  // Row-level shortcut, exact: with only finite ranges a candidate passes the
  // per-nonzero pre-check of R6 only if slack < |a| * useful_range <= maxRange.
  if (a.infMin + a.infMax == 0) {
    p.sideU = p.sideU && p.slackU < a.maxRange;
    p.sideL = p.sideL && p.slackL < a.maxRange;
  }
  if (p.infeasible) {
    p.sideU = false;
    p.sideL = false;
  }
  return p;
}

// R6: the part of a column's range that a candidate must undercut to be
// accepted by R8 (integer: the whole range; continuous: range minus the
// smallest accepted step).
__device__ __forceinline__ double useful_range(double width, bool isInt)
{
  return isInt ? width : width - kContRelThr * fmax(1.0, width);
}

// ---- R3 for one nonzero --------------------------------------------------
// Branch-free: an infinite term enters the sum as a*0 (which leaves the sum
// bit-for-bit unchanged, the sums never hold -0.0) and bumps its counter.
// maxRange collects max |a| * useful_range for the row-level shortcut.
__device__ __forceinline__ void accumulate(RowAct &r, double a, bool isInt, double lo, double hi, double INF)
{
  const bool loInf = is_inf(lo, INF), hiInf = is_inf(hi, INF);
  r.infMin += loInf ? 1 : 0;
  r.infMax += hiInf ? 1 : 0;
  r.minAct = fma(a, loInf ? 0.0 : lo, r.minAct);
  r.maxAct = fma(a, hiInf ? 0.0 : hi, r.maxAct);
  const double w = fabs(a) * useful_range(a > 0.0 ? hi - lo : lo - hi, isInt);
  const double wf = (loInf || hiInf) ? 0.0 : w;
  r.maxRange = wf > r.maxRange ? wf : r.maxRange;
}

// Same result as accumulate(), organised for the common case: both bounds finite (then no
// counters, no zeroing) — the rare infinite bound takes the general path.
__device__ __forceinline__ void accumulate_fast(RowAct &r, double a, bool isInt, const Bnd &x, double INF)
{
  const bool pos = a > 0.0;
  const double lo = pos ? x.lb : x.ub, hi = pos ? x.ub : x.lb;
  if (fabs(x.lb) < INF && fabs(x.ub) < INF) {
    r.minAct = fma(a, lo, r.minAct);
    r.maxAct = fma(a, hi, r.maxAct);
    const double w = fabs(a) * useful_range(x.ub - x.lb, isInt);
    r.maxRange = w > r.maxRange ? w : r.maxRange;
  } else {
    accumulate(r, a, isInt, lo, hi, INF);
  }
}

__device__ __forceinline__ void combine(RowAct &r, const RowAct &o)
{
  r.minAct += o.minAct;
  r.maxAct += o.maxAct;
  r.maxRange = fmax(r.maxRange, o.maxRange);
  r.infMin += o.infMin;
  r.infMax += o.infMax;
}

// ---- R6-R8: candidates of one nonzero ------------------------------------
// Sink must provide  void ub(int j, double v)  and  void lb(int j, double v)
// which publish an ACCEPTED candidate (atomic min / max + dirty mark).
template <class Sink>
__device__ __forceinline__ void offer_ub(Sink &sink, int j, double v, double old, double other, bool isInt,
  const Scal &s)
{
  if (isInt)
    v = floor(v + s.tol);
  v += 0.0;
  if (!(fabs(v) < s.inf))
    return;
  if (fabs(old) < s.inf) {
    const double thr = isInt ? kIntThr : kContRelThr * fmax(1.0, fabs(other) < s.inf ? old - other : fabs(old));
    if (!(v < old - thr))
      return;
  } else if (!(v < old)) {
    return;
  }
  sink.ub(j, v);
}

template <class Sink>
__device__ __forceinline__ void offer_lb(Sink &sink, int j, double v, double old, double other, bool isInt,
  const Scal &s)
{
  if (isInt)
    v = ceil(v - s.tol);
  v += 0.0;
  if (!(fabs(v) < s.inf))
    return;
  if (fabs(old) < s.inf) {
    const double thr = isInt ? kIntThr : kContRelThr * fmax(1.0, fabs(other) < s.inf ? other - old : fabs(old));
    if (!(v > old + thr))
      return;
  } else if (!(v > old)) {
    return;
  }
  sink.lb(j, v);
}

template <class Sink>
__device__ __forceinline__ void tighten_nz(Sink &sink, const RowPlan &p, double a, int j, bool isInt,
  double lo, double hi, const Scal &s)
{
  const bool loInf = is_inf(lo, s.inf), hiInf = is_inf(hi, s.inf);
  const bool finite = !loInf && !hiInf;
  const double l = a > 0.0 ? lo : hi, u = a > 0.0 ? hi : lo;
  const double w = finite ? fabs(a) * useful_range(u - l, isInt) : 0.0;
  bool wantU = false, wantL = false;
  double vU = 0.0, vL = 0.0;
  if (p.sideU) {
    if (p.infMin == 0) {
      if (!finite || p.slackU < w) {
        vU = lo + p.slackU / a;
        wantU = true;
      }
    } else if (loInf) {
      vU = p.slackU / a;
      wantU = true;
    }
  }
  if (p.sideL) {
    if (p.infMax == 0) {
      if (!finite || p.slackL < w) {
        vL = hi - p.slackL / a;
        wantL = true;
      }
    } else if (hiInf) {
      vL = -p.slackL / a;
      wantL = true;
    }
  }
  if (!wantU && !wantL)
    return;
  if (wantU) {  // rowUB side: a>0 bounds x from above, a<0 from below
    if (a > 0.0)
      offer_ub(sink, j, vU, u, l, isInt, s);
    else
      offer_lb(sink, j, vU, l, u, isInt, s);
  }
  if (wantL) {
    if (a > 0.0)
      offer_lb(sink, j, vL, l, u, isInt, s);
    else
      offer_ub(sink, j, vL, u, l, isInt, s);
  }
}

// ---- R9 for one changed column --------------------------------------------
// returns true if the bounds cross beyond tolerance
__device__ __forceinline__ bool settle(double &L, double &U, double oldUb, const Scal &s)
{
  if (L > U) {
    if (L > U + s.feasTol)
      return true;
    if (U < oldUb) {
      const double x = L < oldUb ? L : oldUb;
      U = x;
      L = x;
    } else {
      L = U;
    }
  }
  return false;
}

// ---- R11 for one column ---------------------------------------------------
// returns 0 untouched, 1 changed, 2 crossing (bounds left as they were)
__device__ __forceinline__ int prepass_col(double &l, double &u, bool isInt, const Scal &s)
{
  const double l0 = l, u0 = u;
  int touched = 0;
  if (isInt) {
    if (fabs(l) < s.inf) {
      const double t = ceil(l - s.tol) + 0.0;
      if (t != l) {
        l = t;
        touched = 1;
      }
    }
    if (fabs(u) < s.inf) {
      const double t = floor(u + s.tol) + 0.0;
      if (t != u) {
        u = t;
        touched = 1;
      }
    }
  }
  if (l > u) {
    if (settle(l, u, u0, s)) {
      l = l0;
      u = u0;
      return 2;
    }
    touched = 1;
  }
  return touched;
}

// ---- warp / block reductions with a fixed shape ---------------------------
__device__ __forceinline__ double shfl_xor_f64(double v, int m)
{
  return __shfl_xor_sync(0xffffffffu, v, m);
}

__device__ __forceinline__ void warp_reduce(RowAct &r)
{
#pragma unroll
  for (int m = 16; m >= 1; m >>= 1) {
    RowAct o;
    o.minAct = shfl_xor_f64(r.minAct, m);
    o.maxAct = shfl_xor_f64(r.maxAct, m);
    o.maxRange = shfl_xor_f64(r.maxRange, m);
    o.infMin = 0;
    o.infMax = 0;
    combine(r, o);
  }
  // the counters as whole ints (a lane may hold thousands: rows of any length come through here)
  r.infMin = __reduce_add_sync(0xffffffffu, r.infMin);
  r.infMax = __reduce_add_sync(0xffffffffu, r.infMax);
}

// ---- TMA (bulk async copy) + mbarrier wrappers ------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init()
{
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
  asm volatile(
    "{\n"
    ".reg .pred p;\n"
    "WAIT_%=:\n"
    "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
    "@p bra DONE_%=;\n"
    "bra WAIT_%=;\n"
    "DONE_%=:\n"
    "}\n" ::"r"(smem_u32(bar)),
    "r"(parity)
    : "memory");
}
// contiguous global -> shared copy done by the TMA unit; completes on `bar`
__device__ __forceinline__ void bulk_g2s(void *sdst, const void *gsrc, uint32_t bytes, uint64_t *bar)
{
  asm volatile(
    "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
      smem_u32(sdst)),
    "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
    : "memory");
}

}  // namespace cbp
