// cbp_noalloc.cu — phase 1 (singleton rows) for callers that must not allocate.
//
// Who this is for: code that runs inside a process where device memory is not
// ours to ask for.  Two cases we know of in Cbc's world: (1) a host application
// (or an LP code sharing the GPU) that has handed every free byte of HBM to its
// own pool allocator at start-up, so a late cudaMalloc from a B&B worker thread
// simply fails; (2) node-level propagation on several worker threads, where
// cudaMalloc/cudaFree would stall every other thread's stream each time they
// are called, because both synchronise the whole device.  For those callers
// this entry point takes every array from the caller (device pointers, nothing
// is uploaded or re-tiled) and keeps its working memory in a fixed array that
// is part of the library image.
//
// The price: that array (kScratchBytes, about 4.1 MB) is resident on EVERY GPU
// that loads libcbp.so, from load until the process exits, whether or not this
// entry point is ever called; and problems with more than
// CBP_NOALLOC_MAX_COLS columns are refused.
#include <cuda_runtime.h>

#include <cstdio>
#include <mutex>

#include "../../include/cbp_capi.h"
#include "cbp_device.cuh"

using namespace cbp;

namespace {

constexpr int kMaxCols = CBP_NOALLOC_MAX_COLS;
constexpr size_t kSrcBytes = sizeof(Bnd) * static_cast<size_t>(kMaxCols);  // round-start copy of the bounds
constexpr size_t kDirtyBytes = sizeof(unsigned) * (kMaxCols / 32);         // one mark per changed column
constexpr size_t kResBytes = 256;                                          // changed / fixed / crossing counters
constexpr size_t kScratchBytes = kSrcBytes + kDirtyBytes + kResBytes;

// The working memory.  In host code this name is only a handle for the symbol,
// NOT an address a kernel can use: the device address is looked up once per
// device with cudaGetSymbolAddress (na_scratch_for_current_device) and that
// pointer is what the kernels receive.
static __device__ __align__(256) unsigned char g_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
void *g_scratchPtr[kMaxDevices];  // per device, filled on first use
bool g_scratchReady[kMaxDevices];
std::mutex g_scratchMutex;

// First call on a device: resolve the device address of g_scratch with that
// device current.  Later calls: a table lookup.  Returns nullptr on failure.
void *na_scratch_for_current_device()
{
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    fprintf(stderr, "cbp_singletons_noalloc: cannot identify the current device\n");
    return nullptr;
  }
  std::lock_guard<std::mutex> lock(g_scratchMutex);
  if (!g_scratchReady[dev]) {
    void *ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&ptr, g_scratch);
    if (e != cudaSuccess || !ptr) {
      fprintf(stderr, "cbp_singletons_noalloc: cudaGetSymbolAddress failed on device %d: %s\n", dev,
        cudaGetErrorString(e));
      return nullptr;
    }
    g_scratchPtr[dev] = ptr;
    g_scratchReady[dev] = true;
  }
  return g_scratchPtr[dev];
}

// accepted candidates go straight into the caller's lb / ub arrays
struct NaSink {
  double *lbOut, *ubOut;
  unsigned *dirty;
  int *res;
  __device__ __forceinline__ void mark(int j) { atomicOr(&dirty[j >> 5], 1u << (j & 31)); }
  __device__ __forceinline__ void ub(int j, double v)
  {
    atomic_min_f64(&ubOut[j], v);
    mark(j);
  }
  __device__ __forceinline__ void lb(int j, double v)
  {
    atomic_max_f64(&lbOut[j], v);
    mark(j);
  }
};

// The shared array still holds whatever the previous call left in it: copy
// this call's round-start bounds in, and zero the marks and counters.
__global__ void na_prepare_kernel(double *lb, double *ub, int n,
  Bnd *src, unsigned *dirty, int nWords, int *res)
{
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) {
    Bnd b;
    b.lb = lb[j];
    b.ub = ub[j];
    src[j] = b;
    // the atomics below order doubles by their bit patterns and take -0.0 for the smallest negative
    // number: the caller's arrays are the atomics' target here, so a -0.0 in them becomes +0.0
    if (b.lb == 0.0)
      lb[j] = 0.0;
    if (b.ub == 0.0)
      ub[j] = 0.0;
  }
  if (j < nWords)
    dirty[j] = 0u;
  if (j < 4)
    res[j] = 0;
}

// Variant of sv_singleton_kernel that reads the caller's row-major matrix as
// it is (no re-tiled copy exists here): one thread per row; a row takes part
// iff exactly one of its coefficients is nonzero.
__global__ void na_singleton_rows_kernel(int m, int n, const long long *__restrict__ rowStart,
  const int *__restrict__ rowLen, const int *__restrict__ colIdx, const double *__restrict__ val,
  const char *__restrict__ rowSense, const double *__restrict__ rhs, const double *__restrict__ range,
  const char *__restrict__ colType, const Bnd *__restrict__ src, NaSink sink, Scal s)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m)
    return;
  double rl, ru;
  const double r = rhs[i];
  switch (rowSense[i]) {  // R1
  case 'L': rl = -s.inf; ru = r; break;
  case 'G': rl = r; ru = s.inf; break;
  case 'E': rl = r; ru = r; break;
  case 'R': rl = r - fabs(range ? range[i] : 0.0); ru = r; break;
  default: return;
  }
  const long long kb = rowStart[i], ke = rowLen ? kb + rowLen[i] : rowStart[i + 1];
  long long at = -1;
  int cnt = 0;
  for (long long k = kb; k < ke && cnt < 2; ++k)
    if (val[k] != 0.0) {
      ++cnt;
      at = k;
    }
  if (cnt != 1)
    return;
  const double a = val[at];
  const int j = colIdx[at];
  if (j < 0 || j >= n) {  // a column index outside the problem: report it, touch nothing
    atomicAdd(&sink.res[3], 1);
    return;
  }
  const Bnd b = src[j];
  const bool isInt = colType && colType[j] != 0;
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

// R9 on the marked columns; res[0] changed, res[1] fixed, res[2] crossing (res[3]: column indices out of range)
__global__ void na_settle_kernel(const Bnd *__restrict__ src, double *lb, double *ub,
  const unsigned *__restrict__ dirty, int n, int *res, Scal s)
{
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n || !((dirty[j >> 5] >> (j & 31)) & 1u))
    return;
  double L = lb[j], U = ub[j];
  if (settle(L, U, src[j].ub, s)) {
    atomicAdd(&res[2], 1);
    return;
  }
  lb[j] = L;
  ub[j] = U;
  atomicAdd(&res[0], 1);
  if (L == U)
    atomicAdd(&res[1], 1);
}

// an infeasible phase 1 leaves the caller's bounds as they came in
__global__ void na_restore_kernel(const Bnd *__restrict__ src, double *lb, double *ub, int n)
{
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n)
    return;
  lb[j] = src[j].lb;
  ub[j] = src[j].ub;
}

}  // namespace

extern "C" int cbp_singletons_noalloc(int m, int n, const int64_t *rowStart, const int *rowLen,
  const int *colIdx, const double *val, const char *rowSense, const double *rhs, const double *range,
  const char *colType, double infinity, double *lb, double *ub, double primalTol, int *nTightened,
  int *nFixed, void *cudaStream)
{
  if (m < 0 || n <= 0 || !rowStart || !rowSense || !rhs || !lb || !ub || !nTightened || (m > 0 && (!colIdx || !val)))
    return CBP_ERR_ARG;
  if (n > kMaxCols) {
    fprintf(stderr,
      "cbp_singletons_noalloc: %d columns exceed the fixed working memory (CBP_NOALLOC_MAX_COLS = %d); "
      "use cbp_matrix_upload + cbp_singletons instead\n",
      n, kMaxCols);
    return CBP_ERR_CAPACITY;
  }
  unsigned char *base = static_cast<unsigned char *>(na_scratch_for_current_device());
  if (!base)
    return CBP_ERR_CUDA;
  Bnd *src = reinterpret_cast<Bnd *>(base);
  unsigned *dirty = reinterpret_cast<unsigned *>(base + kSrcBytes);
  int *res = reinterpret_cast<int *>(base + kSrcBytes + kDirtyBytes);
  cudaStream_t st = static_cast<cudaStream_t>(cudaStream);
  Scal s;
  s.inf = infinity;
  s.tol = primalTol;
  s.feasTol = 100.0 * primalTol;
  const int nWords = (n + 31) / 32;
  const int T = 256;
  na_prepare_kernel<<<(n + T - 1) / T, T, 0, st>>>(lb, ub, n, src, dirty, nWords, res);
  if (m > 0) {
    NaSink sink = { lb, ub, dirty, res };
    na_singleton_rows_kernel<<<(m + T - 1) / T, T, 0, st>>>(m, n, reinterpret_cast<const long long *>(rowStart),
      rowLen, colIdx, val, rowSense, rhs, range, colType, src, sink, s);
    na_settle_kernel<<<(n + T - 1) / T, T, 0, st>>>(src, lb, ub, dirty, n, res, s);
  }
  int h[4] = { 0, 0, 0, 0 };
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess)
    e = cudaMemcpyAsync(h, res, sizeof(h), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess)
    e = cudaStreamSynchronize(st);
  if (e == cudaSuccess && (h[2] > 0 || h[3] > 0)) {
    na_restore_kernel<<<(n + T - 1) / T, T, 0, st>>>(src, lb, ub, n);
    e = cudaGetLastError();
    if (e == cudaSuccess)
      e = cudaStreamSynchronize(st);
  }
  if (e != cudaSuccess) {
    fprintf(stderr, "cbp_singletons_noalloc: %s\n", cudaGetErrorString(e));
    return CBP_ERR_CUDA;
  }
  if (h[3] > 0) {
    fprintf(stderr, "cbp_singletons_noalloc: %d singleton rows name a column outside [0, %d)\n", h[3], n);
    return CBP_ERR_ARG;
  }
  if (h[2] > 0) {
    *nTightened = -1;
    if (nFixed)
      *nFixed = 0;
  } else {
    *nTightened = h[0];
    if (nFixed)
      *nFixed = h[1];
  }
  return CBP_OK;
}
