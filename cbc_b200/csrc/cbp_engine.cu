// cbp_engine.cu — host side of libcbp.so: context, matrix re-tiling and upload,
// and the C ABI declared in include/cbp_capi.h.  The kernels live in
// cbp_sv_kernels.cuh (single vector, persistent cooperative loop) and
// cbp_batch_kernels.cuh (B vectors).  No CPU fallback anywhere: every entry
// point either runs on the device or reports an error.
#include <cuda_runtime.h>

#include <algorithm>
#include <climits>
#include <cmath>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <memory>
#include <string>
#include <thread>
#include <utility>
#include <vector>

#include "../../include/cbp_capi.h"
#include "cbp_batch_kernels.cuh"
#include "cbp_sv_kernels.cuh"
#include "cbp_wl_kernels.cuh"

using namespace cbp;

// ----------------------------------------------------------------------------
struct cbp_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  bool ownsStream = false;
  int numSMs = 0;
  int svBlocksPerSM[2][2] = {};  // resident CTAs per SM of the persistent kernel, [tile staging][row-sharded]
  int wlBlocksPerSM = 0;         // ... of the worklist batch kernel
  std::string err;
  cbp_stats stats {};
  cudaEvent_t evA = nullptr, evB = nullptr;
  std::vector<cudaEvent_t> evPool;  // pairs, batch rounds
  // Blocks of freed matrices (their arrays, their batch workspace, its pinned mirror) wait here for the next matrix
  // of about the same size: a branch-and-cut run that gets a new matrix per node (the cuts changed) would otherwise
  // pay cudaMalloc / cudaHostAlloc / cudaFree of the same sizes at every node.  All work on a block is on this
  // context's stream, so the next owner's work is ordered behind the last owner's.
  struct PoolBlock {
    void *p;
    size_t bytes;
    bool host;
  };
  std::vector<PoolBlock> pool;
};

static void *pool_take(cbp_ctx *ctx, size_t bytes, bool host, size_t *got)
{
  int best = -1;
  for (int i = 0; i < static_cast<int>(ctx->pool.size()); ++i) {
    const cbp_ctx::PoolBlock &b = ctx->pool[i];
    if (b.host == host && b.bytes >= bytes && b.bytes <= 2 * bytes + (static_cast<size_t>(64) << 20)
      && (best < 0 || b.bytes < ctx->pool[best].bytes))
      best = i;
  }
  if (best < 0)
    return nullptr;
  void *p = ctx->pool[best].p;
  *got = ctx->pool[best].bytes;
  ctx->pool.erase(ctx->pool.begin() + best);
  return p;
}

// (before an allocation is given up for lack of memory: what this context holds back goes first)
static void pool_flush(cbp_ctx *ctx)
{
  for (const cbp_ctx::PoolBlock &b : ctx->pool) {
    if (b.host)
      cudaFreeHost(b.p);
    else
      cudaFree(b.p);
  }
  ctx->pool.clear();
  cudaGetLastError();
}

static void pool_give(cbp_ctx *ctx, void *p, size_t bytes, bool host)
{
  if (!p)
    return;
  size_t held = 0;
  int same = 0;
  for (const cbp_ctx::PoolBlock &b : ctx->pool)
    if (b.host == host) {
      held += b.bytes;
      ++same;
    }
  if (same >= 6 || held + bytes > (host ? (static_cast<size_t>(64) << 20) : (static_cast<size_t>(16) << 30))) {
    if (host)
      cudaFreeHost(p);
    else
      cudaFree(p);
    return;
  }
  ctx->pool.push_back({ p, bytes, host });
}

struct BatchState {
  int B = 0, Bp = 0, G = 0;
  int W = 32;  // vectors per slice of the slice-major bounds layout
  size_t cells = 0;  // n * Bp
  Bnd *X[2] = { nullptr, nullptr };
  unsigned *dirty[2] = { nullptr, nullptr };
  unsigned *ever = nullptr, *active = nullptr;
  int *ints = nullptr;  // 7 arrays of Bp ints + anyActive
  long long *counts = nullptr;  // B+1
  int *chunkCnt = nullptr;      // [ceil(n / kUpdChunk)][Bp] update counts / offsets
  BtLogEntry *log = nullptr;    // round-0 candidate list (lazy second buffer)
  unsigned logCap = 0;
  int *groupFlag = nullptr;     // [G] + 1 word for the list's counter
  Bnd *baseX = nullptr;         // first round from the base: interleaved base bounds
  unsigned *dmask = nullptr;    // ... and [G][n] which vectors differ from them at a column
  bool baseMode = false;        // the last run started from the base
  bool quiet = false;           // ... and no vector got a single candidate: no marks were ever set
  bool dirty0Clean = false, dmaskClean = false;  // known to be all zero (left so by a quiet run)
  double checkedTol = -1.0;     // R11 check of the base bounds: tolerance it was made for, result
  bool baseNeedsPrepass = true;
  std::vector<int> groupBuilt;  // ... and built the bound buffers of these groups only
  // loaded input (device copies) so the batch can be re-expanded
  double *baseLb = nullptr, *baseUb = nullptr;
  long long *dStart = nullptr;
  unsigned *dVar = nullptr;
  double *dBnd = nullptr;
  long long nDelta = 0;
  bool loaded = false, denseLoaded = false, consumed = false;
  bool hasInf = true;  // some input bound may be infinite (false only when checked)
  bool hasInfKnown = false;  // ... and whether the loaded batch has been checked (on the device, when the full-walk path needs it)
  size_t capCells = 0, capDelta = 0, capStart = 0;
  int capBp = 0;
  // worklist form (cbp_wl_kernels.cuh)
  WlItem *wlItemsS[2] = { nullptr, nullptr }, *wlItemsL[2] = { nullptr, nullptr };
  WlCell *wlK = nullptr, *wlCells = nullptr;
  WlWalk *wlWalk = nullptr;
  unsigned wlWalkCap = 0;
  WlCtl *wlCtl = nullptr;
  int *wlInts = nullptr;  // vecChanged by round parity [2][Bp], active by round parity [2][G]
  // device staging of the update lists for host-pointer fetches (kept between calls: no cudaMalloc per fetch)
  int *stCol = nullptr;
  double *stLb = nullptr, *stUb = nullptr;
  long long stCap = 0;
  int *st2Col = nullptr;        // second set: the unsorted lists of a worklist run's fetch
  double *st2Lb = nullptr, *st2Ub = nullptr;
  long long st2Cap = 0;
  bool lastWasWorklist = false; // the last run was in worklist form and `cells` lists every cell it created
  unsigned wlItemCap = 0, wlCellCap = 0;
  WlPair *wlHKey = nullptr;  // (vector, row) pair table
  unsigned wlHActive = 0;    // slots the next run hashes into (a power of two <= wlHSlots): what the last run took x 3,
                             // so that a batch that touches few pairs keeps its table in L2
  unsigned wlHSlots = 0;
  unsigned *wlUsed = nullptr;   // table slots the running worklist run takes (emptied again after it)
  unsigned wlUsedCap = 0;
  // a worklist run leaves the vectors' update lists in stCol/stLb/stUb and copies what a fetch reads first into
  // pinned host memory before its one synchronisation
  bool updReady = false;        // the last run wrote the lists itself
  bool listsWanted = false;     // a fetch on this matrix has asked for update lists: runs write them from now on
  long long updTotal = 0;
  // control block | per-vector results | list offsets | head of the lists: one device block (wlCtl, ints and counts
  // point into it), so that ONE copy brings all a fetch needs to its pinned mirror
  unsigned char *arena = nullptr;  // the batch workspace below is one device allocation (bt_reserve)
  size_t arenaBytes = 0, hBlobBytes = 0;
  bool stInArena = false, st2InArena = false;  // the two sets of update-list buffers still are the arena's default ones
  unsigned char *blob = nullptr, *hBlob = nullptr;
  size_t blobOffInts = 0, blobOffCounts = 0, blobOffHead = 0, blobBytes = 0;
  long long headCopied = 0;     // list entries of the last run that are in the mirror
  bool wlTableClean = false;    // every slot of the table is free, except the wlPrevUsed listed in wlUsed
  unsigned wlPrevUsed = 0;
  WlRowBase *wlRowBase = nullptr;  // [mKept] the node's sums of the rows
  unsigned wlPrevCells = 0;  // cells the last worklist run left in the masks
  bool wlMarksKnown = false; // dmask / ever / dirty are all zero except for those cells
};

// (the default-sized list buffers are part of the batch arena; grown ones are allocations of their own)
static void bt_free_lists(BatchState &bt, bool first, bool second)
{
  if (first) {
    if (!bt.stInArena) {
      void *q[] = { bt.stCol, bt.stLb, bt.stUb };
      for (void *x : q)
        if (x)
          cudaFree(x);
    }
    bt.stCol = nullptr;
    bt.stLb = bt.stUb = nullptr;
    bt.stCap = 0;
    bt.stInArena = false;
  }
  if (second) {
    if (!bt.st2InArena) {
      void *q[] = { bt.st2Col, bt.st2Lb, bt.st2Ub };
      for (void *x : q)
        if (x)
          cudaFree(x);
    }
    bt.st2Col = nullptr;
    bt.st2Lb = bt.st2Ub = nullptr;
    bt.st2Cap = 0;
    bt.st2InArena = false;
  }
}

struct cbp_matrix {
  cbp_ctx *ctx = nullptr;
  int m = 0, n = 0, mKept = 0;
  long long nnz = 0, nnzPacked = 0;
  double inf = 1e30;
  bool allInt = false;  // no continuous column
  // device, matrix
  double *val = nullptr;
  int *col = nullptr;
  Tile *tiles = nullptr;
  int *rp = nullptr;
  double *rowLB = nullptr, *rowUB = nullptr;
  int *rowOrig = nullptr, *rowNz = nullptr, *rowCnt = nullptr, *rowPos = nullptr;
  LongRow *longRows = nullptr;
  signed char *colType = nullptr;
  int *singRows = nullptr;
  int *shortRowsBt = nullptr, *longRowsBt = nullptr;
  int nShortBt = 0, nLongBt = 0;
  int nTiles = 0, firstLongTile = 0, nLong = 0, nParts = 0, nSing = 0;
  double *sellVal = nullptr;
  int *sellCol = nullptr, *sellOff = nullptr, *sliceOrder = nullptr;
  int nSlices = 0, mShort = 0;
  // column-major copy: rows (kept order) of every column, coefficient beside it
  int *cscStart = nullptr, *cscRow = nullptr;
  double *cscVal = nullptr;
  int firstLongRow = 0;     // kept rows from here on are longer than a tile
  int eventMaxChanged = 0;  // single vector: rounds after a round that changed at most this many columns are event-driven
  // device, single-vector workspace
  RowAct *longPart = nullptr;
  // everything an upload allocates comes out of one device allocation (a second one only if the estimate fell short):
  // thirty cudaMallocs were a quarter of a small upload
  std::vector<std::pair<void *, size_t>> arenaChunks;
  unsigned char *arenaCur = nullptr;
  size_t arenaLeft = 0, arenaHint = 0;
  Bnd *X[2] = { nullptr, nullptr };
  unsigned *dirty[2] = { nullptr, nullptr };
  int nWords = 0;
  SvStatus *st = nullptr;
  double *dLb = nullptr, *dUb = nullptr;  // staging for host-pointer calls
  int *dRes = nullptr;                    // small int scratch
  double *dAct = nullptr;                 // row-activity staging (2m doubles + 2m ints)
  // row-sharded runs over peer memory: one state block per rank, mapped into every rank
  struct P2P {
    // own block: X0 | X1 | dirty0 | dirty1 | proposals [2][world][n] | their marks [2][world][words] | status
    unsigned char *block = nullptr;
    size_t offX1 = 0, offD0 = 0, offD1 = 0, offP = 0, offM = 0, offSt = 0, bytes = 0;
    size_t strideP = 0, strideD = 0;  // elements between two ranks' slots
    int world = 0, rank = 0;
    unsigned char *peer[kMaxPeers] = {};  // every rank's block as seen from here (peer[rank] == block)
    unsigned long long epoch = 0;  // cross-GPU barriers passed so far (the same number on every rank)
    bool attached = false;
    bool broken = false;  // a rank timed out: the group cannot be used again
    bool lastRunHere = false;  // the latest single-vector run used this block (for cbp_sv_round_stamps)
  } p2p;
  BatchState bt;
};

#define CK(call)                                                                         \
  do {                                                                                   \
    cudaError_t e_ = (call);                                                             \
    if (e_ != cudaSuccess) {                                                             \
      ctx->err = std::string(#call) + ": " + cudaGetErrorString(e_);                     \
      return CBP_ERR_CUDA;                                                               \
    }                                                                                    \
  } while (0)

static thread_local std::string g_createErr;

// Test / A-B switches, read from the environment once per process (never inside a round loop).
struct EnvKnobs {
  bool noWorklist, noLazyX1, lazyX1, noBaseRound, noMultiGroup, wlStamps;
  int overlayRounds;
  int wlTableFactor;  // CBP_WL_TABLE_FACTOR: slots per pair the last run took (0 = always the whole table)
  EnvKnobs()
    : noWorklist(getenv("CBP_NO_WORKLIST") != nullptr)
    , noLazyX1(getenv("CBP_NO_LAZY_X1") != nullptr)
    , lazyX1(getenv("CBP_LAZY_X1") != nullptr)
    , noBaseRound(getenv("CBP_NO_BASE_ROUND") != nullptr)
    , noMultiGroup(getenv("CBP_NO_MULTI_GROUP") != nullptr)
    , wlStamps(getenv("CBP_WL_STAMPS") != nullptr)
    , overlayRounds(getenv("CBP_OVERLAY_ROUNDS") ? atoi(getenv("CBP_OVERLAY_ROUNDS")) : INT_MAX)
    , wlTableFactor(getenv("CBP_WL_TABLE_FACTOR") ? atoi(getenv("CBP_WL_TABLE_FACTOR")) : 16)
  {
  }
};
static const EnvKnobs &knobs()
{
  static const EnvKnobs k;
  return k;
}

static constexpr long long kListDefault = 1 << 16;  // update-list entries the batch arena has room for (a fetch grows beyond)
static constexpr long long kUpdPrefetch = 32768;  // update-list entries a worklist run copies to the host unasked
static constexpr size_t kSvSmallSmem = 0;
static constexpr unsigned kBtLogCap = 16u << 20;  // candidate list of a list round: 16M entries (256 MB)  // the short-rows-only variant needs no dynamic shared memory

template <class T>
static cudaError_t dalloc(T **p, size_t count)
{
  return cudaMalloc(reinterpret_cast<void **>(p), std::max<size_t>(count, 1) * sizeof(T));
}

template <class T>
static cudaError_t mat_alloc(cbp_matrix *mat, T **p, size_t count)
{
  const size_t bytes = (std::max<size_t>(count, 1) * sizeof(T) + 255) & ~static_cast<size_t>(255);
  if (bytes > mat->arenaLeft) {
    size_t sz = std::max(bytes, std::max<size_t>(mat->arenaHint, static_cast<size_t>(1) << 20));
    void *q = pool_take(mat->ctx, sz, false, &sz);
    if (!q) {
      cudaError_t e = cudaMalloc(&q, sz);
      if (e != cudaSuccess) {
        pool_flush(mat->ctx);
        e = cudaMalloc(&q, sz);
      }
      if (e != cudaSuccess)
        return e;
    }
    mat->arenaChunks.push_back(std::make_pair(q, sz));
    mat->arenaCur = static_cast<unsigned char *>(q);
    mat->arenaLeft = sz;
    mat->arenaHint = sz / 4;  // (should the estimate have been short: a quarter as much again)
  }
  *p = reinterpret_cast<T *>(mat->arenaCur);
  mat->arenaCur += bytes;
  mat->arenaLeft -= bytes;
  return cudaSuccess;
}

static Scal make_scal(const cbp_matrix *mat, double tol)
{
  Scal s;
  s.inf = mat->inf;
  s.tol = tol;
  s.feasTol = 100.0 * tol;
  return s;
}

static inline int cdiv(long long a, long long b) { return static_cast<int>((a + b - 1) / b); }

// std::vector whose resize() does not zero: the big host arrays of an upload are written in full by the threads that
// fill them (who then also take their page faults), a zero fill by one thread first would cost as much as the fill.
template <class T> struct NoInitAlloc : std::allocator<T> {
  template <class U> struct rebind {
    using other = NoInitAlloc<U>;
  };
  NoInitAlloc() = default;
  template <class U> NoInitAlloc(const NoInitAlloc<U> &) {}
  template <class U> void construct(U *p) { ::new (static_cast<void *>(p)) U; }
  template <class U, class... A> void construct(U *p, A &&...a) { ::new (static_cast<void *>(p)) U(std::forward<A>(a)...); }
};
template <class T> using RawVec = std::vector<T, NoInitAlloc<T>>;

// Host side of an upload in parallel: f(begin, end, thread) over [0, n) cut into one piece per thread (the pieces are
// multiples of `grain`).  Small jobs run on the caller's thread.
template <class F> static void par_for(size_t n, int threads, size_t grain, F f)
{
  if (threads <= 1 || n <= grain) {
    f(static_cast<size_t>(0), n, 0);
    return;
  }
  size_t chunk = (n + threads - 1) / threads;
  chunk = (chunk + grain - 1) / grain * grain;
  std::vector<std::thread> th;
  int t = 0;
  for (size_t b = 0; b < n; b += chunk, ++t)
    th.emplace_back(f, b, std::min(n, b + chunk), t);
  for (std::thread &x : th)
    x.join();
}


// ----------------------------------------------------------------------------
extern "C" {

cbp_ctx *cbp_create(int device)
{
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    g_createErr = std::string("no CUDA device: ") + cudaGetErrorString(e);
    fprintf(stderr, "cbp_create: %s\n", g_createErr.c_str());
    return nullptr;
  }
  if (device < 0)
    cudaGetDevice(&device);
  if (cudaSetDevice(device) != cudaSuccess)
    return nullptr;
  cbp_ctx *ctx = new cbp_ctx();
  ctx->device = device;
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, device);
  ctx->numSMs = prop.multiProcessorCount;
  if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
    delete ctx;
    return nullptr;
  }
  ctx->ownsStream = true;
  cudaEventCreate(&ctx->evA);
  cudaEventCreate(&ctx->evB);
  cudaFuncSetAttribute(sv_fixpoint_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
    static_cast<int>(sizeof(SvSmem)));
  cudaFuncSetAttribute(sv_fixpoint_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
    static_cast<int>(sizeof(SvSmem)));
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctx->svBlocksPerSM[1][0], sv_fixpoint_kernel<true, false>, kThreads,
    sizeof(SvSmem));
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctx->svBlocksPerSM[1][1], sv_fixpoint_kernel<true, true>, kThreads,
    sizeof(SvSmem));
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctx->svBlocksPerSM[0][0], sv_fixpoint_kernel<false, false>, kThreads,
    kSvSmallSmem);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctx->svBlocksPerSM[0][1], sv_fixpoint_kernel<false, true>, kThreads,
    kSvSmallSmem);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctx->wlBlocksPerSM, bt_worklist_kernel, kThreads, 0);
  if (ctx->svBlocksPerSM[0][0] < 1 || ctx->svBlocksPerSM[0][1] < 1 || ctx->svBlocksPerSM[1][0] < 1
    || ctx->svBlocksPerSM[1][1] < 1) {
    fprintf(stderr, "cbp_create: persistent kernel does not fit on this device\n");
    cbp_destroy(ctx);
    return nullptr;
  }
  return ctx;
}

void cbp_destroy(cbp_ctx *ctx)
{
  if (!ctx)
    return;
  cudaSetDevice(ctx->device);
  if (ctx->evA)
    cudaEventDestroy(ctx->evA);
  if (ctx->evB)
    cudaEventDestroy(ctx->evB);
  for (cudaEvent_t e : ctx->evPool)
    cudaEventDestroy(e);
  cudaDeviceSynchronize();
  for (const cbp_ctx::PoolBlock &b : ctx->pool) {
    if (b.host)
      cudaFreeHost(b.p);
    else
      cudaFree(b.p);
  }
  ctx->pool.clear();
  if (ctx->ownsStream && ctx->stream)
    cudaStreamDestroy(ctx->stream);
  delete ctx;
}

int cbp_set_stream(cbp_ctx *ctx, void *cudaStream)
{
  if (!ctx)
    return CBP_ERR_ARG;
  if (ctx->ownsStream && ctx->stream)
    cudaStreamDestroy(ctx->stream);
  ctx->stream = static_cast<cudaStream_t>(cudaStream);
  ctx->ownsStream = false;
  return CBP_OK;
}

const char *cbp_last_error(const cbp_ctx *ctx) { return ctx ? ctx->err.c_str() : g_createErr.c_str(); }

int cbp_get_stats(const cbp_ctx *ctx, cbp_stats *out)
{
  if (!ctx || !out)
    return CBP_ERR_ARG;
  *out = ctx->stats;
  return CBP_OK;
}

int cbp_synchronize(cbp_ctx *ctx)
{
  if (!ctx)
    return CBP_ERR_ARG;
  CK(cudaStreamSynchronize(ctx->stream));
  return CBP_OK;
}

// ----------------------------------------------------------------------------
// Matrix upload.  Rows are split into three classes by nonzero count and laid
// out class by class; inside a class the caller's row order is kept.  Zero
// coefficients, 'N' rows and empty rows are dropped (R1, R3).
// ----------------------------------------------------------------------------
void cbp_matrix_free(cbp_matrix *mat)
{
  if (!mat)
    return;
  cudaSetDevice(mat->ctx->device);
  for (const std::pair<void *, size_t> &q : mat->arenaChunks)  // (the matrix's own arrays and single-vector state)
    pool_give(mat->ctx, q.first, q.second, false);
  pool_give(mat->ctx, mat->bt.arena, mat->bt.arenaBytes, false);
  mat->bt.arena = nullptr;
  bt_free_lists(mat->bt, true, true);
  void *ptrs[] = { mat->bt.dStart, mat->bt.dVar, mat->bt.dBnd };
  for (void *p : ptrs)
    if (p)
      cudaFree(p);
  pool_give(mat->ctx, mat->bt.hBlob, mat->bt.hBlobBytes, true);
  for (int r = 0; r < mat->p2p.world; ++r)
    if (mat->p2p.attached && r != mat->p2p.rank && mat->p2p.peer[r])
      cudaIpcCloseMemHandle(mat->p2p.peer[r]);
  if (mat->p2p.block)
    cudaFree(mat->p2p.block);
  delete mat;
}

cbp_matrix *cbp_matrix_upload(cbp_ctx *ctx, int m, int n, const int64_t *rowStart, const int *rowLen,
  const int *colIdx, const double *val, const char *rowSense, const double *rhs, const double *range,
  const char *colType, double infinity)
{
  if (!ctx)
    return nullptr;
  if (m < 0 || n <= 0 || !rowStart || !rowSense || !rhs || (m > 0 && (!colIdx || !val))) {
    ctx->err = "cbp_matrix_upload: bad argument";
    return nullptr;
  }
  cudaSetDevice(ctx->device);
  const bool upTimes = getenv("CBP_UPLOAD_TIMES") != nullptr;  // tuning aid: where an upload spends its time
  auto upT0 = std::chrono::steady_clock::now();
  auto lap = [&](const char *what) {
    if (!upTimes)
      return;
    const auto t = std::chrono::steady_clock::now();
    fprintf(stderr, "  upload %-28s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(t - upT0).count());
    upT0 = t;
  };
  const double INF = infinity;
  struct KRow {
    int orig, cnt;
    double rl, ru;
  };
  // (re-tiling a matrix is host work proportional to its nonzeros: the passes over them run on several threads —
  // CBP_UPLOAD_THREADS, default min(cores, 16) — and give the same arrays whatever the number of threads)
  long long nnzIn = 0;
  if (m > 0)
    nnzIn = rowLen ? 0 : rowStart[m] - rowStart[0];
  if (rowLen)
    for (int i = 0; i < m; ++i)
      nnzIn += rowLen[i];
  static const int upThreadsEnv = getenv("CBP_UPLOAD_THREADS") ? atoi(getenv("CBP_UPLOAD_THREADS")) : 0;
  const int upThreads = nnzIn < 200000 ? 1
    : std::max(1, upThreadsEnv > 0 ? upThreadsEnv : std::min(16, static_cast<int>(std::thread::hardware_concurrency())));
  std::vector<KRow> cls[3];
  long long nnzKept = 0;
  std::vector<int> rawCnt(static_cast<size_t>(std::max(m, 0)), 0);
  std::atomic<int> badCol(0);
  par_for(static_cast<size_t>(m), upThreads, 1024, [&](size_t b, size_t e2, int) {
    for (size_t i = b; i < e2; ++i) {
      const char sn = rowSense[i];
      if (sn != 'L' && sn != 'G' && sn != 'E' && sn != 'R')
        continue;
      const int64_t s = rowStart[i], e = rowLen ? s + rowLen[i] : rowStart[i + 1];
      int cnt = 0;
      for (int64_t k = s; k < e; ++k)
        if (val[k] != 0.0) {
          if (colIdx[k] < 0 || colIdx[k] >= n)
            badCol.store(1, std::memory_order_relaxed);
          ++cnt;
        }
      rawCnt[i] = cnt;
    }
  });
  if (badCol.load()) {
    ctx->err = "cbp_matrix_upload: column index out of range";
    return nullptr;
  }
  for (int i = 0; i < m; ++i) {
    double rl, ru;
    const double r = rhs[i];
    switch (rowSense[i]) {
    case 'L': rl = -INF; ru = r; break;
    case 'G': rl = r; ru = INF; break;
    case 'E': rl = r; ru = r; break;
    case 'R': rl = r - std::fabs(range ? range[i] : 0.0); ru = r; break;
    default: continue;
    }
    const int cnt = rawCnt[i];
    // a row without nonzeros says rl <= 0 <= ru; keep it only if that can fail (R4)
    if (cnt == 0 && !(ru < 0.0 || rl > 0.0))
      continue;
    const int c = cnt <= kShortLen ? kShort : (cnt <= kTileNnz ? kMedium : kLong);
    cls[c].push_back({ i, cnt, rl, ru });
    nnzKept += cnt;
  }
  lap("classify rows");
  const size_t mKept = cls[0].size() + cls[1].size() + cls[2].size();
  if (nnzKept + static_cast<long long>(kAlign) * (static_cast<long long>(mKept) + 8) > INT_MAX) {
    ctx->err = "cbp_matrix_upload: more than 2^31 nonzeros are not supported";
    return nullptr;
  }
  RawVec<double> hVal;
  RawVec<int> hCol;
  std::vector<double> hRl(mKept), hRu(mKept);
  std::vector<int> hRp, hOrig(mKept), hNz(mKept), hCnt(mKept), hPos(mKept, 0), hSing;
  std::vector<Tile> hTiles;
  std::vector<LongRow> hLong;
  hVal.reserve(nnzKept + nnzKept / 64 + 64);
  hCol.reserve(nnzKept + nnzKept / 64 + 64);
  auto pad_to_align = [&]() {
    while (hVal.size() % kAlign) {
      hVal.push_back(0.0);
      hCol.push_back(0);
    }
  };
  auto emit_row = [&](const KRow &kr, int keptIdx) {
    hOrig[keptIdx] = kr.orig;
    hRl[keptIdx] = kr.rl;
    hRu[keptIdx] = kr.ru;
    hNz[keptIdx] = static_cast<int>(hVal.size());
    hCnt[keptIdx] = kr.cnt;
    const int64_t s = rowStart[kr.orig], e = rowLen ? s + rowLen[kr.orig] : rowStart[kr.orig + 1];
    // rows longer than kBtLongLen are summed piecewise anyway (no CSR-order
    // promise), so they are stored positives first: two passes over the row
    const int passes = kr.cnt > kBtLongLen ? 2 : 1;
    for (int pass = 0; pass < passes; ++pass) {
      for (int64_t k = s; k < e; ++k) {
        if (val[k] == 0.0 || (passes == 2 && (val[k] > 0.0) != (pass == 0)))
          continue;
        hVal.push_back(val[k]);
        const bool isInt = colType && colType[colIdx[k]] != 0;
        hCol.push_back(isInt ? (colIdx[k] | INT_MIN) : colIdx[k]);  // bit 31 = integer column
      }
      if (pass == 0 && passes == 2)
        hPos[keptIdx] = static_cast<int>(hVal.size()) - hNz[keptIdx];
    }
    if (kr.cnt == 1)
      hSing.push_back(keptIdx);
  };
  int kept = 0;
  // short rows: sorted by length inside windows (little padding in the slices,
  // the caller's order survives at window scale), stored once in packed CSR
  // form (batch kernels, phase 1) and once as sliced-ELL (single-vector loop)
  RawVec<double> hSellVal;
  RawVec<int> hSellCol;
  std::vector<int> hSellOff;
  {
    std::vector<KRow> &rows = cls[kShort];
    par_for((rows.size() + kSellWindow - 1) / kSellWindow, upThreads, 1, [&](size_t wb, size_t we, int) {
      for (size_t wi = wb; wi < we; ++wi) {
        const size_t w = wi * kSellWindow, e = std::min(rows.size(), w + kSellWindow);
        std::stable_sort(rows.begin() + w, rows.begin() + e, [](const KRow &a, const KRow &b) { return a.cnt > b.cnt; });
      }
    });
    const int mS = static_cast<int>(rows.size());
    {  // where every short row's entries go in the packed arrays (no padding between short rows)
      int at = 0;
      for (int r = 0; r < mS; ++r) {
        hNz[r] = at;
        at += rows[r].cnt;
      }
      hVal.resize(static_cast<size_t>(at));
      hCol.resize(static_cast<size_t>(at));
    }
    const int nSlices = (mS + 31) / 32;
    hSellOff.assign(nSlices + 1, 0);
    for (int sl = 0; sl < nSlices; ++sl) {
      int len = 0;
      for (int r = sl * 32; r < std::min(mS, sl * 32 + 32); ++r)
        len = std::max(len, rows[r].cnt);
      hSellOff[sl + 1] = hSellOff[sl] + 32 * len;
    }
    hSellVal.resize(static_cast<size_t>(hSellOff[nSlices]) + 32);  // (not zeroed: every entry is written below)
    hSellCol.resize(static_cast<size_t>(hSellOff[nSlices]) + 32);
    for (size_t i = static_cast<size_t>(hSellOff[nSlices]); i < hSellVal.size(); ++i) {
      hSellVal[i] = 0.0;
      hSellCol[i] = 0;
    }
    std::vector<std::vector<int>> singOf(static_cast<size_t>(std::max(upThreads, 1)));
    par_for(static_cast<size_t>(mS), upThreads, 32, [&](size_t rb, size_t re, int t) {  // (whole slices per thread)
      for (size_t r = rb; r < re; ++r) {
        const KRow &kr = rows[r];
        hOrig[r] = kr.orig;
        hRl[r] = kr.rl;
        hRu[r] = kr.ru;
        hCnt[r] = kr.cnt;
        const int64_t s = rowStart[kr.orig], e = rowLen ? s + rowLen[kr.orig] : rowStart[kr.orig + 1];
        size_t at = static_cast<size_t>(hNz[r]);
        const size_t base = static_cast<size_t>(hSellOff[r >> 5]) + (r & 31);
        int k2 = 0;
        for (int64_t k = s; k < e; ++k) {
          if (val[k] == 0.0)
            continue;
          const bool isInt = colType && colType[colIdx[k]] != 0;
          const int c = isInt ? (colIdx[k] | INT_MIN) : colIdx[k];  // bit 31 = integer column
          hVal[at] = val[k];
          hCol[at] = c;
          ++at;
          hSellVal[base + 32 * static_cast<size_t>(k2)] = val[k];
          hSellCol[base + 32 * static_cast<size_t>(k2)] = c;
          ++k2;
        }
        const int len = (hSellOff[(r >> 5) + 1] - hSellOff[r >> 5]) >> 5;
        for (; k2 < len; ++k2) {  // the slice's padding behind this row
          hSellVal[base + 32 * static_cast<size_t>(k2)] = 0.0;
          hSellCol[base + 32 * static_cast<size_t>(k2)] = 0;
        }
        if (kr.cnt == 1)
          singOf[static_cast<size_t>(t)].push_back(static_cast<int>(r));
      }
      if (re == static_cast<size_t>(mS) && (mS & 31)) {  // the lanes of the last slice that have no row
        const int sl = mS >> 5, len = (hSellOff[sl + 1] - hSellOff[sl]) >> 5;
        for (int l = mS & 31; l < 32; ++l)
          for (int k2 = 0; k2 < len; ++k2) {
            hSellVal[static_cast<size_t>(hSellOff[sl]) + 32 * static_cast<size_t>(k2) + l] = 0.0;
            hSellCol[static_cast<size_t>(hSellOff[sl]) + 32 * static_cast<size_t>(k2) + l] = 0;
          }
      }
    });
    for (const std::vector<int> &v : singOf)  // (thread pieces are ascending row ranges: ascending in all)
      hSing.insert(hSing.end(), v.begin(), v.end());
    kept = mS;
  }
  lap("short rows: packed + SELL");
  {  // medium rows: greedy tiles of whole rows, staged through shared memory
    const int c = kMedium;
    size_t i = 0;
    const std::vector<KRow> &rows = cls[c];
    while (i < rows.size()) {
      pad_to_align();
      Tile T {};
      T.nzBegin = static_cast<int>(hVal.size());
      T.rowBegin = kept;
      T.rpOff = static_cast<int>(hRp.size());
      T.kind = c;
      int nz = 0, nr = 0;
      while (i < rows.size() && nr < kTileRows && nz + rows[i].cnt <= kTileNnz) {
        hRp.push_back(nz);
        emit_row(rows[i], kept);
        nz += rows[i].cnt;
        ++nr;
        ++kept;
        ++i;
      }
      hRp.push_back(nz);
      T.nzCount = nz;
      T.rowCount = nr;
      hTiles.push_back(T);
    }
  }
  const int firstLongTile = static_cast<int>(hTiles.size());
  int nParts = 0;
  for (const KRow &kr : cls[2]) {  // long rows: chunks of kTileNnz
    pad_to_align();
    const int begin = static_cast<int>(hVal.size());
    emit_row(kr, kept);
    LongRow lr {};
    lr.row = kept;
    lr.firstPart = nParts;
    lr.nChunks = (kr.cnt + kTileNnz - 1) / kTileNnz;
    for (int ch = 0; ch < lr.nChunks; ++ch) {
      Tile T {};
      T.nzBegin = begin + ch * kTileNnz;
      T.nzCount = std::min(kTileNnz, kr.cnt - ch * kTileNnz);
      T.rowBegin = kept;
      T.rowCount = 1;
      T.rpOff = 0;
      T.kind = kLong;
      T.longIdx = static_cast<int>(hLong.size());
      T.chunkIdx = ch;
      hTiles.push_back(T);
    }
    nParts += lr.nChunks;
    hLong.push_back(lr);
    ++kept;
  }
  for (int i = 0; i < 2 * kAlign; ++i) {  // slack for the 16-byte rounding of the last copy
    hVal.push_back(0.0);
    hCol.push_back(0);
  }

  lap("medium / long rows");
  cbp_matrix *mat = new cbp_matrix();
  mat->ctx = ctx;
  mat->m = m;
  mat->n = n;
  mat->mKept = static_cast<int>(mKept);
  mat->nnz = nnzKept;
  mat->nnzPacked = static_cast<long long>(hVal.size());
  mat->inf = INF;
  mat->nTiles = static_cast<int>(hTiles.size());
  mat->firstLongTile = firstLongTile;
  mat->nLong = static_cast<int>(hLong.size());
  mat->nParts = nParts;
  mat->nSing = static_cast<int>(hSing.size());
  mat->nWords = (n + 31) / 32;
  mat->allInt = colType != nullptr;
  for (int j = 0; j < n && mat->allInt; ++j)
    mat->allInt = colType[j] != 0;
  std::vector<signed char> hType(n, 0);
  if (colType)
    for (int j = 0; j < n; ++j)
      hType[j] = static_cast<signed char>(colType[j]);

  // (what the arrays below will take, within a per cent: packed CSR, sliced-ELL, column-major copy, per-row and
  // per-column arrays, the single-vector state)
  mat->arenaHint = 12 * (hVal.size() + hSellVal.size() + static_cast<size_t>(nnzKept)) + 64 * (mKept + 64)
    + 80 * (static_cast<size_t>(n) + 64) + 32 * (static_cast<size_t>(std::max(m, 1)) + 64) + sizeof(Tile) * (hTiles.size() + 8)
    + 4 * hRp.size() + sizeof(RowAct) * (static_cast<size_t>(nParts) + 8) + sizeof(SvStatus) + 48 * 256 + (64 << 10);
  bool ok = true;
  auto up = [&](auto **dptr, const auto &h) {
    using T = typename std::remove_reference<decltype(h)>::type::value_type;
    if (!ok)
      return;
    if (mat_alloc(mat, reinterpret_cast<T **>(dptr), h.size()) != cudaSuccess) {
      ok = false;
      return;
    }
    if (!h.empty()
      && cudaMemcpyAsync(*dptr, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, ctx->stream)
        != cudaSuccess)
      ok = false;
  };
  up(&mat->val, hVal);
  up(&mat->col, hCol);
  up(&mat->tiles, hTiles);
  up(&mat->rp, hRp);
  up(&mat->rowLB, hRl);
  up(&mat->rowUB, hRu);
  up(&mat->rowOrig, hOrig);
  up(&mat->rowNz, hNz);
  up(&mat->rowCnt, hCnt);
  up(&mat->rowPos, hPos);
  up(&mat->longRows, hLong);
  up(&mat->colType, hType);
  up(&mat->singRows, hSing);
  up(&mat->sellVal, hSellVal);
  up(&mat->sellCol, hSellCol);
  up(&mat->sellOff, hSellOff);
  lap("allocate + copy (CSR, SELL)");
  std::vector<int> hOrder(hSellOff.size() - 1);
  for (size_t i = 0; i < hOrder.size(); ++i)
    hOrder[i] = static_cast<int>(i);
  std::stable_sort(hOrder.begin(), hOrder.end(), [&](int a, int b) {
    return hSellOff[a + 1] - hSellOff[a] > hSellOff[b + 1] - hSellOff[b];
  });
  up(&mat->sliceOrder, hOrder);
  mat->nSlices = static_cast<int>(hSellOff.size()) - 1;
  mat->mShort = static_cast<int>(cls[kShort].size());
  mat->firstLongRow = static_cast<int>(cls[kShort].size() + cls[kMedium].size());
  {  // column-major copy of the kept rows (the CSC of north_star's "CSR and CSC copies in HBM")
    // Every thread takes a range of kept rows: its own count per column, then — columns in parallel — where each
    // thread's entries of a column start (thread ranges ascend, so a column's rows end up in ascending order whatever
    // the number of threads), then the fill through the thread's own cursors.  No atomics, no sorting.
    const int cscT = static_cast<int>(std::max<size_t>(1, std::min<size_t>(static_cast<size_t>(upThreads),
      (static_cast<size_t>(256) << 20) / (sizeof(int) * (static_cast<size_t>(n) + 1)))));
    const size_t rowsPer = (mKept + cscT - 1) / std::max(cscT, 1) + 1;
    std::vector<std::vector<int>> cur(static_cast<size_t>(cscT));
    par_for(static_cast<size_t>(cscT), cscT, 1, [&](size_t tb, size_t te, int) {
      for (size_t t = tb; t < te; ++t) {
        std::vector<int> &h = cur[t];
        h.assign(static_cast<size_t>(n), 0);
        const size_t rb = std::min(mKept, t * rowsPer), re = std::min(mKept, rb + rowsPer);
        for (size_t r = rb; r < re; ++r)
          for (int k = hNz[r]; k < hNz[r] + hCnt[r]; ++k)
            ++h[static_cast<size_t>(hCol[k] & kColMask)];
      }
    });
    std::vector<int> hCscStart(static_cast<size_t>(n) + 1, 0);
    {
      // per column chunk: the chunk's total; then the chunks' bases; then every (thread, column) cursor
      const size_t colChunk = 8192;
      const size_t nChunks = (static_cast<size_t>(n) + colChunk - 1) / colChunk;
      std::vector<long long> chunkSum(nChunks + 1, 0);
      par_for(nChunks, upThreads, 1, [&](size_t cb, size_t ce, int) {
        for (size_t c = cb; c < ce; ++c) {
          long long tot = 0;
          for (size_t j = c * colChunk; j < std::min(static_cast<size_t>(n), (c + 1) * colChunk); ++j)
            for (int t = 0; t < cscT; ++t)
              tot += cur[static_cast<size_t>(t)][j];
          chunkSum[c + 1] = tot;
        }
      });
      for (size_t c = 0; c < nChunks; ++c)
        chunkSum[c + 1] += chunkSum[c];
      par_for(nChunks, upThreads, 1, [&](size_t cb, size_t ce, int) {
        for (size_t c = cb; c < ce; ++c) {
          int at = static_cast<int>(chunkSum[c]);
          for (size_t j = c * colChunk; j < std::min(static_cast<size_t>(n), (c + 1) * colChunk); ++j) {
            hCscStart[j] = at;
            for (int t = 0; t < cscT; ++t) {
              const int cntTj = cur[static_cast<size_t>(t)][j];
              cur[static_cast<size_t>(t)][j] = at;
              at += cntTj;
            }
          }
        }
      });
      hCscStart[static_cast<size_t>(n)] = static_cast<int>(chunkSum[nChunks]);
    }
    RawVec<int> hCscRow;
    RawVec<double> hCscVal;
    hCscRow.resize(static_cast<size_t>(hCscStart[n]) + 1);  // (not zeroed: the fill writes every entry but the spare one)
    hCscVal.resize(static_cast<size_t>(hCscStart[n]) + 1);
    hCscRow[static_cast<size_t>(hCscStart[n])] = 0;
    hCscVal[static_cast<size_t>(hCscStart[n])] = 0.0;
    par_for(static_cast<size_t>(cscT), cscT, 1, [&](size_t tb, size_t te, int) {
      for (size_t t = tb; t < te; ++t) {
        std::vector<int> &h = cur[t];
        const size_t rb = std::min(mKept, t * rowsPer), re = std::min(mKept, rb + rowsPer);
        for (size_t r = rb; r < re; ++r)
          for (int k = hNz[r]; k < hNz[r] + hCnt[r]; ++k) {
            const int at = h[static_cast<size_t>(hCol[k] & kColMask)]++;
            hCscRow[at] = static_cast<int>(r);
            hCscVal[at] = hVal[k];
          }
      }
    });
    lap("column-major copy: build");
    up(&mat->cscStart, hCscStart);
    up(&mat->cscRow, hCscRow);
    up(&mat->cscVal, hCscVal);
    lap("column-major copy: upload");
    // Event-driven rounds pay off while (changed columns) x (rows per column) x (nonzeros per row) stays a small
    // part of the matrix.  Matrices of short rows and short columns only (no row over kShortLen nonzeros, no column
    // over 128 rows): changed <= n m / (8 nnz) — measured on the 1M x 1M mixed instance: an event-driven round costs
    // 13 us + 3.7 us per 1,000 changed columns against 63-75 us for the full scan, so it wins below ~13,000 changed
    // columns.  Otherwise n m / (64 nnz): a medium row is a warp's job of its own in such a round, and a long column is
    // walked by one warp, 32 rows at a time.  CBP_EVENT_DIV replaces the 8 / 64 (0: never event-driven).
    int maxColLen = 0;
    for (int j = 0; j < n; ++j)
      maxColLen = std::max(maxColLen, hCscStart[static_cast<size_t>(j) + 1] - hCscStart[j]);
    const bool lightweight = mat->mShort == static_cast<int>(mKept) && maxColLen <= 128;
    static const double evDivEnv = getenv("CBP_EVENT_DIV") ? atof(getenv("CBP_EVENT_DIV")) : -1.0;
    const double evDiv = evDivEnv >= 0.0 ? evDivEnv : (lightweight ? 8.0 : 64.0);
    mat->eventMaxChanged = (evDiv > 0.0 && nnzKept > 0)
      ? static_cast<int>(std::min<double>(static_cast<double>(n) * static_cast<double>(mKept) / (evDiv * static_cast<double>(nnzKept)), 1.0e9))
      : 0;
  }
  std::vector<int> hShortBt, hLongBt;
  for (int r = 0; r < static_cast<int>(mKept); ++r)
    (hCnt[r] > kBtLongLen ? hLongBt : hShortBt).push_back(r);
  mat->nShortBt = static_cast<int>(hShortBt.size());
  mat->nLongBt = static_cast<int>(hLongBt.size());
  up(&mat->shortRowsBt, hShortBt);
  up(&mat->longRowsBt, hLongBt);
  ok = ok && mat_alloc(mat, &mat->longPart, nParts) == cudaSuccess;
  ok = ok && mat_alloc(mat, &mat->X[0], n) == cudaSuccess && mat_alloc(mat, &mat->X[1], n) == cudaSuccess;
  ok = ok && mat_alloc(mat, &mat->dirty[0], mat->nWords) == cudaSuccess
    && mat_alloc(mat, &mat->dirty[1], mat->nWords) == cudaSuccess;
  ok = ok && mat_alloc(mat, &mat->st, 1) == cudaSuccess;
  ok = ok && mat_alloc(mat, &mat->dLb, n) == cudaSuccess && mat_alloc(mat, &mat->dUb, n) == cudaSuccess;
  ok = ok && mat_alloc(mat, &mat->dRes, 16) == cudaSuccess;
  ok = ok && mat_alloc(mat, &mat->dAct, 3 * static_cast<size_t>(std::max(m, 1))) == cudaSuccess;
  ok = ok && cudaStreamSynchronize(ctx->stream) == cudaSuccess;
  lap("state buffers + sync");
  if (!ok) {
    ctx->err = std::string("cbp_matrix_upload: ") + cudaGetErrorString(cudaGetLastError());
    cbp_matrix_free(mat);
    return nullptr;
  }
  return mat;
}

// ----------------------------------------------------------------------------
// single vector
// ----------------------------------------------------------------------------
static int sv_stage_in(cbp_matrix *mat, const double *lb, const double *ub, int flags,
  const double **dl, const double **du)
{
  cbp_ctx *ctx = mat->ctx;
  if (flags & CBP_FLAG_DEVICE_PTRS) {
    *dl = lb;
    *du = ub;
    return CBP_OK;
  }
  CK(cudaMemcpyAsync(mat->dLb, lb, sizeof(double) * mat->n, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(mat->dUb, ub, sizeof(double) * mat->n, cudaMemcpyHostToDevice, ctx->stream));
  *dl = mat->dLb;
  *du = mat->dUb;
  return CBP_OK;
}

static int sv_stage_out(cbp_matrix *mat, const Bnd *X, double *lb, double *ub, int flags)
{
  cbp_ctx *ctx = mat->ctx;
  const int n = mat->n;
  const bool dev = flags & CBP_FLAG_DEVICE_PTRS;
  sv_finish_kernel<<<cdiv(n, 256), 256, 0, ctx->stream>>>(X, n, dev ? lb : mat->dLb, dev ? ub : mat->dUb);
  ++ctx->stats.kernelLaunches;
  CK(cudaGetLastError());
  if (!dev) {
    CK(cudaMemcpyAsync(lb, mat->dLb, sizeof(double) * n, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(ub, mat->dUb, sizeof(double) * n, cudaMemcpyDeviceToHost, ctx->stream));
  }
  return CBP_OK;
}

// where one run keeps its state: the matrix's own buffers, or its block of a row-sharded group
struct SvState {
  Bnd *X[2];
  unsigned *dirty[2];
  SvStatus *st;
  Bnd *inP[2];  // row-sharded runs only: proposal slots [world] and their marks, per round parity
  unsigned *inD[2];
};

static SvState sv_own_state(cbp_matrix *mat)
{
  SvState v;
  v.X[0] = mat->X[0];
  v.X[1] = mat->X[1];
  v.dirty[0] = mat->dirty[0];
  v.dirty[1] = mat->dirty[1];
  v.st = mat->st;
  v.inP[0] = v.inP[1] = nullptr;
  v.inD[0] = v.inD[1] = nullptr;
  return v;
}

static SvState sv_block_state(const cbp_matrix *mat, unsigned char *base)
{
  SvState v;
  v.X[0] = reinterpret_cast<Bnd *>(base);
  v.X[1] = reinterpret_cast<Bnd *>(base + mat->p2p.offX1);
  v.dirty[0] = reinterpret_cast<unsigned *>(base + mat->p2p.offD0);
  v.dirty[1] = reinterpret_cast<unsigned *>(base + mat->p2p.offD1);
  v.st = reinterpret_cast<SvStatus *>(base + mat->p2p.offSt);
  const cbp_matrix::P2P &q = mat->p2p;
  v.inP[0] = reinterpret_cast<Bnd *>(base + q.offP);
  v.inP[1] = v.inP[0] + q.strideP * q.world;
  v.inD[0] = reinterpret_cast<unsigned *>(base + q.offM);
  v.inD[1] = v.inD[0] + q.strideD * q.world;
  return v;
}

static int sv_reset_state(cbp_matrix *mat, const SvState &v)
{
  cbp_ctx *ctx = mat->ctx;
  SvStatus init;
  memset(&init, 0, sizeof(init));
  init.infeasRow = INT_MAX;
  init.infeasCol = INT_MAX;
  for (int r = 0; r < kMaxPeers; ++r)
    init.inRowFlag[0][r] = init.inRowFlag[1][r] = INT_MAX;
  // only the head of the status block needs initial values; the arrival counter at its end lives on
  CK(cudaMemsetAsync(v.st, 0, offsetof(SvStatus, arrive), ctx->stream));
  CK(cudaMemcpyAsync(v.st, &init, offsetof(SvStatus, roundChanged), cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemsetAsync(v.dirty[0], 0, sizeof(unsigned) * mat->nWords, ctx->stream));
  CK(cudaMemsetAsync(v.dirty[1], 0, sizeof(unsigned) * mat->nWords, ctx->stream));
  return CBP_OK;
}

static SvParams sv_params(cbp_matrix *mat, const SvState &v, const Scal &s, int maxRounds, double timeLimitSec,
  int raw)
{
  SvParams p;
  p.raw = raw;
  p.val = mat->val;
  p.col = mat->col;
  p.tiles = mat->tiles;
  p.rp = mat->rp;
  p.rowLB = mat->rowLB;
  p.rowUB = mat->rowUB;
  p.rowOrig = mat->rowOrig;
  p.longRows = mat->longRows;
  p.colType = mat->colType;
  p.longPart = mat->longPart;
  p.nTiles = mat->nTiles;
  p.firstLongTile = mat->firstLongTile;
  p.nLong = mat->nLong;
  p.n = mat->n;
  p.sellVal = mat->sellVal;
  p.sellCol = mat->sellCol;
  p.sellOff = mat->sellOff;
  p.sliceOrder = mat->sliceOrder;
  p.nSlices = mat->nSlices;
  p.mShort = mat->mShort;
  p.cscStart = mat->cscStart;
  p.cscRow = mat->cscRow;
  p.rowNz = mat->rowNz;
  p.rowCnt = mat->rowCnt;
  p.firstLongRow = mat->firstLongRow;
  p.eventMaxChanged = mat->eventMaxChanged;
  p.X[0] = v.X[0];
  p.X[1] = v.X[1];
  p.dirty[0] = v.dirty[0];
  p.dirty[1] = v.dirty[1];
  p.nWords = mat->nWords;
  p.st = v.st;
  p.inP[0] = v.inP[0];
  p.inP[1] = v.inP[1];
  p.inD[0] = v.inD[0];
  p.inD[1] = v.inD[1];
  p.strideP = mat->p2p.strideP;
  p.strideD = mat->p2p.strideD;
  for (int k = 0; k < 2; ++k) {
    p.myP[k] = v.inP[k] ? v.inP[k] + mat->p2p.rank * mat->p2p.strideP : nullptr;
    p.myD[k] = v.inD[k] ? v.inD[k] + mat->p2p.rank * mat->p2p.strideD : nullptr;
    p.myRowFlag[k] = &v.st->inRowFlag[k][mat->p2p.rank];
  }
  p.s = s;
  p.maxRounds = maxRounds;
  p.budgetNs = timeLimitSec >= 1.0e9 ? ~0ull : static_cast<unsigned long long>(timeLimitSec * 1.0e9);
  p.world = 0;
  p.rank = 0;
  p.rowOffset = 0;
  p.epoch0 = 0;
  return p;
}

// launch the persistent loop, wait, fill `out`, write the bounds back
static int sv_launch_and_collect(cbp_matrix *mat, const SvParams &pIn, const SvState &v, bool p2p, double *lb,
  double *ub, cbp_result *out, int *roundChanged, int roundCap, int flags)
{
  cbp_ctx *ctx = mat->ctx;
  SvParams p = pIn;
  const bool tiles = mat->nTiles > 0;
  const int maxGrid = ctx->svBlocksPerSM[tiles ? 1 : 0][p2p ? 1 : 0] * ctx->numSMs;
  const int want = std::max(std::max(mat->nTiles, cdiv(mat->nSlices, kWarps)), cdiv(mat->nWords, kThreads));
  const int grid = std::max(1, std::min(maxGrid, want));
  void *args[] = { &p };
  while (ctx->evPool.size() < 2) {
    cudaEvent_t e;
    CK(cudaEventCreate(&e));
    ctx->evPool.push_back(e);
  }
  cudaEvent_t k0 = ctx->evPool[0], k1 = ctx->evPool[1];
  CK(cudaEventRecord(k0, ctx->stream));
  mat->p2p.lastRunHere = p2p;
  void *fn = tiles ? (p2p ? reinterpret_cast<void *>(sv_fixpoint_kernel<true, true>)
                          : reinterpret_cast<void *>(sv_fixpoint_kernel<true, false>))
                   : (p2p ? reinterpret_cast<void *>(sv_fixpoint_kernel<false, true>)
                          : reinterpret_cast<void *>(sv_fixpoint_kernel<false, false>));
  CK(cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(kThreads), args, tiles ? sizeof(SvSmem) : kSvSmallSmem,
    ctx->stream));
  CK(cudaEventRecord(k1, ctx->stream));
  ++ctx->stats.kernelLaunches;

  struct Head {
    SvStatus s;
  };
  SvStatus head;
  CK(cudaMemcpyAsync(&head, v.st, offsetof(SvStatus, roundChanged), cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  if (p2p) {
    if (head.p2pError) {
      mat->p2p.broken = true;
      ctx->err = "row-sharded run: a rank did not reach the barrier in time";
      return CBP_ERR_CUDA;
    }
    mat->p2p.epoch = head.epochEnd;
  }
  out->rounds = head.rounds;
  out->stopReason = head.stopReason;
  out->nChanged = head.nChangedTotal;
  if (head.stopReason == CBP_STOP_INFEASIBLE) {
    out->feasible = 0;
    if (head.infeasRow != INT_MAX)
      out->infeasibleRow = head.infeasRow;
    else if (head.infeasCol != INT_MAX)
      out->infeasibleCol = head.infeasCol;
  }
  int rc = sv_stage_out(mat, v.X[head.finalBuf & 1], lb, ub, flags);
  if (rc)
    return rc;
  if (roundChanged && roundCap > 0) {
    const int done = head.rounds - (head.stopReason == CBP_STOP_INFEASIBLE ? 1 : 0);
    const int cnt = std::min(std::min(roundCap, done), kRoundCap);
    if (cnt > 0)
      CK(cudaMemcpyAsync(roundChanged, v.st->roundChanged, sizeof(int) * cnt, cudaMemcpyDeviceToHost,
        ctx->stream));
  }
  CK(cudaEventRecord(ctx->evB, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  float ms = 0.f, kms = 0.f;
  cudaEventElapsedTime(&ms, ctx->evA, ctx->evB);
  cudaEventElapsedTime(&kms, k0, k1);
  out->seconds = ms * 1.0e-3;
  ctx->stats.lastRunMs = ms;
  ctx->stats.lastRoundKernelMs = kms;
  ctx->stats.lastRoundKernelLaunches = 1;
  ctx->stats.lastRoundsTotal = head.rounds;
  return CBP_OK;
}

int cbp_singletons(cbp_matrix *mat, double *lb, double *ub, double primalTol, int *nTightened,
  int *nFixed, int flags)
{
  if (!mat || !lb || !ub || !nTightened)
    return CBP_ERR_ARG;
  cbp_ctx *ctx = mat->ctx;
  CK(cudaSetDevice(ctx->device));
  const Scal s = make_scal(mat, primalTol);
  const int n = mat->n;
  const double *dl, *du;
  int rc = sv_stage_in(mat, lb, ub, flags, &dl, &du);
  if (rc)
    return rc;
  if ((rc = sv_reset_state(mat, sv_own_state(mat))))
    return rc;
  CK(cudaMemsetAsync(mat->dRes, 0, sizeof(int) * 4, ctx->stream));
  sv_prepare_kernel<<<cdiv(n, 256), 256, 0, ctx->stream>>>(dl, du, mat->colType, n, mat->X[0], mat->X[1],
    mat->dirty[0], mat->st, s, 0);
  ++ctx->stats.kernelLaunches;
  if (mat->nSing > 0) {
    sv_singleton_kernel<<<cdiv(mat->nSing, 256), 256, 0, ctx->stream>>>(mat->singRows, mat->nSing,
      mat->rowNz, mat->val, mat->col, mat->rowLB, mat->rowUB, mat->colType, mat->X[0], mat->X[1],
      mat->dirty[0], s);
    sv_singleton_settle_kernel<<<cdiv(mat->nWords, 256), 256, 0, ctx->stream>>>(mat->X[0], mat->X[1],
      mat->dirty[0], mat->nWords, mat->dRes, s);
    ctx->stats.kernelLaunches += 2;
  }
  CK(cudaGetLastError());
  int res[4] = { 0, 0, 0, 0 };
  CK(cudaMemcpyAsync(res, mat->dRes, sizeof(int) * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  if (res[2] > 0) {  // bounds cross: infeasible, bounds left untouched
    *nTightened = -1;
    if (nFixed)
      *nFixed = 0;
    return CBP_OK;
  }
  *nTightened = res[0];
  if (nFixed)
    *nFixed = res[1];
  if (res[0] > 0) {
    if ((rc = sv_stage_out(mat, mat->X[1], lb, ub, flags)))
      return rc;
    CK(cudaStreamSynchronize(ctx->stream));
  }
  return CBP_OK;
}

int cbp_propagate(cbp_matrix *mat, double *lb, double *ub, double primalTol, int maxRounds,
  double timeLimitSec, cbp_result *out, int *roundChanged, int roundCap, int flags)
{
  if (!mat || !lb || !ub || !out)
    return CBP_ERR_ARG;
  cbp_ctx *ctx = mat->ctx;
  CK(cudaSetDevice(ctx->device));
  memset(out, 0, sizeof(*out));
  out->feasible = 1;
  out->infeasibleRow = -1;
  out->infeasibleCol = -1;
  if (!(timeLimitSec > 0.0)) {  // the reference tests the clock before every round (:109-121)
    out->stopReason = CBP_STOP_TIME_LIMIT;
    return CBP_OK;
  }
  const Scal s = make_scal(mat, primalTol);
  const int n = mat->n;
  const double *dl, *du;
  int rc = sv_stage_in(mat, lb, ub, flags, &dl, &du);
  if (rc)
    return rc;
  CK(cudaEventRecord(ctx->evA, ctx->stream));
  const SvState v = sv_own_state(mat);
  if ((rc = sv_reset_state(mat, v)))
    return rc;
  const int raw = (flags & CBP_FLAG_RAW_ROUND) ? 1 : 0;
  sv_prepare_kernel<<<cdiv(n, 256), 256, 0, ctx->stream>>>(dl, du, mat->colType, n, v.X[0], v.X[1], v.dirty[0], v.st,
    s, raw ? 0 : 1);
  ++ctx->stats.kernelLaunches;
  CK(cudaGetLastError());
  const SvParams p = sv_params(mat, v, s, maxRounds, timeLimitSec, raw);
  return sv_launch_and_collect(mat, p, v, false, lb, ub, out, roundChanged, roundCap, flags);
}

// ----------------------------------------------------------------------------
// single vector, rows sharded over the GPUs of one box (peer memory)
// ----------------------------------------------------------------------------
int cbp_p2p_export(cbp_matrix *mat, int world, void *handleOut)
{
  if (!mat || !handleOut || world < 1 || world > kMaxPeers)
    return CBP_ERR_ARG;
  cbp_ctx *ctx = mat->ctx;
  CK(cudaSetDevice(ctx->device));
  cbp_matrix::P2P &q = mat->p2p;
  if (q.block && q.world != world) {
    ctx->err = "cbp_p2p_export: this matrix already belongs to a group of another size";
    return CBP_ERR_ARG;
  }
  if (!q.block) {
    auto up256 = [](size_t x) { return (x + 255) & ~static_cast<size_t>(255); };
    q.world = world;
    q.strideP = up256(sizeof(Bnd) * mat->n) / sizeof(Bnd);
    q.strideD = up256(sizeof(unsigned) * mat->nWords) / sizeof(unsigned);
    q.offX1 = up256(sizeof(Bnd) * mat->n);
    q.offD0 = q.offX1 + up256(sizeof(Bnd) * mat->n);
    q.offD1 = q.offD0 + up256(sizeof(unsigned) * mat->nWords);
    q.offP = q.offD1 + up256(sizeof(unsigned) * mat->nWords);
    q.offM = q.offP + 2 * world * q.strideP * sizeof(Bnd);
    q.offSt = q.offM + 2 * world * q.strideD * sizeof(unsigned);
    q.bytes = q.offSt + up256(sizeof(SvStatus));
    CK(cudaMalloc(reinterpret_cast<void **>(&q.block), q.bytes));
    CK(cudaMemset(q.block, 0, q.bytes));
  }
  static_assert(sizeof(cudaIpcMemHandle_t) == CBP_P2P_HANDLE_BYTES, "handle size");
  cudaIpcMemHandle_t h;
  CK(cudaIpcGetMemHandle(&h, q.block));
  memcpy(handleOut, &h, sizeof(h));
  return CBP_OK;
}

int cbp_p2p_attach(cbp_matrix *mat, int rank, int world, const void *handles)
{
  if (!mat || !handles || world != mat->p2p.world || rank < 0 || rank >= world || !mat->p2p.block)
    return CBP_ERR_ARG;
  cbp_ctx *ctx = mat->ctx;
  CK(cudaSetDevice(ctx->device));
  cbp_matrix::P2P &q = mat->p2p;
  q.rank = rank;
  q.attached = true;
  for (int r = 0; r < world; ++r) {
    if (r == rank) {
      q.peer[r] = q.block;
      continue;
    }
    cudaIpcMemHandle_t h;
    memcpy(&h, static_cast<const unsigned char *>(handles) + static_cast<size_t>(r) * sizeof(h), sizeof(h));
    void *ptr = nullptr;
    CK(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
    q.peer[r] = static_cast<unsigned char *>(ptr);
  }
  return CBP_OK;
}

int cbp_p2p_propagate(cbp_matrix *mat, double *lb, double *ub, double primalTol, int maxRounds, double timeLimitSec,
  int rowOffset, cbp_result *out, int *roundChanged, int roundCap, int flags)
{
  if (!mat || !lb || !ub || !out || !mat->p2p.attached)
    return CBP_ERR_ARG;
  cbp_ctx *ctx = mat->ctx;
  cbp_matrix::P2P &q = mat->p2p;
  if (q.broken) {
    ctx->err = "row-sharded group unusable after a barrier timeout";
    return CBP_ERR_ARG;
  }
  CK(cudaSetDevice(ctx->device));
  memset(out, 0, sizeof(*out));
  out->feasible = 1;
  out->infeasibleRow = -1;
  out->infeasibleCol = -1;
  const Scal s = make_scal(mat, primalTol);
  const double *dl, *du;
  int rc = sv_stage_in(mat, lb, ub, flags, &dl, &du);
  if (rc)
    return rc;
  CK(cudaEventRecord(ctx->evA, ctx->stream));
  const SvState v = sv_block_state(mat, q.block);
  if ((rc = sv_reset_state(mat, v)))
    return rc;
  // the peer tables: every rank's block as this rank sees it
  SvStatus tab;
  memset(&tab, 0, sizeof(tab));
  for (int r = 0; r < q.world; ++r) {
    const SvState pv = sv_block_state(mat, q.peer[r]);
    for (int k = 0; k < 2; ++k) {  // this rank's slot over there
      tab.pushP[k][r] = pv.inP[k] + q.rank * q.strideP;
      tab.pushD[k][r] = pv.inD[k] + q.rank * q.strideD;
    }
    tab.peerSt[r] = pv.st;
  }
  tab.inP[0] = v.inP[0];
  tab.inP[1] = v.inP[1];
  tab.inD[0] = v.inD[0];
  tab.inD[1] = v.inD[1];
  tab.strideP = q.strideP;
  tab.strideD = q.strideD;
  tab.world = q.world;
  tab.rank = q.rank;
  tab.nWords = mat->nWords;
  tab.scal = s;
  const size_t tabOff = offsetof(SvStatus, pushP), tabEnd = offsetof(SvStatus, stamps);
  CK(cudaMemcpyAsync(reinterpret_cast<unsigned char *>(v.st) + tabOff, reinterpret_cast<unsigned char *>(&tab) + tabOff,
    tabEnd - tabOff, cudaMemcpyHostToDevice, ctx->stream));
  // no proposals in this rank's own slots, no marks in any slot; what R11 changes is marked as a
  // (valueless) proposal of round 0 so that every rank counts it there
  for (int k = 0; k < 2; ++k)
    sv_no_proposals_kernel<<<cdiv(mat->n, 256), 256, 0, ctx->stream>>>(v.inP[k] + q.rank * q.strideP, mat->n);
  CK(cudaMemsetAsync(v.inD[0], 0, q.offSt - q.offM, ctx->stream));
  sv_prepare_kernel<<<cdiv(mat->n, 256), 256, 0, ctx->stream>>>(dl, du, mat->colType, mat->n, v.X[0], v.X[1],
    v.inD[0] + q.rank * q.strideD, v.st, s, 1);
  ctx->stats.kernelLaunches += 3;
  CK(cudaGetLastError());
  SvParams p = sv_params(mat, v, s, maxRounds, timeLimitSec, 0);
  p.world = q.world;
  p.rank = q.rank;
  p.rowOffset = rowOffset;
  p.epoch0 = q.epoch;
  return sv_launch_and_collect(mat, p, v, true, lb, ub, out, roundChanged, roundCap, flags);
}

int cbp_sv_round_stamps(cbp_matrix *mat, unsigned *out, int maxRounds)
{
  if (!mat || !out || maxRounds <= 0)
    return CBP_ERR_ARG;
  cbp_ctx *ctx = mat->ctx;
  CK(cudaSetDevice(ctx->device));
  const int cnt = std::min(maxRounds, kStampRounds);
  const SvStatus *st = mat->p2p.lastRunHere ? sv_block_state(mat, mat->p2p.block).st : mat->st;
  CK(cudaMemcpy(out, reinterpret_cast<const char *>(st) + offsetof(SvStatus, stamps), sizeof(unsigned) * 4 * cnt,
    cudaMemcpyDeviceToHost));
  if (getenv("CBP_STAMPS_X"))  // tuning aid: CTA 0's own A1 / rows timestamps instead of columns 1 and 3
    for (int r = 0; r < cnt; ++r) {
      unsigned x[4];
      CK(cudaMemcpy(x, reinterpret_cast<const char *>(st) + offsetof(SvStatus, stampsX) + 16 * r, 16, cudaMemcpyDeviceToHost));
      out[4 * r + 1] = x[0];
      out[4 * r + 3] = x[1];
      if (getenv("CBP_STAMPS_TOP"))
        out[4 * r + 0] = x[2];
    }
  return cnt;
}

int cbp_row_activity(cbp_matrix *mat, const double *lb, const double *ub, double *minAct,
  double *maxAct, int *infMin, int *infMax, int flags)
{
  if (!mat || !lb || !ub || !minAct || !maxAct || !infMin || !infMax)
    return CBP_ERR_ARG;
  cbp_ctx *ctx = mat->ctx;
  CK(cudaSetDevice(ctx->device));
  const double *dl, *du;
  int rc = sv_stage_in(mat, lb, ub, flags, &dl, &du);
  if (rc)
    return rc;
  const int m = mat->m;
  const bool dev = flags & CBP_FLAG_DEVICE_PTRS;
  double *dMin = dev ? minAct : mat->dAct, *dMax = dev ? maxAct : mat->dAct + m;
  int *dIm = dev ? infMin : reinterpret_cast<int *>(mat->dAct + 2 * static_cast<size_t>(m));
  int *dIx = dev ? infMax : dIm + m;
  CK(cudaMemsetAsync(dMin, 0, sizeof(double) * m, ctx->stream));
  CK(cudaMemsetAsync(dMax, 0, sizeof(double) * m, ctx->stream));
  CK(cudaMemsetAsync(dIm, 0, sizeof(int) * m, ctx->stream));
  CK(cudaMemsetAsync(dIx, 0, sizeof(int) * m, ctx->stream));
  if (mat->mKept > 0) {
    row_activity_kernel<<<cdiv(static_cast<long long>(mat->mKept) * 32, 256), 256, 0, ctx->stream>>>(
      mat->rowNz, mat->rowCnt, mat->rowOrig, mat->mKept, mat->val, mat->col, dl, du, mat->inf, dMin, dMax,
      dIm, dIx);
    ++ctx->stats.kernelLaunches;
  }
  CK(cudaGetLastError());
  if (!dev) {
    CK(cudaMemcpyAsync(minAct, dMin, sizeof(double) * m, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(maxAct, dMax, sizeof(double) * m, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(infMin, dIm, sizeof(int) * m, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(infMax, dIx, sizeof(int) * m, cudaMemcpyDeviceToHost, ctx->stream));
  }
  CK(cudaStreamSynchronize(ctx->stream));
  return CBP_OK;
}

// ----------------------------------------------------------------------------
// batch
// ----------------------------------------------------------------------------
static int bt_reserve(cbp_matrix *mat, int B)
{
  cbp_ctx *ctx = mat->ctx;
  BatchState &bt = mat->bt;
  const int Bp = (B + 31) / 32 * 32;
  const size_t cells = static_cast<size_t>(mat->n) * Bp;
  if (cells > bt.capCells || Bp > bt.capBp) {
    bt_free_lists(bt, true, true);
    pool_give(ctx, bt.arena, bt.arenaBytes, false);
    pool_give(ctx, bt.hBlob, bt.hBlobBytes, true);
    bt.arena = nullptr;
    bt.blob = nullptr;
    bt.hBlob = nullptr;
    bt.wlWalk = nullptr;
    bt.wlUsed = nullptr;
    bt.wlTableClean = false;
    bt.wlHKey = nullptr;
    bt.dmask = nullptr;
    bt.wlItemsS[0] = bt.wlItemsS[1] = bt.wlItemsL[0] = bt.wlItemsL[1] = nullptr;
    bt.wlK = bt.wlCells = nullptr;
    bt.wlCtl = nullptr;
    bt.wlInts = nullptr;
    bt.wlMarksKnown = false;
    bt.wlPrevCells = 0;
    bt.X[0] = bt.X[1] = nullptr;
    bt.dirty[0] = bt.dirty[1] = bt.ever = bt.active = nullptr;
    bt.ints = nullptr;
    bt.counts = nullptr;
    bt.chunkCnt = nullptr;
    bt.log = nullptr;
    bt.groupFlag = nullptr;
    bt.capCells = 0;
    bt.capBp = 0;
    const size_t words = cells / 32;
    // list sizes follow the batch: candidates / changed cells / row items per round
    const unsigned logCap = static_cast<unsigned>(std::min<size_t>(kBtLogCap, std::max<size_t>(1u << 20, static_cast<size_t>(Bp) * 2048)));
    bt.wlItemCap = static_cast<unsigned>(std::min<size_t>(1u << 25, std::max<size_t>(1u << 20, static_cast<size_t>(Bp) * 4096)));
    bt.wlCellCap = static_cast<unsigned>(std::min<size_t>(1u << 25, std::max<size_t>(1u << 20, static_cast<size_t>(Bp) * 1024)));
    if (getenv("CBP_WL_ITEM_CAP"))  // tests shrink the item lists to exercise the hand-over to the full-walk path
      bt.wlItemCap = static_cast<unsigned>(std::max<long>(1, atol(getenv("CBP_WL_ITEM_CAP"))));
    // one table slot per (vector, long row) pair that comes up; matrices without rows longer than kShortLen need none
    bt.wlHSlots = 2;
    if (mat->mKept > 0) {
      // (every row's pairs go through the table when sums are exact; only the longer rows' otherwise)
      // (a run empties the slots it took, so the size costs memory, not time: sized for a load factor well under 1/2)
      size_t want = std::max<size_t>(1u << 18, static_cast<size_t>(Bp) * (mat->allInt ? 2048 : (mat->mKept > mat->mShort ? 256 : 0)));
      bt.wlHSlots = 1u << 18;
      while (bt.wlHSlots < want && bt.wlHSlots < (1u << 25))
        bt.wlHSlots <<= 1;
    }
    if (getenv("CBP_WL_SLOTS")) {  // tests shrink the pair table to exercise the hand-over when it fills up
      bt.wlHSlots = 2;
      while (bt.wlHSlots < static_cast<unsigned>(std::max<long>(2, atol(getenv("CBP_WL_SLOTS")))))
        bt.wlHSlots <<= 1;
    }
    bt.wlHActive = bt.wlHSlots;
    bt.wlUsedCap = std::min(bt.wlHSlots, std::max(1u << 20, bt.wlItemCap));
    bt.wlTableClean = false;
    bt.wlWalkCap = mat->nLongBt > 0
      ? static_cast<unsigned>(std::min<size_t>(1u << 22, std::max<size_t>(1u << 16, static_cast<size_t>(Bp) * 256))) : 1u;
    {
      auto up256 = [](size_t x) { return (x + 255) & ~static_cast<size_t>(255); };
      bt.blobOffInts = up256(sizeof(WlCtl));
      bt.blobOffCounts = bt.blobOffInts + up256(sizeof(int) * (static_cast<size_t>(Bp) * 7 + 8));
      bt.blobOffHead = bt.blobOffCounts + up256(sizeof(long long) * (static_cast<size_t>(Bp) + 1));
      bt.blobBytes = bt.blobOffHead + sizeof(WlUpdHead) * static_cast<size_t>(kUpdPrefetch);
      // one device allocation for the whole workspace (a first batch on a new matrix paid some twenty cudaMallocs)
      struct Piece {
        void **ptr;
        size_t bytes;
      };
      auto piece = [](auto **p, size_t count) {
        using T = typename std::remove_pointer<typename std::remove_pointer<decltype(p)>::type>::type;
        return Piece { reinterpret_cast<void **>(p), std::max<size_t>(count, 1) * sizeof(T) };
      };
      const Piece pieces[] = { piece(&bt.wlHKey, bt.wlHSlots), piece(&bt.wlUsed, bt.wlUsedCap), piece(&bt.wlWalk, bt.wlWalkCap),
        piece(&bt.wlItemsS[0], bt.wlItemCap), piece(&bt.wlItemsS[1], bt.wlItemCap), piece(&bt.wlItemsL[0], bt.wlItemCap),
        piece(&bt.wlItemsL[1], bt.wlItemCap), piece(&bt.wlK, bt.wlCellCap), piece(&bt.wlCells, bt.wlCellCap),
        piece(&bt.wlInts, static_cast<size_t>(Bp) * 4 + static_cast<size_t>(Bp / 32) * 2 + 8), piece(&bt.X[0], cells),
        piece(&bt.X[1], cells), piece(&bt.dirty[0], words), piece(&bt.dirty[1], words), piece(&bt.ever, words),
        piece(&bt.active, Bp / 32), piece(&bt.dmask, words),
        piece(&bt.chunkCnt, static_cast<size_t>(cdiv(mat->n, kUpdChunk)) * Bp), piece(&bt.groupFlag, static_cast<size_t>(Bp / 32) + 2),
        piece(&bt.log, logCap), piece(&bt.blob, bt.blobBytes), piece(&bt.wlRowBase, static_cast<size_t>(std::max(mat->mKept, 1))),
        piece(&bt.baseLb, mat->n), piece(&bt.baseUb, mat->n), piece(&bt.baseX, mat->n), piece(&bt.stCol, kListDefault),
        piece(&bt.stLb, kListDefault), piece(&bt.stUb, kListDefault), piece(&bt.st2Col, kListDefault),
        piece(&bt.st2Lb, kListDefault), piece(&bt.st2Ub, kListDefault) };
      size_t total = 0;
      for (const Piece &q : pieces)
        total += up256(q.bytes);
      bt.arenaBytes = total;
      bt.arena = static_cast<unsigned char *>(pool_take(ctx, total, false, &bt.arenaBytes));
      bt.hBlobBytes = bt.blobBytes;
      bt.hBlob = static_cast<unsigned char *>(pool_take(ctx, bt.blobBytes, true, &bt.hBlobBytes));
      if (!bt.arena && cudaMalloc(reinterpret_cast<void **>(&bt.arena), total) != cudaSuccess) {
        bt.arena = nullptr;
        pool_flush(ctx);
      }
      if ((!bt.arena && cudaMalloc(reinterpret_cast<void **>(&bt.arena), total) != cudaSuccess)
        || (!bt.hBlob && cudaHostAlloc(reinterpret_cast<void **>(&bt.hBlob), bt.blobBytes, cudaHostAllocDefault) != cudaSuccess)) {
        ctx->err = std::string("batch workspace: ") + cudaGetErrorString(cudaGetLastError());
        if (bt.arena)
          cudaFree(bt.arena);
        if (bt.hBlob)
          cudaFreeHost(bt.hBlob);
        bt.arena = nullptr;
        bt.hBlob = nullptr;
        return CBP_ERR_NOMEM;
      }
      size_t at = 0;
      for (const Piece &q : pieces) {
        *q.ptr = bt.arena + at;
        at += up256(q.bytes);
      }
      bt.stCap = bt.st2Cap = kListDefault;
      bt.stInArena = bt.st2InArena = true;
      bt.wlCtl = reinterpret_cast<WlCtl *>(bt.blob);
      bt.ints = reinterpret_cast<int *>(bt.blob + bt.blobOffInts);
      bt.counts = reinterpret_cast<long long *>(bt.blob + bt.blobOffCounts);
    }
    bt.capCells = cells;
    bt.capBp = Bp;
    bt.dirty0Clean = bt.dmaskClean = false;
    // CBP_LOG_CAP: tests shrink the list to exercise the fall-back to the two-buffer first round
    bt.logCap = getenv("CBP_LOG_CAP") ? std::min<unsigned>(logCap, static_cast<unsigned>(atol(getenv("CBP_LOG_CAP")))) : logCap;
  }
  bt.B = B;
  bt.Bp = Bp;
  bt.G = Bp / 32;
  bt.cells = cells;
  // slice width: the widest W whose slice (n x W x 16 bytes of bounds) stays well inside L2
  {
    static const double budget = getenv("CBP_L2_BUDGET_MB") ? atof(getenv("CBP_L2_BUDGET_MB")) * 1e6 : 1e15;  // measured so far: W = 32 wins (profiles/README.md)
    const double lineBytes = 16.0 * mat->n;
    bt.W = lineBytes * 32 <= budget ? 32 : (lineBytes * 16 <= budget ? 16 : (lineBytes * 8 <= budget ? 8 : 4));
  }
  return CBP_OK;
}

static BtParams bt_params(cbp_matrix *mat, double tol)
{
  BatchState &bt = mat->bt;
  BtParams p;
  p.val = mat->val;
  p.col = mat->col;
  p.rowNz = mat->rowNz;
  p.rowCnt = mat->rowCnt;
  p.rowPos = mat->rowPos;
  p.rowLB = mat->rowLB;
  p.rowUB = mat->rowUB;
  p.rowOrig = mat->rowOrig;
  p.colType = mat->colType;
  p.shortRowsBt = mat->shortRowsBt;
  p.longRowsBt = mat->longRowsBt;
  p.nShortBt = mat->nShortBt;
  p.nLongBt = mat->nLongBt;
  p.mKept = mat->mKept;
  p.n = mat->n;
  p.Bp = bt.Bp;
  p.G = bt.G;
  p.W = bt.W;
  p.X[0] = bt.X[0];
  p.X[1] = bt.X[1];
  p.dirty[0] = bt.dirty[0];
  p.dirty[1] = bt.dirty[1];
  p.ever = bt.ever;
  p.active = bt.active;
  int *q = bt.ints;
  p.vecInfRow = q;
  p.vecInfCol = q + bt.Bp;
  p.vecChanged = q + 2 * bt.Bp;
  p.vecRounds = q + 3 * bt.Bp;
  p.vecTotal = q + 4 * bt.Bp;
  p.vecStop = q + 5 * bt.Bp;
  p.vecFinal = q + 6 * bt.Bp;
  p.anyActive = q + 7 * bt.Bp;
  p.log = bt.log;
  p.logCap = bt.logCap;
  p.groupFlag = bt.groupFlag;
  p.baseX = bt.baseX;
  p.dmask = bt.dmask;
  p.logCount = reinterpret_cast<unsigned *>(bt.groupFlag + bt.G);
  p.s = make_scal(mat, tol);
  return p;
}

int cbp_batch_load_delta(cbp_matrix *mat, const double *baseLb, const double *baseUb, int B,
  const int64_t *deltaStart, const uint32_t *deltaVar, const double *deltaBound, int flags)
{
  if (!mat || !baseLb || !baseUb || B <= 0 || !deltaStart)
    return CBP_ERR_ARG;
  cbp_ctx *ctx = mat->ctx;
  CK(cudaSetDevice(ctx->device));
  int rc = bt_reserve(mat, B);
  if (rc)
    return rc;
  BatchState &bt = mat->bt;
  const cudaMemcpyKind kind = (flags & CBP_FLAG_DEVICE_PTRS) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
  long long nDelta = 0;
  if (flags & CBP_FLAG_DEVICE_PTRS) {
    CK(cudaMemcpyAsync(&nDelta, deltaStart + B, sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
  } else {
    nDelta = deltaStart[B];
  }
  if (nDelta < 0 || (nDelta > 0 && (!deltaVar || !deltaBound)))
    return CBP_ERR_ARG;
  // every diff must name a column of this matrix (bits 30 and 31 are flags, CbcPartialNodeInfo.cpp:139,208)
  if (!(flags & CBP_FLAG_DEVICE_PTRS)) {
    for (long long k = 0; k < nDelta; ++k)
      if (diff_col(deltaVar[k]) >= mat->n) {
        ctx->err = "cbp_batch_load_delta: a bound diff names a column outside the matrix";
        return CBP_ERR_ARG;
      }
  } else if (nDelta > 0) {
    int bad = 0;
    CK(cudaMemsetAsync(mat->dRes, 0, sizeof(int), ctx->stream));
    bt_delta_check_kernel<<<cdiv(nDelta, 256), 256, 0, ctx->stream>>>(deltaVar, nDelta, mat->n, mat->dRes);
    ++ctx->stats.kernelLaunches;
    CK(cudaMemcpyAsync(&bad, mat->dRes, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (bad) {
      ctx->err = "cbp_batch_load_delta: a bound diff names a column outside the matrix";
      return CBP_ERR_ARG;
    }
  }
  if (static_cast<size_t>(nDelta) > bt.capDelta || !bt.dStart) {
    void *old[] = { bt.dVar, bt.dBnd };
    for (void *p : old)
      if (p)
        cudaFree(p);
    bt.dVar = nullptr;
    bt.dBnd = nullptr;
    if (dalloc(&bt.dVar, nDelta) != cudaSuccess || dalloc(&bt.dBnd, nDelta) != cudaSuccess) {
      ctx->err = "batch workspace: deltas";
      return CBP_ERR_NOMEM;
    }
    bt.capDelta = static_cast<size_t>(nDelta);
  }
  if (!bt.dStart || static_cast<size_t>(B) + 1 > bt.capStart) {
    if (bt.dStart)
      cudaFree(bt.dStart);
    bt.dStart = nullptr;
    bt.capStart = 0;
    if (dalloc(&bt.dStart, static_cast<size_t>(B) + 1) != cudaSuccess) {
      ctx->err = "batch workspace: delta offsets";
      return CBP_ERR_NOMEM;
    }
    bt.capStart = static_cast<size_t>(B) + 1;
  }
  bt.checkedTol = -1.0;
  CK(cudaMemcpyAsync(bt.baseLb, baseLb, sizeof(double) * mat->n, kind, ctx->stream));
  CK(cudaMemcpyAsync(bt.baseUb, baseUb, sizeof(double) * mat->n, kind, ctx->stream));
  CK(cudaMemcpyAsync(bt.dStart, deltaStart, sizeof(long long) * (B + 1), kind, ctx->stream));
  if (nDelta > 0) {
    CK(cudaMemcpyAsync(bt.dVar, deltaVar, sizeof(unsigned) * nDelta, kind, ctx->stream));
    CK(cudaMemcpyAsync(bt.dBnd, deltaBound, sizeof(double) * nDelta, kind, ctx->stream));
  }
  bt.nDelta = nDelta;
  bt.loaded = true;
  bt.denseLoaded = false;
  bt.consumed = false;
  // Are all input bounds finite?  (Then the full-walk rows kernels drop their infinity tests.)  Found out on the device,
  // and only if that path is taken: the worklist kernel looks for itself.
  bt.hasInf = true;
  bt.hasInfKnown = false;
  bt.lastWasWorklist = false;
  return CBP_OK;
}

// the buffers a worklist run writes the vectors' update lists into (a fetch grows them when a run needs more)
static int bt_list_buffers(cbp_matrix *mat)
{
  BatchState &bt = mat->bt;
  if (bt.stCap > 0 && bt.st2Cap > 0)
    return CBP_OK;
  bt_free_lists(bt, true, true);
  const long long want = kListDefault;
  if (dalloc(&bt.stCol, want) != cudaSuccess || dalloc(&bt.stLb, want) != cudaSuccess || dalloc(&bt.stUb, want) != cudaSuccess
    || dalloc(&bt.st2Col, want) != cudaSuccess || dalloc(&bt.st2Lb, want) != cudaSuccess || dalloc(&bt.st2Ub, want) != cudaSuccess) {
    mat->ctx->err = "batch workspace (update lists)";
    return CBP_ERR_NOMEM;
  }
  bt.stCap = bt.st2Cap = want;
  return CBP_OK;
}

// A base + diffs batch in worklist form (cbp_wl_kernels.cuh): a few set-up launches, ONE cooperative launch for all
// rounds, one read-back of the control block.  *status: kWlOk, or why the caller must run the full-walk path instead.
static int bt_wl_run(cbp_matrix *mat, const BtParams &p, int limit, double timeLimitSec, int *status)
{
  cbp_ctx *ctx = mat->ctx;
  BatchState &bt = mat->bt;
  cudaStream_t st = ctx->stream;
  const size_t words = bt.cells / 32;
  if (!bt.wlMarksKnown) {  // first use, or the full-walk path ran in between: start from empty masks
    CK(cudaMemsetAsync(bt.dirty[0], 0, sizeof(unsigned) * words, st));
    CK(cudaMemsetAsync(bt.dirty[1], 0, sizeof(unsigned) * words, st));
    CK(cudaMemsetAsync(bt.ever, 0, sizeof(unsigned) * words, st));
    CK(cudaMemsetAsync(bt.dmask, 0, sizeof(unsigned) * words, st));
    bt.wlPrevCells = 0;
  }
  bt.wlMarksKnown = false;  // (until this run has ended in good order)
  bt.dmaskClean = bt.dirty0Clean = false;
  WlParams w;
  w.bt = p;
  w.cscStart = mat->cscStart;
  w.cscRow = mat->cscRow;
  w.cscVal = mat->cscVal;
  w.sellVal = mat->sellVal;
  w.sellCol = mat->sellCol;
  w.sellOff = mat->sellOff;
  w.nSlices = mat->nSlices;
  w.mShort = mat->mShort;
  w.tab = bt.wlHKey;
  w.hMask = bt.wlHActive - 1;
  w.hShift = 64;
  for (unsigned t = bt.wlHActive; t > 1u; t >>= 1)
    --w.hShift;
  if (w.hShift > 63) {  // (a one-slot table: no shift by 64)
    w.hShift = 63;
  }
  w.rowBase = bt.wlRowBase;
  w.exactOk = mat->allInt ? 1 : 0;  // (which rows' sums are exact — integer data, finite bounds — the kernel finds out)
  w.used = bt.wlUsed;
  w.usedCap = bt.wlUsedCap;
  if (!bt.wlTableClean) {  // first use, or the last run took more slots than its list holds
    CK(cudaMemsetAsync(bt.wlHKey, 0, static_cast<size_t>(bt.wlHSlots) * sizeof(WlPair), st));
    bt.wlPrevUsed = 0;
  }
  w.prevUsed = bt.wlPrevUsed;
  bt.wlTableClean = false;
  w.dStart = bt.dStart;
  w.dVar = bt.dVar;
  w.dBnd = bt.dBnd;
  w.B = bt.B;
  for (int k = 0; k < 2; ++k) {
    w.itemsS[k] = bt.wlItemsS[k];
    w.itemsL[k] = bt.wlItemsL[k];
    w.vecChanged2[k] = bt.wlInts + static_cast<size_t>(k) * bt.Bp;
    w.active2[k] = reinterpret_cast<unsigned *>(bt.wlInts + 2 * static_cast<size_t>(bt.Bp) + static_cast<size_t>(k) * bt.G);
  }
  w.K = bt.wlK;
  w.cells = bt.wlCells;
  w.walk = bt.wlWalk;
  w.walkCap = bt.wlWalkCap;
  w.itemCap = bt.wlItemCap;
  w.kCap = bt.wlCellCap;
  w.cellCap = bt.wlCellCap;
  w.prevCells = bt.wlPrevCells;
  w.ctl = bt.wlCtl;
  // the update lists (P7): unsorted into st2*, sorted into st*; buffers sized by an earlier fetch (none yet: the
  // first fetch takes the stand-alone path and sizes them)
  if (bt.listsWanted) {
    const int rcl = bt_list_buffers(mat);
    if (rcl)
      return rcl;
  }
  bt.updReady = false;
  w.updStart = bt.counts;
  w.updCnt = bt.wlInts + 2 * static_cast<size_t>(bt.Bp) + 2 * static_cast<size_t>(bt.G) + 8;  // (its own 2 Bp ints)
  w.updColTmp = bt.st2Col;
  w.updLbTmp = bt.st2Lb;
  w.updUbTmp = bt.st2Ub;
  w.updCol = bt.stCol;
  w.updLb = bt.stLb;
  w.updUb = bt.stUb;
  w.updCap = bt.listsWanted ? std::min(bt.stCap, bt.st2Cap) : 0;  // (a caller that never fetches lists does not pay for them)
  w.updHead = reinterpret_cast<WlUpdHead *>(bt.blob + bt.blobOffHead);
  w.updHeadCap = kUpdPrefetch;
  w.limit = limit;
  w.budgetNs = timeLimitSec >= 1.0e9 ? ~0ull : static_cast<unsigned long long>(timeLimitSec * 1.0e9);
  wl_setup_kernel<<<cdiv(std::max<size_t>(std::max(mat->n, bt.Bp), sizeof(WlCtl) / sizeof(unsigned)), 256), 256, 0, st>>>(w,
    bt.baseLb, bt.baseUb, bt.baseX);
  ++ctx->stats.kernelLaunches;
  const int grid = std::max(1, ctx->wlBlocksPerSM * ctx->numSMs);
  void *args[] = { &w };
  CK(cudaEventRecord(ctx->evPool[0], st));
  CK(cudaLaunchCooperativeKernel(reinterpret_cast<void *>(bt_worklist_kernel), dim3(grid), dim3(kThreads), args, 0, st));
  CK(cudaEventRecord(ctx->evPool[1], st));
  ++ctx->stats.kernelLaunches;
  CK(cudaEventRecord(ctx->evB, st));
  // the control block, and with it what a fetch reads first: the per-vector results, the lists' offsets and as much of
  // the lists' head as the last run would have needed (and a quarter more)
  const long long pre = w.updCap > 0 ? std::min<long long>(kUpdPrefetch, std::max<long long>(256, bt.updTotal + bt.updTotal / 4 + 64)) : 0;
  CK(cudaMemcpyAsync(bt.hBlob, bt.blob, bt.blobOffHead + sizeof(WlUpdHead) * static_cast<size_t>(pre), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));  // the one host synchronisation of the run
  const WlCtl &h = *reinterpret_cast<const WlCtl *>(bt.hBlob);
  bt.headCopied = pre;
  *status = h.status;
  bt.wlTableClean = !h.usedOverflow;
  bt.wlPrevUsed = std::min(h.nUsed, bt.wlUsedCap);
  {  // the next run's table: three slots per pair this one took (all of it after a run that ran out of room)
    unsigned want = 1u << 10;
    const unsigned f = static_cast<unsigned>(std::max(1, knobs().wlTableFactor));
    while (want < bt.wlHSlots && want / f < h.nUsed)
      want <<= 1;
    bt.wlHActive = (h.status == kWlOverflow || h.usedOverflow || knobs().wlTableFactor <= 0) ? bt.wlHSlots
                                                                                             : std::min(want, bt.wlHSlots);
  }
  if (knobs().wlStamps) {
    fprintf(stderr, "worklist run: status %d rounds %d items %llu pairs %llu cells %u table %u of %u; phase ends (us):", h.status,
      h.rounds, h.itemsTotal, h.pairs, h.nCells, h.nUsed, w.hMask + 1u);
    for (unsigned i = 0; i < std::min(h.nStamps, 96u); ++i)
      fprintf(stderr, " %.1f", h.stamps[i] * 1e-3);
    fprintf(stderr, "\n");
  }
  float kms = 0.f, ms = 0.f;
  cudaEventElapsedTime(&kms, ctx->evPool[0], ctx->evPool[1]);
  cudaEventElapsedTime(&ms, ctx->evA, ctx->evB);
  if (h.status == kWlBaseNotFixpoint) {
    bt.wlMarksKnown = true;  // (the kernel left before touching anything but the previous run's cells)
    bt.wlPrevCells = 0;
    return CBP_OK;
  }
  if (h.status != kWlOk)
    return CBP_OK;
  bt.wlMarksKnown = true;
  bt.wlPrevCells = std::min(h.nCells, bt.wlCellCap);
  bt.lastWasWorklist = true;
  bt.updReady = h.updDone != 0;
  bt.updTotal = static_cast<long long>(h.nUpd);
  bt.baseMode = true;
  bt.quiet = false;
  bt.groupBuilt.assign(bt.G, 0);
  ctx->stats.lastRunMs = ms;
  ctx->stats.lastRoundKernelMs = kms;
  ctx->stats.lastRoundKernelLaunches = std::max(1, h.rounds);
  ctx->stats.lastFromBase = 1;
  ctx->stats.lastGroupsBuilt = 0;
  ctx->stats.lastMode = 2;
  ctx->stats.lastPairsWalked = static_cast<int64_t>(h.pairs);
  ctx->stats.lastHostSyncs = 1;
  ctx->stats.lastItems = static_cast<int64_t>(h.itemsTotal);
  ctx->stats.lastRoundsTotal = static_cast<long long>(h.roundsTotal);  // rounds summed over the vectors
  return CBP_OK;
}

int cbp_batch_run(cbp_matrix *mat, double primalTol, int maxRounds, double timeLimitSec)
{
  if (!mat || !mat->bt.loaded)
    return CBP_ERR_ARG;
  cbp_ctx *ctx = mat->ctx;
  CK(cudaSetDevice(ctx->device));
  BatchState &bt = mat->bt;
  const BtParams p = bt_params(mat, primalTol);
  const int limit = maxRounds > 0 ? maxRounds : INT_MAX;
  const size_t words = bt.cells / 32;
  cudaStream_t st = ctx->stream;
  while (ctx->evPool.size() < 2) {
    cudaEvent_t e;
    CK(cudaEventCreate(&e));
    ctx->evPool.push_back(e);
  }
  CK(cudaEventRecord(ctx->evA, st));
  int hostSyncs = 0;
  // Base + diffs batches first try the worklist form: one persistent kernel for all rounds, only the (row, vector)
  // pairs that a changed column touches.  It hands back when the node bounds are not a propagation fixpoint (then
  // every row can act for every vector) or when a list overflows.  CBP_NO_WORKLIST=1: tests and A/B runs.
  if (!bt.denseLoaded && !bt.consumed && bt.W == 32 && !knobs().noWorklist && !(timeLimitSec <= 0.0)) {
    int wlStatus = kWlOk;
    const int rcw = bt_wl_run(mat, p, limit, timeLimitSec, &wlStatus);
    if (rcw)
      return rcw;
    ++hostSyncs;
    if (wlStatus == kWlOk)
      return CBP_OK;
    CK(cudaEventRecord(ctx->evA, st));
  }
  bt.wlMarksKnown = false;  // the full-walk path below leaves marks the worklist form knows nothing about
  bt.lastWasWorklist = false;
  if (!bt.denseLoaded && !bt.hasInfKnown) {  // no infinite bound in base or diffs: kernel variants without infinity tests
    int bad = 0;
    CK(cudaMemsetAsync(mat->dRes, 0, sizeof(int), st));
    bt_has_inf_kernel<<<std::max(1, std::min(1024, cdiv(std::max<long long>(mat->n, bt.nDelta), 256))), 256, 0, st>>>(bt.baseLb,
      bt.baseUb, mat->n, bt.dBnd, bt.nDelta, mat->inf, mat->dRes);
    ++ctx->stats.kernelLaunches;
    CK(cudaMemcpyAsync(&bad, mat->dRes, sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    ++hostSyncs;
    bt.hasInf = bad != 0;
    bt.hasInfKnown = true;
  }
  if (!bt.dirty0Clean)
    CK(cudaMemsetAsync(bt.dirty[0], 0, sizeof(unsigned) * words, st));
  bt.dirty0Clean = false;
  bool restZeroed = false;  // dirty[1] and ever: not needed by a batch that ends quietly after its first round
  auto zero_rest = [&]() -> int {
    if (!restZeroed) {
      CK(cudaMemsetAsync(bt.dirty[1], 0, sizeof(unsigned) * words, st));
      CK(cudaMemsetAsync(bt.ever, 0, sizeof(unsigned) * words, st));
      restZeroed = true;
    }
    return CBP_OK;
  };
  bt.quiet = false;
  bt_init_state_kernel<<<cdiv(bt.Bp, 256), 256, 0, st>>>(p, bt.B);
  ++ctx->stats.kernelLaunches;
  const int cellGrid = cdiv(static_cast<long long>(bt.cells), 256);
  bool fullPrepass = true, lazy = false, fromBase = false;
  int rc0 = 0;
  if (!bt.denseLoaded) {
    // base + deltas: expand here, so a loaded batch can be run any number of times
    if (bt.consumed) {
      ctx->err = "cbp_batch_run: this batch was loaded dense and has already been run";
      return CBP_ERR_ARG;
    }
    if (bt.checkedTol != primalTol) {  // (the loaded base bounds do not change between runs)
      int need = 0;
      CK(cudaMemsetAsync(p.anyActive, 0, sizeof(int), st));
      bt_base_check_kernel<<<cdiv(mat->n, 256), 256, 0, st>>>(bt.baseLb, bt.baseUb, mat->colType, mat->n, p.s,
        p.anyActive);
      ++ctx->stats.kernelLaunches;
      CK(cudaMemcpyAsync(&need, p.anyActive, sizeof(int), cudaMemcpyDeviceToHost, st));
      CK(cudaStreamSynchronize(st));
      ++hostSyncs;
      bt.baseNeedsPrepass = need != 0;
      bt.checkedTol = primalTol;
    }
    CK(cudaMemsetAsync(bt.groupFlag, 0, sizeof(int) * (bt.G + 2), st));
    fullPrepass = bt.baseNeedsPrepass;
    // Lazy second buffer: most vectors of a node batch reach their fixpoint in the first round without
    // a single accepted candidate, so the other bound buffer is built after that round and only for the
    // vector groups that have candidates (collected in a list) or R11 changes.
    // (worth it for shallow changes per vector — strong-branching arms, probes; deep dives tighten so much
    // in their first round that the direct two-buffer round is the cheaper one)
    lazy = !fullPrepass && bt.W == 32 && !knobs().noLazyX1 && (bt.nDelta <= 256ll * bt.B || knobs().lazyX1);
    // ... and when that is so, neither buffer is built before the first round: it runs "from the base"
    // (shared base bounds + the cells the diffs touch), and the buffers are built afterwards for the
    // groups that go on
    fromBase = lazy && !knobs().noBaseRound;
    if (!fromBase && (rc0 = zero_rest()))
      return rc0;
    if (fromBase) {
      bt_base_pack_kernel<<<cdiv(mat->n, 256), 256, 0, st>>>(bt.baseLb, bt.baseUb, mat->n, bt.baseX);
      if (!bt.dmaskClean)
        CK(cudaMemsetAsync(bt.dmask, 0, sizeof(unsigned) * words, st));
      bt.dmaskClean = false;
    } else {
      bt_expand_kernel<<<cellGrid, 256, 0, st>>>(bt.baseLb, bt.baseUb, mat->n, bt.W, bt.cells, bt.X[0],
        lazy ? nullptr : bt.X[1]);
    }
    ++ctx->stats.kernelLaunches;
    if (bt.nDelta > 0) {
      bt_delta_kernel<<<cdiv(bt.B, 128), 128, 0, st>>>(bt.dStart, bt.dVar, bt.dBnd, bt.B, p, fullPrepass ? 0 : 1,
        fromBase ? 2 : (lazy ? 1 : 0));
      ++ctx->stats.kernelLaunches;
    }
  } else {
    if (bt.consumed) {
      ctx->err = "cbp_batch_run: a dense batch can be run once";
      return CBP_ERR_ARG;
    }
    bt.consumed = true;
    if ((rc0 = zero_rest()))
      return rc0;
  }
  if (fullPrepass) {
    bt_prepass_kernel<<<cellGrid, 256, 0, st>>>(p, bt.B);
    ++ctx->stats.kernelLaunches;
  }
  bt.baseMode = fromBase;
  bt.groupBuilt.assign(bt.G, fromBase ? 0 : 1);
  bt_prepass_state_kernel<<<cdiv(bt.Bp, 256), 256, 0, st>>>(p);
  ++ctx->stats.kernelLaunches;
  CK(cudaGetLastError());
  const long long warps = static_cast<long long>(bt.G) * mat->nShortBt;
  const int rowsGrid = cdiv(warps, kWarps);
  const long long longGrid = static_cast<long long>(bt.G) * mat->nLongBt;
  if (longGrid > INT_MAX || warps / kWarps > INT_MAX) {
    ctx->err = "cbp_batch_run: batch too large for one launch";
    return CBP_ERR_ARG;
  }
  // settle / carry: blockIdx.y = vector group, x strides over blocks of 32 columns
  const dim3 markGrid(std::max(1, std::min(cdiv(mat->n, 32 * kWarps), 4096 / std::max(1, bt.G) + 1)), bt.G);
  double roundKernelMs = 0.0;
  int roundKernelLaunches = 0;
  int any = 1;
  bool overlay = fromBase;  // the bound buffers hold only the cells diffs or candidates have touched
  while (ctx->evPool.size() < 2) {
    cudaEvent_t e;
    CK(cudaEventCreate(&e));
    ctx->evPool.push_back(e);
  }
  for (int round = 0; any && !(timeLimitSec <= 0.0); ++round) {
    const int cur = round & 1;
    CK(cudaMemsetAsync(p.anyActive, 0, sizeof(int), st));
    if (overlay && round > 0)  // a new list (the first round's flags were seeded by the diffs' R11)
      CK(cudaMemsetAsync(bt.groupFlag, 0, sizeof(int) * (bt.G + 2), st));
    CK(cudaEventRecord(ctx->evPool[0], st));
    // kernel variants: <no infinite bound in this batch, no continuous column in this matrix, candidates to a list>
    const int variant = (bt.hasInf ? 0 : 2) | (mat->allInt ? 1 : 0);
    auto launch_rows = [&](bool toList, bool base) {
      const int mode = base ? 2 : (toList ? 1 : 0);  // base implies the list
      if (longGrid > 0 && mode == 2 && !bt.hasInf && !knobs().noMultiGroup) {
        // no infinite bound anywhere: four vector groups per warp share every staged matrix entry
        constexpr int V = kBaseGroups;
        const long long lg = static_cast<long long>(cdiv(bt.G, V)) * mat->nLongBt;
        bt_rows_long_base_kernel<V><<<static_cast<int>(lg), kBaseWarps * 32, 0, st>>>(p, cur);
        ++ctx->stats.kernelLaunches;
      } else if (longGrid > 0) {
        const int lg = static_cast<int>(longGrid);
        const int nt = kLongWarps * 32;
#define CBP_LONG(A, B_)                                                              \
  if (mode == 2)                                                                     \
    bt_rows_long_kernel<A, B_, true, true><<<lg, nt, 0, st>>>(p, cur);               \
  else if (mode == 1)                                                                \
    bt_rows_long_kernel<A, B_, true><<<lg, nt, 0, st>>>(p, cur);                     \
  else                                                                               \
    bt_rows_long_kernel<A, B_, false><<<lg, nt, 0, st>>>(p, cur);
        switch (variant) {
        case 0: CBP_LONG(false, false) break;
        case 1: CBP_LONG(false, true) break;
        case 2: CBP_LONG(true, false) break;
        default: CBP_LONG(true, true) break;
        }
#undef CBP_LONG
        ++ctx->stats.kernelLaunches;
      }
      if (warps > 0 && mode == 2 && !bt.hasInf && !knobs().noMultiGroup) {
        constexpr int V = kBaseGroups;
        const dim3 sg(cdiv(mat->nShortBt, kWarps), cdiv(bt.G, V));
        bt_rows_short_base_kernel<V><<<sg, kThreads, 0, st>>>(p, cur);
        ++ctx->stats.kernelLaunches;
      } else if (warps > 0 && bt.W == 32) {
#define CBP_SHORT(A, B_)                                                             \
  if (mode == 2)                                                                     \
    bt_rows_kernel<A, B_, true, true><<<rowsGrid, kThreads, 0, st>>>(p, cur);        \
  else if (mode == 1)                                                                \
    bt_rows_kernel<A, B_, true><<<rowsGrid, kThreads, 0, st>>>(p, cur);              \
  else                                                                               \
    bt_rows_kernel<A, B_, false><<<rowsGrid, kThreads, 0, st>>>(p, cur);
        switch (variant) {
        case 0: CBP_SHORT(false, false) break;
        case 1: CBP_SHORT(false, true) break;
        case 2: CBP_SHORT(true, false) break;
        default: CBP_SHORT(true, true) break;
        }
#undef CBP_SHORT
        ++ctx->stats.kernelLaunches;
      } else if (warps > 0) {
        const int W = bt.W;  // vectors per warp = slice width of the layout
        const int rowsPerWarp = 32 / W;
        const long long narrowWarps = static_cast<long long>(bt.G) * rowsPerWarp * cdiv(mat->nShortBt, rowsPerWarp);
        const int narrowGrid = cdiv(narrowWarps, kWarps);
#define CBP_NARROW(WW)                                                                                      \
  switch (variant) {                                                                                        \
  case 0: bt_rows_narrow_kernel<WW, false, false><<<narrowGrid, kThreads, 0, st>>>(p, cur); break;          \
  case 1: bt_rows_narrow_kernel<WW, false, true><<<narrowGrid, kThreads, 0, st>>>(p, cur); break;           \
  case 2: bt_rows_narrow_kernel<WW, true, false><<<narrowGrid, kThreads, 0, st>>>(p, cur); break;           \
  default: bt_rows_narrow_kernel<WW, true, true><<<narrowGrid, kThreads, 0, st>>>(p, cur); break;           \
  }
        if (W == 16) {
          CBP_NARROW(16)
        } else if (W == 8) {
          CBP_NARROW(8)
        } else {
          CBP_NARROW(4)
        }
#undef CBP_NARROW
        ++ctx->stats.kernelLaunches;
      }
    };
    const bool listRound = overlay || (lazy && round == 0);
    launch_rows(listRound, overlay);
    CK(cudaEventRecord(ctx->evPool[1], st));
    if (overlay && round == 0) {  // in case the round turns out quiet (the kernel checks for itself): saves a round trip
      bt_state_kernel<<<cdiv(bt.Bp, 256), 256, 0, st>>>(p, cur, round, limit, 0, bt.groupFlag + bt.G);
      ++ctx->stats.kernelLaunches;
    }
    if (listRound) {
      // did anybody get a candidate, and did the list hold them all?  (first round: also, which groups did R11 touch)
      std::vector<int> flags(bt.G + 1);
      CK(cudaMemcpyAsync(flags.data(), bt.groupFlag, sizeof(int) * (bt.G + 1), cudaMemcpyDeviceToHost, st));
      CK(cudaStreamSynchronize(st));
      ++hostSyncs;
      const unsigned count = static_cast<unsigned>(flags[bt.G]);
      bool anyFlag = false;
      for (int g = 0; g < bt.G; ++g)
        anyFlag = anyFlag || flags[g] != 0;
      if (overlay && round == 0 && !anyFlag && count == 0) {
        // nobody has a candidate and R11 touched nothing: no mark was set, every vector stops here
        // (fixpoint, or infeasible row); nothing to settle, nothing to carry, no update to report
        // (the guarded state kernel above has done the bookkeeping)
        if (bt.nDelta > 0) {
          bt_delta_unmask_kernel<<<cdiv(bt.B, 128), 128, 0, st>>>(bt.dStart, bt.dVar, bt.B, p);
          ++ctx->stats.kernelLaunches;
        }
        bt.dmaskClean = true;
        bt.dirty0Clean = true;
        bt.quiet = true;
        float ms = 0.f;
        CK(cudaEventSynchronize(ctx->evPool[1]));
        cudaEventElapsedTime(&ms, ctx->evPool[0], ctx->evPool[1]);
        roundKernelMs += ms;
        ++roundKernelLaunches;
        break;
      }
      if ((rc0 = zero_rest()))
        return rc0;
      // CBP_OVERLAY_ROUNDS: tests leave the overlay form after that many rounds to exercise the change-over
      const int overlayRounds = knobs().overlayRounds;
      if (count > bt.logCap || (overlay && round >= overlayRounds)) {
        // too many for the list: build the whole buffers and redo the round directly
        if (overlay) {
          bt_overlay_expand_kernel<<<cellGrid, 256, 0, st>>>(p);
          overlay = false;  // the classic two-buffer rounds from here on
          bt.groupBuilt.assign(bt.G, 1);
        } else {
          bt_copy_groups_kernel<<<cellGrid, 256, 0, st>>>(p, nullptr);
        }
        ++ctx->stats.kernelLaunches;
        launch_rows(false, false);
      } else if (overlay) {
        // the bounds stay in overlay form: cells the candidates touch for the first time are created, then the
        // list is applied exactly as the direct sink would have done
        if (count > 0) {
          bt_overlay_touch_kernel<<<cdiv(count, 256), 256, 0, st>>>(p, count);
          bt_apply_log_kernel<<<cdiv(count, 256), 256, 0, st>>>(p, count, cur);
          ctx->stats.kernelLaunches += 2;
        }
      } else if (anyFlag) {
        bt_copy_groups_kernel<<<cellGrid, 256, 0, st>>>(p, bt.groupFlag);
        ++ctx->stats.kernelLaunches;
        if (count > 0) {
          bt_apply_log_kernel<<<cdiv(count, 256), 256, 0, st>>>(p, count, cur);
          ++ctx->stats.kernelLaunches;
        }
      }
    }
    bt_settle_kernel<<<markGrid, kThreads, 0, st>>>(p, cur);
    int timeUp = 0;
    if (timeLimitSec < 1.0e9) {  // a real limit: look at the clock between rounds
      CK(cudaEventRecord(ctx->evB, st));
      CK(cudaEventSynchronize(ctx->evB));
      float ms = 0.f;
      cudaEventElapsedTime(&ms, ctx->evA, ctx->evB);
      timeUp = ms * 1.0e-3 >= timeLimitSec;
    }
    bt_state_kernel<<<cdiv(bt.Bp, 256), 256, 0, st>>>(p, cur, round, limit, timeUp);
    bt_carry_kernel<<<markGrid, kThreads, 0, st>>>(p, cur);
    ctx->stats.kernelLaunches += 3;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(&any, p.anyActive, sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    ++hostSyncs;
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->evPool[0], ctx->evPool[1]);
    roundKernelMs += ms;
    ++roundKernelLaunches;
  }
  CK(cudaEventRecord(ctx->evB, st));
  CK(cudaStreamSynchronize(st));
  float ms = 0.f;
  cudaEventElapsedTime(&ms, ctx->evA, ctx->evB);
  ctx->stats.lastRunMs = ms;
  ctx->stats.lastRoundKernelMs = roundKernelMs;
  ctx->stats.lastRoundKernelLaunches = roundKernelLaunches;
  // rounds summed over vectors (for throughput accounting); a quiet run: one round each
  long long tot = bt.B;
  if (!bt.quiet) {
    std::vector<int> rounds(bt.Bp);
    CK(cudaMemcpy(rounds.data(), p.vecRounds, sizeof(int) * bt.Bp, cudaMemcpyDeviceToHost));
    tot = 0;
    for (int b = 0; b < bt.B; ++b)
      tot += rounds[b];
  }
  ctx->stats.lastRoundsTotal = tot;
  ctx->stats.lastMode = fromBase ? 1 : 0;
  ctx->stats.lastPairsWalked = 0;
  ctx->stats.lastItems = 0;
  ctx->stats.lastHostSyncs = hostSyncs + 1;
  ctx->stats.lastFromBase = fromBase ? 1 : 0;
  int built = 0;
  for (int g = 0; g < bt.G; ++g)
    built += bt.groupBuilt[g] ? 1 : 0;
  ctx->stats.lastGroupsBuilt = built;
  return CBP_OK;
}

static int bt_fetch_results(cbp_matrix *mat, cbp_result *out, bool devOut)
{
  cbp_ctx *ctx = mat->ctx;
  BatchState &bt = mat->bt;
  std::vector<int> hv;
  const int *h = reinterpret_cast<const int *>(bt.hBlob + bt.blobOffInts);  // (a worklist run has brought them over already)
  if (!bt.lastWasWorklist) {
    hv.resize(static_cast<size_t>(bt.Bp) * 7);
    CK(cudaMemcpyAsync(hv.data(), bt.ints, sizeof(int) * hv.size(), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    h = hv.data();
  }
  std::vector<cbp_result> tmp;
  cbp_result *dstp = out;
  if (devOut) {
    tmp.resize(bt.B);
    dstp = tmp.data();
  }
  const int Bp = bt.Bp;
  for (int b = 0; b < bt.B; ++b) {
    cbp_result r;
    memset(&r, 0, sizeof(r));
    const int ir = h[b], ic = h[Bp + b];
    r.rounds = h[3 * Bp + b];
    r.nChanged = h[4 * Bp + b];
    r.stopReason = h[5 * Bp + b];
    r.feasible = r.stopReason != CBP_STOP_INFEASIBLE;
    r.infeasibleRow = -1;
    r.infeasibleCol = -1;
    if (!r.feasible) {
      if (ir != INT_MAX)
        r.infeasibleRow = ir;
      else if (ic != INT_MAX)
        r.infeasibleCol = ic;
    }
    if (r.stopReason == CBP_STOP_NOT_RUN && r.rounds == 0)
      r.stopReason = CBP_STOP_TIME_LIMIT;
    r.seconds = ctx->stats.lastRunMs * 1.0e-3;
    dstp[b] = r;
  }
  if (devOut)
    CK(cudaMemcpy(out, tmp.data(), sizeof(cbp_result) * bt.B, cudaMemcpyHostToDevice));
  return CBP_OK;
}

int cbp_batch_fetch(cbp_matrix *mat, cbp_result *out, int64_t updCap, int64_t *updStart, int *updCol,
  double *updLb, double *updUb, int flags)
{
  if (!mat || !mat->bt.loaded || !out)
    return CBP_ERR_ARG;
  cbp_ctx *ctx = mat->ctx;
  CK(cudaSetDevice(ctx->device));
  BatchState &bt = mat->bt;
  const bool dev = flags & CBP_FLAG_DEVICE_PTRS;
  int rc = bt_fetch_results(mat, out, dev);
  if (rc)
    return rc;
  if (!updStart)
    return CBP_OK;
  if (!bt.listsWanted) {  // from now on the runs on this matrix write the lists themselves
    bt.listsWanted = true;
    const int rcl = bt_list_buffers(mat);
    if (rcl)
      return rcl;
  }
  if (bt.quiet) {  // the run set no mark at all: no vector has an update
    if (dev)
      CK(cudaMemsetAsync(updStart, 0, sizeof(long long) * (bt.B + 1), ctx->stream));
    else
      memset(updStart, 0, sizeof(long long) * (bt.B + 1));
    return CBP_OK;
  }
  const BtParams p = bt_params(mat, 0.0);
  if (bt.lastWasWorklist && bt.updReady) {  // the run wrote the lists itself: copies only
    const long long run = bt.updTotal;
    if (dev)
      CK(cudaMemcpyAsync(updStart, bt.counts, sizeof(long long) * (bt.B + 1), cudaMemcpyDeviceToDevice, ctx->stream));
    else
      memcpy(updStart, bt.hBlob + bt.blobOffCounts, sizeof(long long) * (bt.B + 1));
    if (run > updCap || (run > 0 && (!updCol || !updLb || !updUb))) {
      ctx->err = "cbp_batch_fetch: update buffers too small";
      return CBP_ERR_CAPACITY;
    }
    if (dev) {
      if (run > 0) {
        CK(cudaMemcpyAsync(updCol, bt.stCol, sizeof(int) * run, cudaMemcpyDeviceToDevice, ctx->stream));
        CK(cudaMemcpyAsync(updLb, bt.stLb, sizeof(double) * run, cudaMemcpyDeviceToDevice, ctx->stream));
        CK(cudaMemcpyAsync(updUb, bt.stUb, sizeof(double) * run, cudaMemcpyDeviceToDevice, ctx->stream));
      }
      CK(cudaStreamSynchronize(ctx->stream));
    } else if (run == 0) {
    } else if (run <= bt.headCopied) {
      const WlUpdHead *hd = reinterpret_cast<const WlUpdHead *>(bt.hBlob + bt.blobOffHead);
      for (long long i = 0; i < run; ++i) {
        updCol[i] = hd[i].col;
        updLb[i] = hd[i].lb;
        updUb[i] = hd[i].ub;
      }
    } else {
      CK(cudaMemcpyAsync(updCol, bt.stCol, sizeof(int) * run, cudaMemcpyDeviceToHost, ctx->stream));
      CK(cudaMemcpyAsync(updLb, bt.stLb, sizeof(double) * run, cudaMemcpyDeviceToHost, ctx->stream));
      CK(cudaMemcpyAsync(updUb, bt.stUb, sizeof(double) * run, cudaMemcpyDeviceToHost, ctx->stream));
      CK(cudaStreamSynchronize(ctx->stream));
    }
    return CBP_OK;
  }
  if (bt.lastWasWorklist) {
    // the run listed every cell it created: the update lists come from that list (count, fill, order per vector)
    // instead of two scans over the n x B / 32 words of marks
    const unsigned nCells = bt.wlPrevCells;
    int *cnt32 = bt.wlInts;  // (the run is over: its per-vector ints are free)  [0, Bp) counts, [Bp, 2Bp) cursors
    CK(cudaMemsetAsync(cnt32, 0, sizeof(int) * 2 * static_cast<size_t>(bt.Bp), ctx->stream));
    if (nCells > 0) {
      wl_upd_count_kernel<<<cdiv(nCells, 256), 256, 0, ctx->stream>>>(bt.wlCells, nCells, p, cnt32);
      ++ctx->stats.kernelLaunches;
    }
    std::vector<int> c32(bt.Bp);
    CK(cudaMemcpyAsync(c32.data(), cnt32, sizeof(int) * bt.Bp, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    std::vector<long long> off(bt.B + 1);
    long long run = 0;
    for (int b = 0; b < bt.B; ++b) {
      off[b] = run;
      run += c32[b];
    }
    off[bt.B] = run;
    if (dev)
      CK(cudaMemcpy(updStart, off.data(), sizeof(long long) * (bt.B + 1), cudaMemcpyHostToDevice));
    else
      memcpy(updStart, off.data(), sizeof(long long) * (bt.B + 1));
    if (run > updCap || (run > 0 && (!updCol || !updLb || !updUb))) {
      ctx->err = "cbp_batch_fetch: update buffers too small";
      return CBP_ERR_CAPACITY;
    }
    if (run == 0)
      return CBP_OK;
    auto grow = [&](int *&c, double *&l, double *&u, long long &cap) -> int {
      if (run <= cap)
        return CBP_OK;
      bt_free_lists(bt, &c == &bt.stCol, &c == &bt.st2Col);
      const long long want = std::max<long long>(run + run / 2, 1 << 16);
      if (dalloc(&c, want) != cudaSuccess || dalloc(&l, want) != cudaSuccess || dalloc(&u, want) != cudaSuccess) {
        ctx->err = "cbp_batch_fetch: staging";
        return CBP_ERR_NOMEM;
      }
      cap = want;
      return CBP_OK;
    };
    int rcg = grow(bt.st2Col, bt.st2Lb, bt.st2Ub, bt.st2Cap);
    if (rcg)
      return rcg;
    int *oCol = updCol;
    double *oLb = updLb, *oUb = updUb;
    if (!dev) {
      if ((rcg = grow(bt.stCol, bt.stLb, bt.stUb, bt.stCap)))
        return rcg;
      oCol = bt.stCol;
      oLb = bt.stLb;
      oUb = bt.stUb;
    }
    CK(cudaMemcpyAsync(bt.counts, off.data(), sizeof(long long) * (bt.B + 1), cudaMemcpyHostToDevice, ctx->stream));
    wl_upd_fill_kernel<<<cdiv(nCells, 256), 256, 0, ctx->stream>>>(bt.wlCells, nCells, p, bt.counts, cnt32 + bt.Bp, bt.st2Col,
      bt.st2Lb, bt.st2Ub);
    wl_upd_sort_kernel<<<bt.B, kThreads, 0, ctx->stream>>>(p, bt.counts, bt.B, bt.st2Col, bt.st2Lb,
      bt.st2Ub, oCol, oLb, oUb);
    ctx->stats.kernelLaunches += 2;
    CK(cudaGetLastError());
    if (!dev) {
      CK(cudaMemcpyAsync(updCol, oCol, sizeof(int) * run, cudaMemcpyDeviceToHost, ctx->stream));
      CK(cudaMemcpyAsync(updLb, oLb, sizeof(double) * run, cudaMemcpyDeviceToHost, ctx->stream));
      CK(cudaMemcpyAsync(updUb, oUb, sizeof(double) * run, cudaMemcpyDeviceToHost, ctx->stream));
    }
    CK(cudaStreamSynchronize(ctx->stream));
    return CBP_OK;
  }
  const int nChunks = cdiv(mat->n, kUpdChunk);
  const dim3 updGrid(bt.G, cdiv(nChunks, 4)), updBlock(128);
  bt_updates_kernel<false><<<updGrid, updBlock, 0, ctx->stream>>>(p, bt.B, bt.chunkCnt, nullptr, nullptr, nullptr,
    nullptr);
  bt_updates_prefix_kernel<<<cdiv(bt.B, 128), 128, 0, ctx->stream>>>(bt.chunkCnt, nChunks, bt.Bp, bt.B, bt.counts);
  ctx->stats.kernelLaunches += 2;
  std::vector<long long> cnt(bt.B + 1);
  CK(cudaMemcpyAsync(cnt.data(), bt.counts, sizeof(long long) * bt.B, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  long long run = 0;
  for (int b = 0; b < bt.B; ++b) {
    const long long c = cnt[b];
    cnt[b] = run;
    run += c;
  }
  cnt[bt.B] = run;
  if (dev)
    CK(cudaMemcpy(updStart, cnt.data(), sizeof(long long) * (bt.B + 1), cudaMemcpyHostToDevice));
  else
    memcpy(updStart, cnt.data(), sizeof(long long) * (bt.B + 1));
  if (run > updCap || (run > 0 && (!updCol || !updLb || !updUb))) {
    ctx->err = "cbp_batch_fetch: update buffers too small";
    return CBP_ERR_CAPACITY;
  }
  if (run == 0)
    return CBP_OK;
  CK(cudaMemcpyAsync(bt.counts, cnt.data(), sizeof(long long) * (bt.B + 1), cudaMemcpyHostToDevice,
    ctx->stream));
  int *dCol = updCol;
  double *dLb = updLb, *dUb = updUb;
  if (!dev) {
    if (run > bt.stCap) {
      bt_free_lists(bt, true, false);
      const long long cap = std::max<long long>(run + run / 2, 1 << 16);
      if (dalloc(&bt.stCol, cap) != cudaSuccess || dalloc(&bt.stLb, cap) != cudaSuccess || dalloc(&bt.stUb, cap) != cudaSuccess) {
        ctx->err = "cbp_batch_fetch: staging";
        return CBP_ERR_NOMEM;
      }
      bt.stCap = cap;
    }
    dCol = bt.stCol;
    dLb = bt.stLb;
    dUb = bt.stUb;
  }
  bt_updates_kernel<true><<<updGrid, updBlock, 0, ctx->stream>>>(p, bt.B, bt.chunkCnt, bt.counts, dCol, dLb, dUb);
  ++ctx->stats.kernelLaunches;
  cudaError_t e = cudaGetLastError();
  if (!dev) {
    if (e == cudaSuccess)
      e = cudaMemcpyAsync(updCol, dCol, sizeof(int) * run, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess)
      e = cudaMemcpyAsync(updLb, dLb, sizeof(double) * run, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess)
      e = cudaMemcpyAsync(updUb, dUb, sizeof(double) * run, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess)
      e = cudaStreamSynchronize(ctx->stream);
  } else if (e == cudaSuccess) {
    e = cudaStreamSynchronize(ctx->stream);
  }
  CK(e);
  return CBP_OK;
}

int cbp_batch_fetch_dense(cbp_matrix *mat, double *lb, double *ub, int flags)
{
  if (!mat || !mat->bt.loaded || !lb || !ub)
    return CBP_ERR_ARG;
  cbp_ctx *ctx = mat->ctx;
  CK(cudaSetDevice(ctx->device));
  BatchState &bt = mat->bt;
  const bool dev = flags & CBP_FLAG_DEVICE_PTRS;
  const BtParams p = bt_params(mat, 0.0);
  const size_t total = static_cast<size_t>(bt.B) * mat->n;
  double *dl = lb, *du = ub;
  if (!dev) {
    CK(dalloc(&dl, total));
    if (dalloc(&du, total) != cudaSuccess) {
      cudaFree(dl);
      ctx->err = "cbp_batch_fetch_dense: staging";
      return CBP_ERR_NOMEM;
    }
  }
  if (bt.baseMode && !bt.groupBuilt.empty() && !bt.groupBuilt[0]) {
    // the run stayed in overlay form (base bounds + touched cells): build the full buffers now
    bt_overlay_expand_kernel<<<cdiv(static_cast<long long>(bt.cells), 256), 256, 0, ctx->stream>>>(p);
    ++ctx->stats.kernelLaunches;
    bt.groupBuilt.assign(bt.G, 1);
  }
  dim3 grid(cdiv(mat->n, 32), cdiv(bt.B, 32)), block(32, 8);
  bt_transpose_out_kernel<<<grid, block, 0, ctx->stream>>>(p, bt.B, dl, du);
  ++ctx->stats.kernelLaunches;
  cudaError_t e = cudaGetLastError();
  if (!dev) {
    if (e == cudaSuccess)
      e = cudaMemcpyAsync(lb, dl, sizeof(double) * total, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess)
      e = cudaMemcpyAsync(ub, du, sizeof(double) * total, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess)
      e = cudaStreamSynchronize(ctx->stream);
    cudaFree(dl);
    cudaFree(du);
  } else if (e == cudaSuccess) {
    e = cudaStreamSynchronize(ctx->stream);
  }
  CK(e);
  return CBP_OK;
}

int cbp_propagate_batch(cbp_matrix *mat, int B, double *lb, double *ub, double primalTol,
  int maxRounds, double timeLimitSec, cbp_result *out, int flags)
{
  if (!mat || B <= 0 || !lb || !ub || !out)
    return CBP_ERR_ARG;
  cbp_ctx *ctx = mat->ctx;
  CK(cudaSetDevice(ctx->device));
  int rc = bt_reserve(mat, B);
  if (rc)
    return rc;
  BatchState &bt = mat->bt;
  const bool dev = flags & CBP_FLAG_DEVICE_PTRS;
  const size_t total = static_cast<size_t>(B) * mat->n;
  double *dl = lb, *du = ub;
  if (!dev) {
    CK(dalloc(&dl, total));
    if (dalloc(&du, total) != cudaSuccess) {
      cudaFree(dl);
      ctx->err = "cbp_propagate_batch: staging";
      return CBP_ERR_NOMEM;
    }
    cudaError_t e = cudaMemcpyAsync(dl, lb, sizeof(double) * total, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess)
      e = cudaMemcpyAsync(du, ub, sizeof(double) * total, cudaMemcpyHostToDevice, ctx->stream);
    if (e != cudaSuccess) {
      cudaFree(dl);
      cudaFree(du);
      CK(e);
    }
  }
  dim3 grid(cdiv(mat->n, 32), bt.Bp / 32), block(32, 8);
  bt_transpose_in_kernel<<<grid, block, 0, ctx->stream>>>(dl, du, B, bt_params(mat, primalTol), 0.0, 0.0);
  ++ctx->stats.kernelLaunches;
  bt.loaded = true;
  bt.denseLoaded = true;
  bt.consumed = false;
  bt.hasInf = true;
  rc = cbp_batch_run(mat, primalTol, maxRounds, timeLimitSec);
  if (rc == CBP_OK)
    rc = bt_fetch_results(mat, out, dev);
  if (rc == CBP_OK) {
    const BtParams p = bt_params(mat, 0.0);
    dim3 g2(cdiv(mat->n, 32), cdiv(B, 32));
    bt_transpose_out_kernel<<<g2, block, 0, ctx->stream>>>(p, B, dl, du);
    ++ctx->stats.kernelLaunches;
    cudaError_t e = cudaGetLastError();
    if (!dev) {
      if (e == cudaSuccess)
        e = cudaMemcpyAsync(lb, dl, sizeof(double) * total, cudaMemcpyDeviceToHost, ctx->stream);
      if (e == cudaSuccess)
        e = cudaMemcpyAsync(ub, du, sizeof(double) * total, cudaMemcpyDeviceToHost, ctx->stream);
    }
    if (e == cudaSuccess)
      e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) {
      ctx->err = std::string("cbp_propagate_batch: ") + cudaGetErrorString(e);
      rc = CBP_ERR_CUDA;
    }
  }
  if (!dev) {
    cudaFree(dl);
    cudaFree(du);
  }
  bt.loaded = rc == CBP_OK;
  return rc;
}

int cbp_propagate_batch_delta(cbp_matrix *mat, const double *baseLb, const double *baseUb, int B,
  const int64_t *deltaStart, const uint32_t *deltaVar, const double *deltaBound, double primalTol,
  int maxRounds, double timeLimitSec, cbp_result *out, int64_t updCap, int64_t *updStart, int *updCol,
  double *updLb, double *updUb, int flags)
{
  int rc = cbp_batch_load_delta(mat, baseLb, baseUb, B, deltaStart, deltaVar, deltaBound, flags);
  if (rc)
    return rc;
  if (updStart)  // (the run then writes the lists itself: fetching them is a copy)
    mat->bt.listsWanted = true;
  if ((rc = cbp_batch_run(mat, primalTol, maxRounds, timeLimitSec)))
    return rc;
  return cbp_batch_fetch(mat, out, updCap, updStart, updCol, updLb, updUb, flags);
}

}  // extern "C"
