/*
 * cbp_capi.h — C ABI of the B200 bound-propagation engine (libcbp.so).
 *
 * This is the drop-in boundary for the hot path of
 * /root/reference/src/CbcBoundPropagation.cpp: everything `run()` does between
 * extracting the problem data (:90-105) and applying the tightened bounds
 * (:165-194) happens behind these entry points, on the GPU.  Plain pointers
 * and sizes only; no C++ or torch types cross.  All entry points return 0 on
 * success and a negative CBP_ERR_* code on failure (cbp_last_error gives the
 * text); there is no CPU fallback: without a usable CUDA device cbp_create
 * fails.
 *
 * Pointer arguments are HOST pointers unless CBP_FLAG_DEVICE_PTRS is given, in
 * which case every bound / result array of that call is a device pointer on
 * the context's device.
 */
#ifndef CBP_CAPI_H
#define CBP_CAPI_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define CBP_API __attribute__((visibility("default")))
#else
#define CBP_API
#endif

typedef struct cbp_ctx cbp_ctx;
typedef struct cbp_matrix cbp_matrix;

enum {
  CBP_OK = 0,
  CBP_ERR_ARG = -1,     /* bad argument */
  CBP_ERR_CUDA = -2,    /* a CUDA call failed; see cbp_last_error */
  CBP_ERR_NOMEM = -3,   /* device or host allocation failed */
  CBP_ERR_CAPACITY = -4 /* a caller-provided output buffer is too small */
};

enum {
  CBP_FLAG_DEVICE_PTRS = 1, /* bound/result arrays are device pointers */
  /* cbp_propagate only: run the row phase of exactly ONE round over this matrix's rows and return
   * the raw Jacobi proposals (max of accepted lower, min of accepted upper candidates per column),
   * without R11, without settling crossings (R9) and without bookkeeping.  out->feasible == 0 only
   * for a violated row (infeasibleRow).  For row-sharded multi-GPU runs: every rank does this on its
   * row block, the proposals are combined with one max/min all-reduce, then settled once. */
  CBP_FLAG_RAW_ROUND = 2
};

/* Stop reasons; numbering of CbcBoundPropagation::StopReason
 * (src/CbcBoundPropagation.hpp:60-66). */
enum {
  CBP_STOP_NOT_RUN = 0,
  CBP_STOP_FIXPOINT = 1,
  CBP_STOP_MAX_ROUNDS = 2,
  CBP_STOP_TIME_LIMIT = 3,
  CBP_STOP_INFEASIBLE = 4
};

/* What one propagation run reports; the fields are what
 * CbcBoundPropagation keeps after the round loop
 * (src/CbcBoundPropagation.cpp:129-133,166-189,197-201). */
typedef struct cbp_result {
  int feasible;      /* 0 iff infeasibility was proved (run() returns false) */
  int rounds;        /* nRoundsRun_: rounds executed, the empty fixpoint round included */
  int nChanged;      /* nBoundPropFixed_: sum over rounds of changed columns */
  int infeasibleRow; /* CoinBoundPropagation::infeasibleRow(), -1 if n/a (:132) */
  int infeasibleCol; /* CoinBoundPropagation::infeasibleCol(), -1 if n/a (:133) */
  int stopReason;    /* CBP_STOP_* */
  double seconds;    /* device time of the run */
} cbp_result;

/* Counters for benchmarks: kernels launched by this context so far and the
 * device time of the most recent run's kernels, taken with CUDA events on
 * the context's stream. */
typedef struct cbp_stats {
  int64_t kernelLaunches;
  double lastRunMs;       /* whole run (all kernels of the last propagate call) */
  double lastRoundKernelMs; /* time spent in the round kernel(s) of the last run */
  int lastRoundKernelLaunches;
  int64_t lastRoundsTotal; /* rounds summed over the vectors of the last run */
  int lastFromBase;        /* batch: the first round ran from the shared base bounds (no per-vector buffers) */
  int lastGroupsBuilt;     /* batch: vector groups (of 32) whose bound buffers were built afterwards */
  int lastMode;            /* batch: 0 classic two-buffer rounds, 1 overlay rounds that walk every (row, vector) pair,
                              2 worklist form (one persistent kernel, only the pairs a changed column touches) */
  int lastHostSyncs;       /* batch: host synchronisations between the start and the end of the run */
  int64_t lastPairsWalked; /* batch, worklist form: (nonzero, vector) pairs walked by all rounds */
  int64_t lastItems;       /* batch, worklist form: (row, vector) items evaluated by all rounds */
} cbp_stats;

/* -- context ----------------------------------------------------------- */

/* One context per calling thread / stream; device -1 = current device.
 * Replaces nothing in the reference (it has no device); needed because
 * run() is re-entered from Cbc worker threads on distinct solver clones
 * (src/CbcThread.cpp:1249,1319). */
CBP_API cbp_ctx *cbp_create(int device);
CBP_API void cbp_destroy(cbp_ctx *ctx);
/* Use an existing cudaStream_t (e.g. torch's current stream) for all work. */
CBP_API int cbp_set_stream(cbp_ctx *ctx, void *cudaStream);
CBP_API const char *cbp_last_error(const cbp_ctx *ctx);
CBP_API int cbp_get_stats(const cbp_ctx *ctx, cbp_stats *out);
CBP_API int cbp_synchronize(cbp_ctx *ctx);

/* -- problem data -------------------------------------------------------- */

/* Upload the data run() reads once per call: getMatrixByRow() (row-major
 * CoinPackedMatrix: vectorStarts / vectorLengths / indices / elements),
 * getRowSense(), getRightHandSide(), getRowRange(), getColType(false),
 * getInfinity()  (src/CbcBoundPropagation.cpp:91-99).  rowLen may be NULL when
 * the rows are stored without gaps (rowStart has m+1 entries).  The matrix is
 * re-tiled on the host and kept in HBM until cbp_matrix_free. */
CBP_API cbp_matrix *cbp_matrix_upload(cbp_ctx *ctx, int m, int n, const int64_t *rowStart,
  const int *rowLen, const int *colIdx, const double *val, const char *rowSense,
  const double *rhs, const double *range, const char *colType, double infinity);
CBP_API void cbp_matrix_free(cbp_matrix *mat);

/* -- the path itself ------------------------------------------------------- */

/* Phase 1, OsiSolverInterface::tightenBoundsFromSingletonRows(int &nFixed)
 * as called at src/CbcBoundPropagation.cpp:51: *nTightened < 0 means
 * infeasible, otherwise the number of columns whose bounds changed; *nFixed of
 * them ended fixed.  lb/ub (n doubles) are updated in place. */
CBP_API int cbp_singletons(cbp_matrix *mat, double *lb, double *ub, double primalTol,
  int *nTightened, int *nFixed, int flags);

/* Phase 2, the whole round loop of src/CbcBoundPropagation.cpp:107-195 (each
 * round = one CoinBoundPropagation construction, :124-127) run on the device
 * without host round trips.  maxRounds <= 0 means "until fixpoint" (Level
 * Fixpoint, :88).  lb/ub (n doubles) are updated in place with every completed
 * feasible round.  roundChanged (nullable, roundCap ints) receives the
 * per-round changed-column counts that run() prints at logLevel >= 2 (:191). */
CBP_API int cbp_propagate(cbp_matrix *mat, double *lb, double *ub, double primalTol,
  int maxRounds, double timeLimitSec, cbp_result *out, int *roundChanged, int roundCap,
  int flags);

/* B independent runs of the same loop over one matrix (new surface; per
 * vector it is exactly cbp_propagate).  lb/ub are B x n, row-major per vector. */
CBP_API int cbp_propagate_batch(cbp_matrix *mat, int B, double *lb, double *ub,
  double primalTol, int maxRounds, double timeLimitSec, cbp_result *out, int flags);

/* Batch in the node bound-diff wire format of CbcPartialNodeInfo
 * (src/CbcNode.cpp:265-273: variable | 0x80000000 = upper bound; the column is
 * bits 0-29, bit 30 is the "can go odd way" flag of applyBounds() and is
 * ignored, src/CbcPartialNodeInfo.cpp:139,196,208; a column >= n is refused
 * with CBP_ERR_ARG): vector b is
 * baseLb/baseUb with deltaVar/deltaBound[deltaStart[b] .. deltaStart[b+1])
 * applied in order.  Outputs: out[B]; and, as CoinBoundPropagation's
 * updatedBounds() would list them (src/CbcBoundPropagation.cpp:166,181-186),
 * for vector b the columns whose final bounds differ from its input:
 * updCol/updLb/updUb[updStart[b] .. updStart[b+1]).  updCap is the capacity of
 * the three upd arrays; if it is too small CBP_ERR_CAPACITY is returned and
 * updStart[B] holds the required size.  upd* may all be NULL. */
CBP_API int cbp_propagate_batch_delta(cbp_matrix *mat, const double *baseLb, const double *baseUb,
  int B, const int64_t *deltaStart, const uint32_t *deltaVar, const double *deltaBound,
  double primalTol, int maxRounds, double timeLimitSec, cbp_result *out,
  int64_t updCap, int64_t *updStart, int *updCol, double *updLb, double *updUb, int flags);

/* Device-resident batch, split into its three steps so a caller (or a
 * benchmark) can keep the inputs in HBM: load = upload base + deltas; run =
 * expand them into the per-vector bound arrays and do the round loop (may be
 * repeated: every run starts again from base + deltas); fetch = results and
 * changed bounds of the last run.  A run in worklist form (cbp_stats.lastMode
 * == 2) leaves every vector's result in pinned host memory when it returns,
 * and — once a fetch on this matrix has asked for update lists (updStart !=
 * NULL) — the lists too, ordered by column; fetching them is then a host-side
 * copy. */
CBP_API int cbp_batch_load_delta(cbp_matrix *mat, const double *baseLb, const double *baseUb,
  int B, const int64_t *deltaStart, const uint32_t *deltaVar, const double *deltaBound, int flags);
CBP_API int cbp_batch_run(cbp_matrix *mat, double primalTol, int maxRounds, double timeLimitSec);
CBP_API int cbp_batch_fetch(cbp_matrix *mat, cbp_result *out, int64_t updCap, int64_t *updStart,
  int *updCol, double *updLb, double *updUb, int flags);
/* Dense final bounds of the resident batch, B x n row-major per vector. */
CBP_API int cbp_batch_fetch_dense(cbp_matrix *mat, double *lb, double *ub, int flags);

/* Single bound vector over a matrix whose ROWS are sharded over the GPUs of one box (one process per
 * GPU; every rank uploads its own block of rows with cbp_matrix_upload and holds the whole bound
 * vector).  The round loop of src/CbcBoundPropagation.cpp:107-195 then runs as one persistent kernel
 * per rank.  Inside a round a rank collects the candidates its rows accept, copies the columns it
 * changed into its slot of every other rank's state block (NVLink peer mappings, plain stores), meets
 * the other ranks at a barrier kept in peer memory, and settles its own full copy of the bounds from
 * what all ranks sent — the exchange is part of the kernel, there is no separate collective and no
 * host round trip per round.
 *   cbp_p2p_export   : allocate this rank's state block for a group of `world` ranks (<= 8), return its
 *                      CUDA IPC handle (64 bytes);
 *   cbp_p2p_attach   : map all ranks' blocks (handles = world x 64 bytes, in rank order);
 *   cbp_p2p_propagate: collective over the group: every rank calls it with the same lb / ub, tolerance
 *                      and limits; rowOffset = index of this rank's first row in the whole matrix
 *                      (infeasibleRow is reported in whole-matrix numbering).  Every rank gets the same
 *                      result and the same bounds.  The kernels find each other on their own: no host
 *                      barrier is needed around the call.  A rank that waits ~30 s for the others to start,
 *                      or ~4 s at a barrier inside the loop, gives up; then every rank returns an error and
 *                      the group is unusable. */
#define CBP_P2P_HANDLE_BYTES 64
CBP_API int cbp_p2p_export(cbp_matrix *mat, int world, void *handleOut);
CBP_API int cbp_p2p_attach(cbp_matrix *mat, int rank, int world, const void *handles);
CBP_API int cbp_p2p_propagate(cbp_matrix *mat, double *lb, double *ub, double primalTol, int maxRounds,
  double timeLimitSec, int rowOffset, cbp_result *out, int *roundChanged, int roundCap, int flags);

/* Phase 1 for callers that cannot allocate device memory.  This entry point has no counterpart in
 * the reference (which has no device at all); it exists for two kinds of caller: a process whose
 * host application or LP code has already taken all free HBM into its own pool, where a late
 * cudaMalloc from a worker thread fails; and node-level propagation on several Cbc worker threads,
 * where cudaMalloc / cudaFree would stall the other threads' streams because both synchronise the
 * device.  It does what cbp_singletons does (same rules, same results) but takes EVERY array from
 * the caller as a device pointer on the current device — the row-major matrix exactly as
 * getMatrixByRow() lays it out (rowLen may be NULL), row sense / rhs / range, colType, and lb / ub
 * (n doubles each, updated in place; left untouched when *nTightened is -1) — and allocates
 * nothing: its working memory is a fixed array inside the library image (about 4.1 MB, resident on
 * every GPU that loads libcbp.so until the process exits).  Limits and rules:
 *   - n must not exceed CBP_NOALLOC_MAX_COLS; larger problems get CBP_ERR_CAPACITY and a message
 *     on stderr (use cbp_matrix_upload + cbp_singletons for those);
 *   - the first call on each device looks up the device address of that array; make that first call
 *     before starting any stream capture on that device;
 *   - all calls on one device share the array (each call clears what it needs), so two calls whose
 *     kernels could overlap on the GPU — different streams, different threads — race: serialise them;
 *   - work is queued on cudaStream (a cudaStream_t, may be 0) and the call returns after it has
 *     finished; nTightened / nFixed are host ints with the meaning they have in cbp_singletons. */
#define CBP_NOALLOC_MAX_COLS 262144
CBP_API int cbp_singletons_noalloc(int m, int n, const int64_t *rowStart, const int *rowLen,
  const int *colIdx, const double *val, const char *rowSense, const double *rhs, const double *range,
  const char *colType, double infinity, double *lb, double *ub, double primalTol, int *nTightened,
  int *nFixed, void *cudaStream);

/* Diagnostic: device timestamps (ns since the loop started, as seen by CTA 0) of the last
 * cbp_propagate: for each of the first rounds, after the row phase, after the long-row phase,
 * after the settle phase and after the stop decision; out gets 4 values per round.  Returns the
 * number of rounds written (at most 64) or a negative error. */
CBP_API int cbp_sv_round_stamps(cbp_matrix *mat, unsigned *out, int maxRounds);

/* Min/max row activities with exact infinite-bound counters for one bound
 * vector (the arithmetic of src/CbcSolverHeuristics.cpp:282-318); rows in the
 * caller's order, skipped ('N' / empty) rows report zeros. */
CBP_API int cbp_row_activity(cbp_matrix *mat, const double *lb, const double *ub,
  double *minAct, double *maxAct, int *infMin, int *infMax, int flags);

#ifdef __cplusplus
}
#endif
#endif
