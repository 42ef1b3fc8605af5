// CbpDeviceCache — per-thread device context plus a small cache of uploaded
// matrices for the C++ drop-ins.  run() is re-entered from several Cbc worker
// threads on distinct solver clones (/root/reference/src/CbcThread.cpp:1249,1319)
// and node-level calls present the same matrix thousands of times
// (src/CbcModel.cpp:18263-18271), so every thread owns one cbp_ctx (one CUDA
// stream) and keeps the last matrices it uploaded (16; CBP_CACHE_ENTRIES).
// A call finds its matrix again in two steps ("matrix pointer +
// nnz + a cheap fingerprint"):
//  1. same array addresses, same dimensions, same LIGHT key = evenly spaced
//     64-byte samples of every array run() reads (256 of each per-row /
//     per-column array, 64 of each per-nonzero array, head and tail included):
//     a few microseconds, what a node call pays.
//  2. otherwise (a solver clone, a rebuilt row copy) every array is hashed in
//     full and the entry is found by content, its addresses updated.
// A value changed IN PLACE between two calls, at an address the samples miss,
// is not noticed by step 1.  Osi implementations rebuild their cached row copy
// and row sense / rhs / range arrays when the model changes, which changes the
// addresses; a caller that edits such arrays in place calls
// cbpInvalidateMatrix() at that point, or runs with
// CBP_CACHE_ROW_HASH=1 (per-row and per-column arrays hashed in full on every
// call) or CBP_CACHE_FULL_HASH=1 (every array, every call).
// A node's matrix with cut rows appended is a different matrix here (full
// upload): appending the new rows' tiles to the device copy is not implemented.
#ifndef CbpDeviceCache_hpp
#define CbpDeviceCache_hpp

#include <cstdint>

#include "cbp_capi.h"

struct CbpProblemView {
  int m, n;
  const int *rowStart32;  // CoinBigIndex is int in this build
  const int *rowLen;
  const int *colIdx;
  const double *val;
  const char *rowSense;
  const double *rhs;
  const double *range;
  const char *colType;
  double infinity;
};

// Returns the device copy of `p` for the calling thread (uploading it when the
// fingerprint is new) or nullptr if no device / upload failed (reason printed
// to stderr once).
cbp_matrix *cbpAcquireMatrix(const CbpProblemView &p);

// Drops every cached matrix of the calling thread (tests use it).
void cbpReleaseThreadCache();

// Drops the calling thread's device copy of the matrix whose element array is at `elements`
// (nullptr: all of them): for callers that change coefficients in place.
void cbpInvalidateMatrix(const void *elements);

// How the calling thread's calls found their matrix (benchmarks, tests).
struct CbpCacheStats {
  long long calls, pointerHits, contentHits, fullHashes, uploads;
};
CbpCacheStats cbpCacheStats();

#endif
