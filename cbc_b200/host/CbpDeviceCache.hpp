// CbpDeviceCache — per-thread device context plus a small cache of uploaded
// matrices for the C++ drop-ins.  run() is re-entered from several Cbc worker
// threads on distinct solver clones (/root/reference/src/CbcThread.cpp:1249,1319)
// and node-level calls present the same matrix thousands of times
// (src/CbcModel.cpp:18263-18271), so every thread owns one cbp_ctx (one CUDA
// stream) and keeps the last few matrices it uploaded, found again by a
// fingerprint of the problem data (dimensions + hashes of every array run()
// reads once per call, src/CbcBoundPropagation.cpp:91-99).
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

#endif
