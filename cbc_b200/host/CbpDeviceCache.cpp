#include "CbpDeviceCache.hpp"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace {

// 64-bit multiply-xor hash over four independent lanes (32 bytes per step: the loop runs at memory speed);
// not cryptographic, only has to tell one problem's arrays from another's
uint64_t hashBytes(const void *data, size_t bytes, uint64_t seed)
{
  const unsigned char *p = static_cast<const unsigned char *>(data);
  const uint64_t K = 0x9E3779B97F4A7C15ull;
  uint64_t h0 = seed, h1 = seed ^ K, h2 = seed + K, h3 = ~seed;
  size_t i = 0;
  for (; i + 32 <= bytes; i += 32) {
    uint64_t w[4];
    memcpy(w, p + i, 32);
    h0 = (h0 ^ w[0]) * K;
    h1 = (h1 ^ w[1]) * K;
    h2 = (h2 ^ w[2]) * K;
    h3 = (h3 ^ w[3]) * K;
    h0 ^= h0 >> 29;
    h1 ^= h1 >> 31;
    h2 ^= h2 >> 27;
    h3 ^= h3 >> 33;
  }
  uint64_t h = ((h0 * 3) ^ (h1 * 5) ^ (h2 * 7) ^ (h3 * 11));
  for (; i + 8 <= bytes; i += 8) {
    uint64_t w;
    memcpy(&w, p + i, 8);
    h = (h ^ w) * K;
    h ^= h >> 29;
  }
  uint64_t tail = 0;
  if (i < bytes)
    memcpy(&tail, p + i, bytes - i);
  h = (h ^ tail ^ static_cast<uint64_t>(bytes)) * K;
  return h ^ (h >> 32);
}

// evenly spaced 64-byte pieces of an array, head and tail among them (pieces x 64 bytes however long it is)
uint64_t hashSampled(const void *data, size_t bytes, uint64_t h, size_t pieces = 64)
{
  const size_t piece = 64;
  if (bytes <= piece * pieces)
    return hashBytes(data, bytes, h);
  const unsigned char *p = static_cast<const unsigned char *>(data);
  for (size_t s = 0; s < pieces; ++s) {
    size_t off = (bytes - piece) / (pieces - 1) * s;
    off &= ~static_cast<size_t>(7);
    h = hashBytes(p + off, piece, h);
  }
  return hashBytes(p + bytes - piece, piece, h);
}

struct Entry {
  // where the caller keeps the arrays (the cheap way to recognise "the same matrix again")
  const void *val, *col, *start;
  int m, n;
  long long nnz;
  uint64_t light;  // per-row / per-column arrays in full + samples of the per-nonzero arrays
  uint64_t full;   // every array in full
  cbp_matrix *mat;
  unsigned long stamp;
};

struct ThreadCache {
  cbp_ctx *ctx = nullptr;
  bool failed = false;
  std::vector<Entry> entries;
  unsigned long clock = 0;
  CbpCacheStats stats {};
  ~ThreadCache()
  {
    for (Entry &e : entries)
      cbp_matrix_free(e.mat);
    if (ctx)
      cbp_destroy(ctx);
  }
};

thread_local ThreadCache tls;

size_t maxEntries()
{
  static const size_t n = getenv("CBP_CACHE_ENTRIES") ? static_cast<size_t>(std::max(1, atoi(getenv("CBP_CACHE_ENTRIES")))) : 16;
  return n;
}

long long nnzOf(const CbpProblemView &p)
{
  return p.m > 0 ? (p.rowLen ? static_cast<long long>(p.rowStart32[p.m - 1]) + p.rowLen[p.m - 1]
                             : static_cast<long long>(p.rowStart32[p.m]))
                 : 0;
}

// 0: samples of every array (default); 1: the per-row and per-column arrays in full, samples of the per-nonzero
// arrays; 2: every array in full on every call
int cachePolicy()
{
  static const int v = getenv("CBP_CACHE_FULL_HASH") ? 2 : (getenv("CBP_CACHE_ROW_HASH") ? 1 : 0);
  return v;
}

// what a node call can afford on every call: dimensions, and samples of every array run() reads — 256 pieces of each
// per-row / per-column array (sense, rhs, range, row starts / lengths, column types: what Cbc may edit in place
// between calls), 64 of each per-nonzero array — or those arrays in full (cachePolicy)
uint64_t lightKey(const CbpProblemView &p, long long nnz)
{
  uint64_t h = 0xCBF29CE484222325ull;
  const long long dims[4] = { p.m, p.n, nnz, p.rowLen ? 1 : 0 };
  h = hashBytes(dims, sizeof(dims), h);
  h = hashBytes(&p.infinity, sizeof(double), h);
  const size_t rowPieces = cachePolicy() >= 1 ? ~static_cast<size_t>(0) / 64 : 256;
  h = hashSampled(p.rowStart32, sizeof(int) * (p.rowLen ? p.m : p.m + 1), h, rowPieces);
  if (p.rowLen)
    h = hashSampled(p.rowLen, sizeof(int) * p.m, h, rowPieces);
  h = hashSampled(p.rowSense, p.m, h, rowPieces);
  h = hashSampled(p.rhs, sizeof(double) * p.m, h, rowPieces);
  if (p.range)
    h = hashSampled(p.range, sizeof(double) * p.m, h, rowPieces);
  if (p.colType)
    h = hashSampled(p.colType, p.n, h, rowPieces);
  h = hashSampled(p.colIdx, sizeof(int) * nnz, h);
  h = hashSampled(p.val, sizeof(double) * nnz, h);
  return h;
}

uint64_t fullKey(const CbpProblemView &p, long long nnz, uint64_t light)
{
  uint64_t h = light;
  h = hashBytes(p.rowStart32, sizeof(int) * (p.rowLen ? p.m : p.m + 1), h);
  if (p.rowLen)
    h = hashBytes(p.rowLen, sizeof(int) * p.m, h);
  h = hashBytes(p.rowSense, p.m, h);
  h = hashBytes(p.rhs, sizeof(double) * p.m, h);
  if (p.range)
    h = hashBytes(p.range, sizeof(double) * p.m, h);
  if (p.colType)
    h = hashBytes(p.colType, p.n, h);
  h = hashBytes(p.colIdx, sizeof(int) * nnz, h);
  h = hashBytes(p.val, sizeof(double) * nnz, h);
  return h;
}

}  // namespace

cbp_matrix *cbpAcquireMatrix(const CbpProblemView &p)
{
  ThreadCache &c = tls;
  if (c.failed)
    return nullptr;
  if (!c.ctx) {
    c.ctx = cbp_create(-1);
    if (!c.ctx) {
      fprintf(stderr, "CbcBoundPropagation: no usable CUDA device (%s); bound propagation is skipped\n",
        cbp_last_error(nullptr));
      c.failed = true;
      return nullptr;
    }
  }
  const bool alwaysFull = cachePolicy() == 2;
  const long long nnz = nnzOf(p);
  const uint64_t light = lightKey(p, nnz);
  ++c.clock;
  ++c.stats.calls;
  // 1. the same arrays at the same addresses with the same light key: the matrix of the last call (a node re-presents
  //    its solver's matrix thousands of times, src/CbcModel.cpp:18263-18271)
  if (!alwaysFull)
    for (Entry &e : c.entries)
      if (e.val == p.val && e.col == p.colIdx && e.start == p.rowStart32 && e.m == p.m && e.n == p.n && e.nnz == nnz
        && e.light == light) {
        e.stamp = c.clock;
        ++c.stats.pointerHits;
        return e.mat;
      }
  // 2. new addresses (a solver clone, a rebuilt row copy): every array in full, matched by content
  const uint64_t full = fullKey(p, nnz, light);
  ++c.stats.fullHashes;
  for (Entry &e : c.entries) {
    if (e.m != p.m || e.n != p.n || e.nnz != nnz || e.light != light)
      continue;
    if (e.full == full) {
      e.val = p.val;
      e.col = p.colIdx;
      e.start = p.rowStart32;
      e.stamp = c.clock;
      ++c.stats.contentHits;
      return e.mat;
    }
  }
  std::vector<int64_t> starts(static_cast<size_t>(p.m) + 1, 0);
  for (int i = 0; i < p.m; ++i)
    starts[i] = p.rowStart32[i];
  starts[p.m] = p.rowLen ? nnz : (p.m > 0 ? p.rowStart32[p.m] : 0);
  cbp_matrix *mat = cbp_matrix_upload(c.ctx, p.m, p.n, starts.data(), p.rowLen, p.colIdx, p.val, p.rowSense,
    p.rhs, p.range, p.colType, p.infinity);
  if (!mat) {
    fprintf(stderr, "CbcBoundPropagation: matrix upload failed (%s); bound propagation is skipped\n",
      cbp_last_error(c.ctx));
    return nullptr;
  }
  ++c.stats.uploads;
  if (c.entries.size() >= maxEntries()) {  // evict the least recently used
    size_t victim = 0;
    for (size_t i = 1; i < c.entries.size(); ++i)
      if (c.entries[i].stamp < c.entries[victim].stamp)
        victim = i;
    cbp_matrix_free(c.entries[victim].mat);
    c.entries.erase(c.entries.begin() + victim);
  }
  c.entries.push_back({ p.val, p.colIdx, p.rowStart32, p.m, p.n, nnz, light, full, mat, c.clock });
  return mat;
}

void cbpReleaseThreadCache()
{
  for (Entry &e : tls.entries)
    cbp_matrix_free(e.mat);
  tls.entries.clear();
}

void cbpInvalidateMatrix(const void *elements)
{
  std::vector<Entry> &v = tls.entries;
  for (size_t i = 0; i < v.size();)
    if (!elements || v[i].val == elements) {
      cbp_matrix_free(v[i].mat);
      v.erase(v.begin() + i);
    } else {
      ++i;
    }
}

CbpCacheStats cbpCacheStats() { return tls.stats; }
