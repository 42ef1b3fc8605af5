#include "CbpDeviceCache.hpp"

#include <cstdio>
#include <cstring>
#include <vector>

namespace {

// 64-bit multiply-xor hash, 8 bytes per step; not cryptographic, only has to
// tell one problem's arrays from another's
uint64_t hashBytes(const void *data, size_t bytes, uint64_t h)
{
  const unsigned char *p = static_cast<const unsigned char *>(data);
  const uint64_t K = 0x9E3779B97F4A7C15ull;
  size_t i = 0;
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

struct Entry {
  uint64_t key;
  int m, n;
  long long nnz;
  cbp_matrix *mat;
  unsigned long stamp;
};

struct ThreadCache {
  cbp_ctx *ctx = nullptr;
  bool failed = false;
  std::vector<Entry> entries;
  unsigned long clock = 0;
  ~ThreadCache()
  {
    for (Entry &e : entries)
      cbp_matrix_free(e.mat);
    if (ctx)
      cbp_destroy(ctx);
  }
};

thread_local ThreadCache tls;
const size_t kMaxEntries = 4;

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
  const long long nnz = p.m > 0 ? (p.rowLen ? static_cast<long long>(p.rowStart32[p.m - 1]) + p.rowLen[p.m - 1]
                                            : static_cast<long long>(p.rowStart32[p.m]))
                                : 0;
  uint64_t h = 0xCBF29CE484222325ull;
  h = hashBytes(p.rowStart32, sizeof(int) * (p.rowLen ? p.m : p.m + 1), h);
  if (p.rowLen)
    h = hashBytes(p.rowLen, sizeof(int) * p.m, h);
  h = hashBytes(p.colIdx, sizeof(int) * nnz, h);
  h = hashBytes(p.val, sizeof(double) * nnz, h);
  h = hashBytes(p.rowSense, p.m, h);
  h = hashBytes(p.rhs, sizeof(double) * p.m, h);
  if (p.range)
    h = hashBytes(p.range, sizeof(double) * p.m, h);
  if (p.colType)
    h = hashBytes(p.colType, p.n, h);
  h = hashBytes(&p.infinity, sizeof(double), h);
  ++c.clock;
  for (Entry &e : c.entries)
    if (e.key == h && e.m == p.m && e.n == p.n && e.nnz == nnz) {
      e.stamp = c.clock;
      return e.mat;
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
  if (c.entries.size() >= kMaxEntries) {  // evict the least recently used
    size_t victim = 0;
    for (size_t i = 1; i < c.entries.size(); ++i)
      if (c.entries[i].stamp < c.entries[victim].stamp)
        victim = i;
    cbp_matrix_free(c.entries[victim].mat);
    c.entries.erase(c.entries.begin() + victim);
  }
  c.entries.push_back({ h, p.m, p.n, nnz, mat, c.clock });
  return mat;
}

void cbpReleaseThreadCache()
{
  for (Entry &e : tls.entries)
    cbp_matrix_free(e.mat);
  tls.entries.clear();
}
