// Cost of N scattered 32-byte slot operations (stores / 64-bit CAS / f64 RED + int max) on a table of S slots, as the
// worklist kernel's pair table sees them.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scatter_bench scatter_bench.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <vector>

struct __align__(32) Slot {
  unsigned long long key;
  double a, b;
  int stamp, pad;
};

__global__ void k_store(Slot *t, const unsigned *idx, unsigned n)
{
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    uint4 *q = reinterpret_cast<uint4 *>(&t[idx[i]]);
    q[0] = make_uint4(0, 0, 0, 0);
    q[1] = make_uint4(0, 0, 0, 0);
  }
}
__global__ void k_cas(Slot *t, const unsigned *idx, unsigned n, unsigned *sink)
{
  unsigned acc = 0;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const unsigned long long seen = atomicCAS(&t[idx[i]].key, 0ull, static_cast<unsigned long long>(i) + 1ull);
    acc += seen == 0ull;
  }
  if (acc == 0xffffffffu)
    *sink = acc;
}
__global__ void k_red(Slot *t, const unsigned *idx, unsigned n, unsigned *sink)
{
  unsigned acc = 0;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    Slot *s = &t[idx[i]];
    atomicAdd(&s->a, 1.0);
    atomicAdd(&s->b, -1.0);
    acc += atomicMax(&s->stamp, 3) < 3;
  }
  if (acc == 0xffffffffu)
    *sink = acc;
}
__global__ void k_ctr(unsigned *ctr, unsigned n, unsigned *sink)  // n warp-level reservations on one counter
{
  unsigned acc = 0;
  const unsigned warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
  for (unsigned i = warp; i < n; i += nw) {
    unsigned base = 0;
    if ((threadIdx.x & 31) == 31)
      base = atomicAdd(ctr, 17u);
    acc += __shfl_sync(0xffffffffu, base, 31);
  }
  if (acc == 0xffffffffu)
    *sink = acc;
}

int main()
{
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int grid = 2 * sms, threads = 256;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  unsigned *sink;
  cudaMalloc(&sink, 256);
  const size_t slotsList[] = { 1u << 19, 1u << 23, 1u << 25 };
  const unsigned nList[] = { 27000u, 1100000u };
  for (size_t slots : slotsList) {
    Slot *t;
    if (cudaMalloc(&t, slots * sizeof(Slot)) != cudaSuccess) {
      printf("alloc failed\n");
      return 1;
    }
    for (unsigned n : nList) {
      std::vector<unsigned> h(n);
      unsigned long long x = 88172645463325252ull;
      for (unsigned i = 0; i < n; ++i) {
        x ^= x << 13;
        x ^= x >> 7;
        x ^= x << 17;
        h[i] = static_cast<unsigned>(x % slots);
      }
      unsigned *idx;
      cudaMalloc(&idx, n * 4);
      cudaMemcpy(idx, h.data(), n * 4, cudaMemcpyHostToDevice);
      for (int which = 0; which < 3; ++which) {
        float best = 1e30f;
        for (int rep = 0; rep < 6; ++rep) {
          cudaMemset(t, 0, slots * sizeof(Slot));
          cudaDeviceSynchronize();
          cudaEventRecord(e0);
          if (which == 0)
            k_store<<<grid, threads>>>(t, idx, n);
          else if (which == 1)
            k_cas<<<grid, threads>>>(t, idx, n, sink);
          else
            k_red<<<grid, threads>>>(t, idx, n, sink);
          cudaEventRecord(e1);
          cudaEventSynchronize(e1);
          float ms;
          cudaEventElapsedTime(&ms, e0, e1);
          if (ms < best)
            best = ms;
        }
        printf("table %4zu MB, %7u ops, %-22s %8.2f us\n", slots * sizeof(Slot) >> 20, n,
          which == 0 ? "32-byte stores" : which == 1 ? "64-bit CAS" : "2 x RED f64 + atomicMax", best * 1000.f);
      }
      cudaFree(idx);
    }
    cudaFree(t);
  }
  unsigned *ctr;
  cudaMalloc(&ctr, 256);
  for (unsigned n : { 128u, 1000u, 4500u, 13500u }) {
    float best = 1e30f;
    for (int rep = 0; rep < 6; ++rep) {
      cudaMemset(ctr, 0, 256);
      cudaDeviceSynchronize();
      cudaEventRecord(e0);
      k_ctr<<<grid, threads>>>(ctr, n, sink);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      if (ms < best)
        best = ms;
    }
    printf("%6u warp reservations on one counter: %8.2f us\n", n, best * 1000.f);
  }
  { // empty kernel for the launch floor
    float best = 1e30f;
    for (int rep = 0; rep < 6; ++rep) {
      cudaDeviceSynchronize();
      cudaEventRecord(e0);
      k_ctr<<<grid, threads>>>(ctr, 0, sink);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      if (ms < best)
        best = ms;
    }
    printf("empty kernel: %8.2f us\n", best * 1000.f);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
