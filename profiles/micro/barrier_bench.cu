// Grid-wide barrier cost on one GPU: cooperative_groups grid.sync() against a counter barrier (one release-add per
// CTA, acquire-spin on the same word).  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o barrier_bench barrier_bench.cu
#include <cooperative_groups.h>
#include <cstdio>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;

__device__ __forceinline__ unsigned ld_acquire(const unsigned *p)
{
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release(unsigned *p, unsigned v)
{
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__global__ void k_cg(int n, unsigned *sink)
{
  cg::grid_group g = cg::this_grid();
  unsigned acc = 0;
  for (int i = 0; i < n; ++i) {
    acc += i;
    g.sync();
  }
  if (acc == 0xffffffffu)
    *sink = acc;
}

__global__ void k_ctr(int n, unsigned *count, unsigned *sink)
{
  unsigned target = 0, acc = 0;
  for (int i = 0; i < n; ++i) {
    acc += i;
    __syncthreads();
    if (threadIdx.x == 0) {
      target += gridDim.x;
      red_release(count, 1u);
      while (ld_acquire(count) < target) {
      }
    }
    __syncthreads();
  }
  if (acc == 0xffffffffu)
    *sink = acc;
}

int main()
{
  int dev = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  unsigned *d;
  cudaMalloc(&d, 256);
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  for (int perSm = 1; perSm <= 2; ++perSm)
    for (int threads = 256; threads <= 512; threads *= 2) {
      if (perSm * threads > 1024)
        continue;
      const int grid = perSm * sms;
      for (int which = 0; which < 2; ++which) {
        float best = 1e30f;
        for (int rep = 0; rep < 5; ++rep) {
          int n = 2000;
          cudaMemset(d, 0, 256);
          unsigned *cnt = d, *sink = d + 32;
          void *args0[] = { &n, &sink };
          void *args1[] = { &n, &cnt, &sink };
          cudaEventRecord(a);
          if (which == 0)
            cudaLaunchCooperativeKernel((void *)k_cg, dim3(grid), dim3(threads), args0, 0, 0);
          else
            cudaLaunchCooperativeKernel((void *)k_ctr, dim3(grid), dim3(threads), args1, 0, 0);
          cudaEventRecord(b);
          cudaEventSynchronize(b);
          float ms = 0;
          cudaEventElapsedTime(&ms, a, b);
          if (ms < best)
            best = ms;
        }
        printf("%s grid %d x %d threads: %.3f us per barrier (%s)\n", which ? "counter" : "cg.sync", grid, threads,
          best * 1000.f / 2000.f, cudaGetErrorString(cudaGetLastError()));
      }
    }
  return 0;
}
