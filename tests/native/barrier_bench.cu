// Microbenchmark: cycles per grid barrier for the variants considered for k_grow (development tool, B200).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o barrier_bench barrier_bench.cu && ./barrier_bench
#include <cooperative_groups.h>
#include <cstdio>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;

__device__ __forceinline__ unsigned long long ld_acq(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long ld_rlx(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release(unsigned long long* p) {
  asm volatile("red.release.gpu.global.add.u64 [%0], 1;" ::"l"(p) : "memory");
}

template <int V>
__global__ void k(unsigned long long* bar, int iters, long long* out, int n_part, unsigned* marks) {
  cg::grid_group grid = cg::this_grid();
  unsigned long long target = 0;
  const bool part = (int)blockIdx.x < n_part;
  __syncthreads();
  const long long t0 = clock64();
  if (part)
    for (int it = 0; it < iters; ++it) {
      // a little "phase work": a few REDs like the conflict marks
      atomicMax(&marks[(blockIdx.x * 256 + threadIdx.x + it) & 65535], (unsigned)it);
      if (V == 0) {  // fence + atomicAdd + ld.acquire spin + fence
        __syncthreads();
        target += n_part;
        if (threadIdx.x == 0) {
          __threadfence();
          atomicAdd(bar, 1ull);
          while (ld_acq(bar) < target) {}
          __threadfence();
        }
        __syncthreads();
      } else if (V == 1) {  // red.release + ld.acquire spin, no fences
        __syncthreads();
        target += n_part;
        if (threadIdx.x == 0) {
          red_release(bar);
          while (ld_acq(bar) < target) {}
        }
        __syncthreads();
      } else if (V == 2) {
        grid.sync();
      } else if (V == 3) {  // fence + red (no return) + relaxed spin + one fence
        __syncthreads();
        target += n_part;
        if (threadIdx.x == 0) {
          __threadfence();
          asm volatile("red.relaxed.gpu.global.add.u64 [%0], 1;" ::"l"(bar) : "memory");
          while (ld_rlx(bar) < target) {}
          __threadfence();
        }
        __syncthreads();
      }
    }
  const long long t1 = clock64();
  if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
}

template <int V>
void run(const char* name, int n_part, int grid) {
  unsigned long long* bar;
  long long* out;
  unsigned* marks;
  cudaMalloc(&bar, 256);
  cudaMalloc(&out, 8 * grid);
  cudaMalloc(&marks, 4 * 65536);
  cudaMemset(bar, 0, 256);
  cudaMemset(marks, 0, 4 * 65536);
  int iters = 2000;
  void* args[] = {&bar, &iters, &out, &n_part, &marks};
  cudaLaunchCooperativeKernel((const void*)k<V>, dim3(V == 2 ? n_part : grid), dim3(256), args, 0, 0);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[256];
  cudaMemcpy(h, out, 8 * (V == 2 ? n_part : grid), cudaMemcpyDeviceToHost);
  printf("%-40s participants %3d: %8.0f cycles / barrier (%s)\n", name, n_part, (double)h[0] / iters, cudaGetErrorString(e));
  cudaFree(bar); cudaFree(out); cudaFree(marks);
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  const int G = p.multiProcessorCount;
  for (int n : {G, 64, 32, 8}) {
    run<0>("fence+atomicAdd+acq spin+fence", n, G);
    run<1>("red.release + ld.acquire spin", n, G);
    run<3>("fence+red.relaxed+relaxed spin+fence", n, G);
    run<2>("cooperative_groups grid.sync", n, G);
  }
  return 0;
}
