// Microbenchmark: dependent-load latency through L2 (ld.global.cg pointer chase) and atomic round trips on B200,
// the numbers the k_grow sub-round critical paths are made of (development tool).
#include <cstdio>
#include <cuda_runtime.h>
#include <vector>
#include <algorithm>
#include <random>

__global__ void chase(const int* __restrict__ next, int start, int hops, long long* out, int* sink) {
  int i = start;
  const long long t0 = clock64();
  for (int k = 0; k < hops; ++k) i = __ldcg(&next[i]);
  const long long t1 = clock64();
  if (threadIdx.x == 0) { out[blockIdx.x] = (t1 - t0) / hops; sink[blockIdx.x] = i; }
}
__global__ void atomchase(int* a, int hops, long long* out) {
  int v = 0;
  const long long t0 = clock64();
  for (int k = 0; k < hops; ++k) v = atomicAdd(&a[(v & 1023) * 32], 1) & 1023;
  const long long t1 = clock64();
  if (threadIdx.x == 0) out[blockIdx.x] = (t1 - t0) / hops + (v == -1);
}
__global__ void fencechase(int* a, int hops, long long* out) {
  const long long t0 = clock64();
  for (int k = 0; k < hops; ++k) { a[blockIdx.x * 32 + (k & 7)] = k; __threadfence(); }
  const long long t1 = clock64();
  if (threadIdx.x == 0) out[blockIdx.x] = (t1 - t0) / hops;
}

int main() {
  long long* out; int* sink;
  cudaMalloc(&out, 8 * 256); cudaMalloc(&sink, 4 * 256);
  for (size_t mb : {1, 8, 64, 512}) {
    size_t n = mb * (1 << 20) / 4 / 32;  // one int per 128 B line
    std::vector<int> perm(n); for (size_t i = 0; i < n; ++i) perm[i] = (int)i;
    std::mt19937 g(1); std::shuffle(perm.begin(), perm.end(), g);
    std::vector<int> nx(n * 32, 0);
    for (size_t i = 0; i < n; ++i) nx[(size_t)perm[i] * 32] = perm[(i + 1) % n] * 32;
    int* d; cudaMalloc(&d, nx.size() * 4); cudaMemcpy(d, nx.data(), nx.size() * 4, cudaMemcpyHostToDevice);
    for (int rep = 0; rep < 2; ++rep) {
      for (int ctas : {1, 40}) {
        chase<<<ctas, 32>>>(d, perm[0] * 32, 4000, out, sink);
        cudaDeviceSynchronize();
        long long h[64]; cudaMemcpy(h, out, 8 * ctas, cudaMemcpyDeviceToHost);
        printf("chase %4zu MB  ctas %2d rep %d: %lld cycles / hop\n", mb, ctas, rep, h[0]);
      }
    }
    cudaFree(d);
  }
  int* a; cudaMalloc(&a, 4 * 1024 * 32); cudaMemset(a, 0, 4 * 1024 * 32);
  for (int ctas : {1, 40}) {
    atomchase<<<ctas, 32>>>(a, 2000, out); cudaDeviceSynchronize();
    long long h[64]; cudaMemcpy(h, out, 8 * ctas, cudaMemcpyDeviceToHost);
    printf("atomicAdd round trip ctas %2d: %lld cycles\n", ctas, h[0]);
    fencechase<<<ctas, 32>>>(a, 2000, out); cudaDeviceSynchronize();
    cudaMemcpy(h, out, 8 * ctas, cudaMemcpyDeviceToHost);
    printf("store + __threadfence ctas %2d: %lld cycles\n", ctas, h[0]);
  }
  return 0;
}
