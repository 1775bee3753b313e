// Microbenchmark (development tool, B200): read bandwidth of an L2-resident buffer, the roofline of k_grow's neighbour scan
// (the sorted copy of a <= 1 M-node tree is 32 MB: it lives in the 126 MB L2, DRAM sees almost nothing of the scan).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o l2_peak l2_peak.cu && ./l2_peak      (prints one JSON line)
// 16-byte ld.global.cg loads (the scan's own access: __ldcg of double2), 8 in flight per thread, every CTA walks the
// whole buffer with a CTA-specific phase so that all SMs hit all L2 slices; best of 5 per size.
#include <cstdio>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(256) k_read(const double2* __restrict__ buf, size_t n, int passes, double* sink) {
  double acc = 0.0;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (int p = 0; p < passes; ++p) {
    size_t i = ((size_t)blockIdx.x * 7919u % gridDim.x) * blockDim.x + threadIdx.x;
    for (; i < n; i += 8 * stride) {
      double2 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = (i + u * stride < n) ? __ldcg(&buf[i + u * stride]) : make_double2(0.0, 0.0);
#pragma unroll
      for (int u = 0; u < 8; ++u) acc += v[u].x + v[u].y;
    }
  }
  if (acc == 123.456) *sink = acc;
}

int main() {
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, 0);
  double best_all = 0.0;
  std::printf("{\"device\": \"%s\", \"l2_bytes\": %d, \"sizes\": [", prop.name, prop.l2CacheSize);
  const size_t sizes_mb[] = {8, 16, 32, 64};
  for (int s = 0; s < 4; ++s) {
    const size_t bytes = sizes_mb[s] << 20, n = bytes / sizeof(double2);
    double2* buf;
    double* sink;
    cudaMalloc(&buf, bytes);
    cudaMalloc(&sink, 8);
    cudaMemset(buf, 0, bytes);
    const int grid = prop.multiProcessorCount * 8, passes = 20;
    k_read<<<grid, 256>>>(buf, n, 2, sink);  // warm the L2
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    double best = 0.0;
    for (int rep = 0; rep < 5; ++rep) {
      cudaEventRecord(e0);
      k_read<<<grid, 256>>>(buf, n, passes, sink);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms = 0.f;
      cudaEventElapsedTime(&ms, e0, e1);
      const size_t per_pass = n * sizeof(double2);
      const double gbs = (double)per_pass * passes / (ms * 1e-3) / 1e9;
      if (gbs > best) best = gbs;
    }
    std::printf("%s{\"mb\": %zu, \"read_GBps\": %.1f}", s ? ", " : "", sizes_mb[s], best);
    if (sizes_mb[s] == 32) best_all = best;
    cudaFree(buf);
    cudaFree(sink);
  }
  std::printf("], \"l2_read_GBps\": %.1f, \"method\": \"16-byte ld.global.cg, 8 in flight per thread, 8 CTAs x 256 threads per SM, "
              "32 MB buffer (the sorted copy of a 1 M-node tree), best of 5\"}\n", best_all);
  return cudaGetLastError() == cudaSuccess ? 0 : 1;
}
