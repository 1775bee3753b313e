// Microbenchmark: one CTA, a burst of independent L2 loads per thread (like the first round trip of k_grow's decide),
// time until all are consumed.  Development tool.
#include <cstdio>
#include <cuda_runtime.h>

__global__ void fill(double* a, size_t n) {
  for (size_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) a[i] = (double)i;
}

template <int NL>
__global__ void burst(const double* __restrict__ a, size_t stride, int active, long long* out, double* sink, int reps) {
  const int t = threadIdx.x;
  long long acc = 0;
  double s = 0.0;
  for (int r = 0; r < reps; ++r) {
    __syncthreads();
    const long long t0 = clock64();
    double v[NL];
    if (t < active) {
#pragma unroll
      for (int k = 0; k < NL; ++k)
        asm volatile("ld.global.cg.f64 %0, [%1];" : "=d"(v[k]) : "l"(&a[k * stride + (size_t)(r & 7) * 4096 + blockIdx.x * 512 + t]));
#pragma unroll
      for (int k = 0; k < NL; ++k) s += v[k];
    }
    if (s == -1.0) sink[0] = s;
    const long long t1 = clock64();
    acc += t1 - t0;
  }
  if (t == 0) out[blockIdx.x] = acc / reps;
  sink[1 + blockIdx.x * 256 + t] = s;
}

int main() {
  for (size_t stride : {(size_t)1 << 17, (size_t)1 << 16}) {  // 8 MB / 1 MB / 64 KB between arrays
  printf("stride %zu KB\n", stride * 8 / 1024);
  double* a; long long* out; double* sink;
  cudaMalloc(&a, 16 * stride * 8); cudaMalloc(&out, 8 * 64); cudaMalloc(&sink, 8 * (1 + 64 * 256));
  fill<<<296, 256>>>(a, 16 * stride); cudaDeviceSynchronize();
  long long h[64];
  for (int ctas : {1, 40}) for (int active : {1, 32, 64, 256}) {
    burst<1><<<ctas, 256>>>(a, stride, active, out, sink, 200); cudaDeviceSynchronize();
    cudaMemcpy(h, out, 8 * ctas, cudaMemcpyDeviceToHost); long long l1 = h[0];
    burst<6><<<ctas, 256>>>(a, stride, active, out, sink, 200); cudaDeviceSynchronize();
    cudaMemcpy(h, out, 8 * ctas, cudaMemcpyDeviceToHost); long long l6 = h[0];
    burst<13><<<ctas, 256>>>(a, stride, active, out, sink, 200); cudaDeviceSynchronize();
    cudaMemcpy(h, out, 8 * ctas, cudaMemcpyDeviceToHost); long long l13 = h[0];
    printf("ctas %2d active threads %3d: 1 load %lld cyc, 6 loads %lld cyc, 13 loads %lld cyc\n", ctas, active, l1, l6, l13);
  }
  cudaFree(a); cudaFree(out); cudaFree(sink);
  }
  return 0;
}
