// fp_peak.cu — FMA-chain microbenchmark: measured FP64 / FP32 vector-pipe peaks of the device the roofline fractions of the
// steering / collision math are quoted against ("the builder must measure them on the box").
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp_peak fp_peak.cu ; prints one JSON line.
#include <cstdio>
#include <cuda_runtime.h>

template <typename T, int ILP>
__global__ void __launch_bounds__(256) k_fma(T* out, T a, T b, int iters) {
  T x[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) x[i] = (T)(threadIdx.x + i);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) x[i] = x[i] * a + b;  // contracted to FMA (default -fmad=true in this file)
  }
  T s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename T>
static double run(int sms, int ctas_per_sm, int iters) {
  constexpr int ILP = 8;
  const int grid = sms * ctas_per_sm, block = 256;
  T* out;
  cudaMalloc(&out, sizeof(T) * grid * block);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  k_fma<T, ILP><<<grid, block>>>(out, (T)1.0000001, (T)1e-9, iters);  // warm-up
  cudaDeviceSynchronize();
  double best = 0;
  for (int rep = 0; rep < 5; ++rep) {
    cudaEventRecord(e0);
    k_fma<T, ILP><<<grid, block>>>(out, (T)1.0000001, (T)1e-9, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    double tf = 2.0 * ILP * (double)iters * grid * block / (ms * 1e-3) / 1e12;
    if (tf > best) best = tf;
  }
  cudaFree(out);
  return best;
}

int main() {
  cudaDeviceProp p;
  if (cudaGetDeviceProperties(&p, 0) != cudaSuccess) { printf("{\"error\": \"no device\"}\n"); return 1; }
  double f64 = run<double>(p.multiProcessorCount, 8, 1 << 14);
  double f32 = run<float>(p.multiProcessorCount, 8, 1 << 16);
  printf("{\"device\": \"%s\", \"sms\": %d, \"fp64_fma_tflops\": %.3f, \"fp32_fma_tflops\": %.3f, \"method\": \"8 independent FMA chains per thread, 8 CTAs x 256 threads per SM, best of 5\"}\n",
         p.name, p.multiProcessorCount, f64, f32);
  return 0;
}
