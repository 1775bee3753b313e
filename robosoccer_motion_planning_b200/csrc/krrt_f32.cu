// krrt_f32.cu — the binary32 build of the batched planner (k_plan_batch2 + its redo pass k_plan_batch) and of the
// stateless Cost / Environment operators: THE SAME SOURCE as the binary64 library (krrt_math.cuh, krrt_warp.cuh,
// krrt_batch.cuh, krrt_batch2.cuh) compiled with `double` -> `float` (krrt_real.cuh says how and why).  north_star's
// "fp32 mode": float states (8-byte position / velocity pairs, 42 B per node of shared memory against 74), fp32 steering
// and collision tests; results widen to the binary64 records of the C ABI.  Checked bit for bit against the CPU
// checker's float build and within 1e-4 against the binary64 planner on costs and
// arrival times of identical edges (tests/test_f32.py).
// The single-tree planner (k_grow) exists in binary64 only: it is bound by the latency of its commit chain, not by
// arithmetic, and krrt_grow on a KRRT_F32 context returns KRRT_ERR_UNSUPPORTED.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include <algorithm>

#include "krrt_f32_api.h"
#include "krrt_real.cuh"

#define KRRT_F32_BUILD 1
#undef KR
#define KR(x) x##f
#define krrt krrt_f32
#define double float
#define double2 float2
#define double4 float4
#define make_double2 make_float2
#define make_double4 make_float4
#include "krrt_batch2.cuh"

namespace krrt_f32 {

__global__ void __launch_bounds__(128) k_steer_batch(const krrt_io_real* __restrict__ xi, const krrt_io_real* __restrict__ xf,
                                                       int n, double r, krrt_io_real* __restrict__ t,
                                                       krrt_io_real* __restrict__ c) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const krrt_io_real* a = xi + 4 * (size_t)i;
  const krrt_io_real* b = xf + 4 * (size_t)i;
  const State4 x0 = {(double)a[0], (double)a[1], (double)a[2], (double)a[3]};
  const State4 x1 = {(double)b[0], (double)b[1], (double)b[2], (double)b[3]};
  double tau, cost;
  steer(x0, x1, r, tau, cost);
  t[i] = tau;
  c[i] = cost;
}

// one warp per edge: S x m disc tests over the lanes
__global__ void __launch_bounds__(128) k_edge_collide_batch(const krrt_io_real* __restrict__ xi,
                                                              const krrt_io_real* __restrict__ xf,
                                                              const krrt_io_real* __restrict__ tau,
                                                              const krrt_io_real* __restrict__ t0, int n,
                                                              const krrt_io_real* __restrict__ obs, int m, int dynamic,
                                                              double R2, int S, unsigned char* __restrict__ hit) {
  __shared__ double4 s_obs[KRRT_MAX_OBS_DEV];
  for (int j = threadIdx.x; j < m; j += blockDim.x)
    s_obs[j] = make_double4((double)obs[4 * j], (double)obs[4 * j + 1], (double)obs[4 * j + 2], (double)obs[4 * j + 3]);
  __syncthreads();
  const int e = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (e >= n) return;
  const krrt_io_real* a = xi + 4 * (size_t)e;
  const krrt_io_real* b = xf + 4 * (size_t)e;
  const State4 x0 = {(double)a[0], (double)a[1], (double)a[2], (double)a[3]};
  const State4 x1 = {(double)b[0], (double)b[1], (double)b[2], (double)b[3]};
  const EnvView E = {s_obs, m, dynamic, R2, S};
  const bool h = warp_edge_collide(x0, x1, (double)tau[e], (double)t0[e], E);
  if (lane_id() == 0) hit[e] = h ? 1 : 0;
}

// ---- Connector slot in binary32 (the kernels of krrt_ops.cuh with the I/O arrays kept binary64)
__device__ __forceinline__ int solve_last_row(double tau, double dt) {
  int k = (int)floor(tau / dt);
  if (k < 0) k = 0;
  while ((double)k * dt < tau) ++k;
  while (k > 0 && !((double)(k - 1) * dt < tau)) --k;
  return k;
}
__device__ __forceinline__ State4 load_io_state(const krrt_io_real* p) {
  const State4 s = {(double)p[0], (double)p[1], (double)p[2], (double)p[3]};
  return s;
}
__global__ void k_traj_count(const krrt_io_real* __restrict__ states, int n_edges, double r, double dt,
                             krrt_io_real* tau, int* rows) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n_edges) return;
  double t, c;
  steer(load_io_state(states + 4 * (size_t)e), load_io_state(states + 4 * (size_t)(e + 1)), r, t, c);
  tau[e] = t;
  rows[e] = solve_last_row(t, dt) + 1;
}
__global__ void __launch_bounds__(128) k_traj_fill(const krrt_io_real* __restrict__ states, const krrt_io_real* __restrict__ t0s,
                                                     const krrt_io_real* __restrict__ tau, const int* __restrict__ offs,
                                                     double dt, krrt_io_real* out, int cap_rows,
                                                     krrt_io_real* __restrict__ inputs) {
  const int e = blockIdx.x;
  const State4 x0 = load_io_state(states + 4 * (size_t)e), x1 = load_io_state(states + 4 * (size_t)(e + 1));
  const double T = (double)tau[e], t0 = (double)t0s[e];
  const int o = offs[e], cnt = offs[e + 1] - o;
  const Edge ed = make_edge(x0, x1, T, t0);
  for (int k = threadIdx.x; k < cnt; k += blockDim.x) {
    if (o + k >= cap_rows) break;
    const double t = (k == cnt - 1) ? T : (double)k * dt;
    const double h = (T == KR(0.0)) ? KR(0.0) : t / T;
    krrt_io_real* row = out + 5 * (size_t)(o + k);
    row[0] = axis_pos(ed.kx, h);
    row[1] = axis_pos(ed.ky, h);
    row[2] = axis_vel(ed.kx, h, x0.vx, T);
    row[3] = axis_vel(ed.ky, h, x0.vy, T);
    row[4] = t0 + t;
    if (inputs) {
      inputs[2 * (size_t)(o + k) + 0] = axis_acc(ed.kx, h, T);
      inputs[2 * (size_t)(o + k) + 1] = axis_acc(ed.ky, h, T);
    }
  }
}
__global__ void __launch_bounds__(128) k_edge_samples(const krrt_io_real* __restrict__ xi, const krrt_io_real* __restrict__ xf,
                                                        const krrt_io_real* __restrict__ t0, int n, int S, double r,
                                                        krrt_io_real* __restrict__ rows) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= (long long)n * S) return;
  const int e = (int)(g / S), k = (int)(g - (long long)e * S);
  const State4 x0 = load_io_state(xi + 4 * (size_t)e), x1 = load_io_state(xf + 4 * (size_t)e);
  double tau, c;
  steer(x0, x1, r, tau, c);
  const double h = (S > 1) ? (double)k / (double)(S - 1) : KR(0.0);
  const double t0e = (double)t0[e];
  const Edge ed = make_edge(x0, x1, tau, t0e);
  krrt_io_real* row = rows + 5 * g;
  row[0] = axis_pos(ed.kx, h);
  row[1] = axis_pos(ed.ky, h);
  row[2] = axis_vel(ed.kx, h, x0.vx, tau);
  row[3] = axis_vel(ed.ky, h, x0.vy, tau);
  row[4] = t0e + tau * h;
}

static BatchJob make_job(const krrt_f32_api::BatchArgs& a) {
  BatchJob J;
  J.starts = a.starts; J.goals = a.goals; J.obstacles = a.obstacles;
  J.out = reinterpret_cast<ResultRec*>(a.out);
  J.next_query = a.next_query;
  J.list = a.list; J.list_count = a.list_count;
  J.stats = a.stats;
  J.rho2 = a.rho2;
  J.q = a.q; J.m = a.m; J.dynamic = a.dynamic; J.n_iter = a.n_iter; J.cap = a.cap;
  J.query_base = a.query_base;
  const double R = (double)a.R, tol = (double)a.goal_tol;
  J.r = (double)a.r; J.R2 = R * R; J.goal_tol2 = tol * tol; J.S = a.S;
  J.goal_mode = a.goal_mode;
  J.chunk_done = a.chunk_done;
  J.chunk_queries = a.chunk_queries > 0 ? a.chunk_queries : 1;
  for (int k = 0; k < 4; ++k) { J.sp.lo[k] = (double)a.lo[k]; J.sp.hi[k] = (double)a.hi[k]; }
  J.sp.p_direct = (double)a.p_direct;
  J.sp.seed = a.seed;
  return J;
}

}  // namespace krrt_f32

namespace krrt_f32_api {

size_t batch2_warp_smem_bytes(int cap, int m) { return krrt_f32::batch2_warp_smem_bytes(cap, m); }
size_t batch_warp_smem_bytes(int cap, int m) { return krrt_f32::batch_warp_smem_bytes(cap, m); }
int batch2_max_warps() { return krrt_f32::B2_MAX_WARPS; }

int launch_batch2(const BatchArgs& a, int grid, int warps, size_t smem, cudaStream_t st) {
  krrt_f32::Batch2Job J2;
  J2.b = krrt_f32::make_job(a);
  J2.redo_list = a.redo_list;
  J2.redo_count = a.redo_count;
  cudaError_t e = cudaFuncSetAttribute(krrt_f32::k_plan_batch2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  krrt_f32::k_plan_batch2<<<grid, warps * 32, smem, st>>>(J2);
  return (int)cudaGetLastError();
}

int launch_batch(const BatchArgs& a, int grid, int warps, size_t smem, cudaStream_t st) {
  const krrt_f32::BatchJob J = krrt_f32::make_job(a);
  cudaError_t e = cudaFuncSetAttribute(krrt_f32::k_plan_batch, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  krrt_f32::k_plan_batch<<<grid, warps * 32, smem, st>>>(J);
  return (int)cudaGetLastError();
}

// rho(n) = min(rho_max, gamma (ln n / n)^(1/4)), every operation in binary32 (host libm: the CPU checker's float build
// calls the same logf / sqrtf)
krrt_io_real near_radius2(krrt_io_real gamma, krrt_io_real rho_max, int n) {
  if (n < 2) return KR(0.0);
  const double q = logf((double)n) / (double)n;
  double rho = (double)gamma * sqrtf(sqrtf(q));
  if (rho > (double)rho_max) rho = (double)rho_max;
  return rho * rho;
}

int steer_batch(const krrt_io_real* d_xi, const krrt_io_real* d_xf, int n, krrt_io_real r, krrt_io_real* d_t,
                krrt_io_real* d_c, cudaStream_t st) {
  if (n <= 0) return 0;
  krrt_f32::k_steer_batch<<<(n + 127) / 128, 128, 0, st>>>(d_xi, d_xf, n, (double)r, d_t, d_c);
  return (int)cudaGetLastError();
}

int edge_collide_batch(const krrt_io_real* d_xi, const krrt_io_real* d_xf, const krrt_io_real* d_tau,
                       const krrt_io_real* d_t0, int n, const krrt_io_real* d_obs, int m, int dynamic, krrt_io_real R, int S,
                       unsigned char* d_hit, cudaStream_t st) {
  if (n <= 0) return 0;
  const double Rf = (double)R;
  krrt_f32::k_edge_collide_batch<<<(n * 32 + 127) / 128, 128, 0, st>>>(d_xi, d_xf, d_tau, d_t0, n, d_obs, m, dynamic, Rf * Rf, S,
                                                                      d_hit);
  return (int)cudaGetLastError();
}

int traj_count(const krrt_io_real* d_states, int n_edges, krrt_io_real r, krrt_io_real dt, krrt_io_real* d_tau, int* d_rows,
               cudaStream_t st) {
  if (n_edges <= 0) return 0;
  krrt_f32::k_traj_count<<<(n_edges + 127) / 128, 128, 0, st>>>(d_states, n_edges, (double)r, (double)dt, d_tau, d_rows);
  return (int)cudaGetLastError();
}
int traj_fill(const krrt_io_real* d_states, const krrt_io_real* d_t0, const krrt_io_real* d_tau, const int* d_offs, int n_edges,
              krrt_io_real dt, krrt_io_real* d_out, int cap_rows, krrt_io_real* d_inputs, cudaStream_t st) {
  if (n_edges <= 0) return 0;
  krrt_f32::k_traj_fill<<<n_edges, 128, 0, st>>>(d_states, d_t0, d_tau, d_offs, (double)dt, d_out, cap_rows, d_inputs);
  return (int)cudaGetLastError();
}
int edge_samples(const krrt_io_real* d_xi, const krrt_io_real* d_xf, const krrt_io_real* d_t0, int n, int S, krrt_io_real r,
                 krrt_io_real* d_rows, cudaStream_t st) {
  if (n <= 0) return 0;
  const long long threads = (long long)n * S;
  krrt_f32::k_edge_samples<<<(unsigned)((threads + 127) / 128), 128, 0, st>>>(d_xi, d_xf, d_t0, n, S, (double)r, d_rows);
  return (int)cudaGetLastError();
}

}  // namespace krrt_f32_api
