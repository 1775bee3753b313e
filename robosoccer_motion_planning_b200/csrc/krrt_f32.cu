// krrt_f32.cu — the binary32 build of the batched planner (k_plan_batch2 + its redo pass k_plan_batch) and of the
// stateless Cost / Environment operators: THE SAME SOURCE as the binary64 library (krrt_math.cuh, krrt_warp.cuh,
// krrt_batch.cuh, krrt_batch2.cuh) compiled with `double` -> `float` (krrt_real.cuh says how and why).  north_star's
// "fp32 mode": float states (8-byte position / velocity pairs, 42 B per node of shared memory against 74), fp32 steering
// and collision tests; results widen to the binary64 records of the C ABI.  Checked bit for bit against the CPU
// checker's float build and within 1e-4 against the binary64 planner on costs and
// arrival times of identical edges (tests/test_f32.py).
// The single-tree planner too: k_grow and the kernels around it (krrt_grow.cuh, krrt_kernels.cuh) are compiled in float
// from the same source further down, with a small runtime of their own (Tree: arena, stream, control block, logged
// calls, read-outs); krrt_capi.cu forwards the entry points of a KRRT_F32 context here.  Only the development
// diagnostics (phase timers, sub-round and scan statistics) stay binary64.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <string>
#include <vector>

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
#include "krrt_kernels.cuh"
#include "krrt_grow.cuh"

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

// =====================================================================================================================
// The single-tree planner in binary32: k_set_start / k_begin_grow / k_grow / k_goal_connect of namespace krrt_f32 (the
// binary64 sources compiled in float, as everything above) behind a small runtime of their own.  It mirrors what
// krrt_capi.cu does for a binary64 context — one arena, one stream, one cooperative launch per grow call, one read-back of
// the control block — without the diagnostics, the logged call and the single-CTA launch of tiny calls.
// (`double` is `float` in this file: the C ABI's arrays are krrt_io_real.)
namespace krrt_f32 {

__global__ void k_path_extract(TreeView T, int n, int idx, double* states, double* times, int cap, int* out_len) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  if (idx < 0 || idx >= n) {
    *out_len = 0;
    return;
  }
  int len = 0;
  for (int i = idx; i >= 0 && len < n; i = T.parent[i]) ++len;
  int p = len - 1;
  for (int i = idx; i >= 0 && p >= 0; i = T.parent[i], --p) {
    if (p < cap) {
      const double2 a = T.pos[i], v = T.vel[i];
      states[4 * p + 0] = a.x; states[4 * p + 1] = a.y; states[4 * p + 2] = v.x; states[4 * p + 3] = v.y;
      times[p] = T.tim[i];
    }
  }
  *out_len = len;
}

}  // namespace krrt_f32

namespace krrt_f32_api {

using namespace krrt_f32;

struct Tree {
  krrt_config cfg;
  cudaStream_t st = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  int sms = 0, grow_grid = 0;
  unsigned char* arena = nullptr;
  size_t arena_bytes = 0, arena_used = 0;
  TreeView T{};
  double* d_rho2 = nullptr;
  double4* d_obs = nullptr;
  Ctl* d_ctl = nullptr;
  Ctl* h_ctl = nullptr;
  bool ctl_valid = false, started = false;
  int n_upper = 0;
  Batch Bt{};
  Wave W{};
  Params P{};
  GridBar* d_bar = nullptr;
  GridIndex gi{};
  GoalScratch goal_scr{};
  double* d_user = nullptr;
  size_t user_cap = 0;
  double* d_scr = nullptr;   // path read-out
  int* d_len = nullptr;
  size_t scr_cap = 0;
  unsigned char* d_iterlog = nullptr;  // logged calls: the IterLog views + the device records (one allocation)
  IterLog ilog = {nullptr, nullptr, nullptr, nullptr, nullptr, 0, 0};
  IterRec* d_iterrec = nullptr;
  int wave_min_n = 32;
  int* d_near = nullptr;     // tree_near: [count, nearest, d2, pad] + the index list; grows, never shrinks
  size_t near_cap = 0;       // ints
  int64_t launches = 0, last_iters = 0;
  int last_full = 0;
  float last_ms = 0.f;
  std::string err;
};

#define TCK(call)                                                             \
  do {                                                                        \
    cudaError_t e_ = (call);                                                  \
    if (e_ != cudaSuccess) {                                                  \
      t->err = std::string(#call) + ": " + cudaGetErrorString(e_);            \
      return KRRT_ERR_CUDA;                                                   \
    }                                                                         \
  } while (0)
#define TREQ(cond, code, msg) \
  do {                        \
    if (!(cond)) {            \
      t->err = (msg);         \
      return (code);          \
    }                         \
  } while (0)

template <typename Tp>
static cudaError_t talloc(Tree* t, Tp** p, size_t count) {
  const size_t bytes = (sizeof(Tp) * std::max<size_t>(count, 1) + 255) / 256 * 256;
  if (t->arena_used + bytes > t->arena_bytes) return cudaErrorMemoryAllocation;
  *p = reinterpret_cast<Tp*>(t->arena + t->arena_used);
  t->arena_used += bytes;
  return cudaSuccess;
}

static void fill_params(Tree* t) {
  const krrt_config& c = t->cfg;
  Params& P = t->P;
  const double R = (double)c.R, tol = (double)c.goal_tol;
  P.r = (double)c.r;
  P.R2 = R * R;
  P.S = c.S;
  P.capacity = c.capacity;
  for (int k = 0; k < 4; ++k) {
    P.sp.lo[k] = (double)c.lo[k];
    P.sp.hi[k] = (double)c.hi[k];
  }
  P.sp.p_direct = (double)c.p_direct;
  P.sp.seed = c.seed;
  P.query_id = c.query_id;
  P.goal_tol2 = tol * tol;
  P.goal_mode = c.goal_mode;
  P.rho2 = t->d_rho2;
  P.obs = t->d_obs;
}

static int upload_radius_table(Tree* t) {
  const size_t cap = (size_t)t->cfg.capacity;
  std::vector<double> rho2(cap + 1);
  for (size_t n = 0; n <= cap; ++n) rho2[n] = (double)near_radius2(t->cfg.gamma, t->cfg.rho_max, (int)n);
  TCK(cudaMemcpyAsync(t->d_rho2, rho2.data(), sizeof(double) * (cap + 1), cudaMemcpyHostToDevice, t->st));
  TCK(cudaStreamSynchronize(t->st));
  return KRRT_OK;
}

static int create_body(Tree* t) {
  const krrt_config& c = t->cfg;
  TCK(cudaSetDevice(c.device));
  cudaDeviceProp prop;
  TCK(cudaGetDeviceProperties(&prop, c.device));
  t->sms = prop.multiProcessorCount;
  TCK(cudaStreamCreateWithFlags(&t->st, cudaStreamNonBlocking));
  TCK(cudaEventCreate(&t->ev0));
  TCK(cudaEventCreate(&t->ev1));
  const size_t cap = (size_t)c.capacity;
  const size_t bcn = (size_t)c.batch * c.near_cap;
  const size_t need = cap * (8 * 2 + 4 * 4 + 8 + 4 * 3 + 4 * 3) + cap * (8 * 2 + 4 * 2) + cap * 8 + (GRID_CELLS_MAX + 1) * 4 +
                      (cap + 1) * 4 + bcn * (4 + 4 * 5) + (size_t)1024 * GROW_QCAP * 4 + 2 * 1024 * 4 + 512 +
                      (size_t)c.batch * (16 + 1 + 1 + 4 + 4 + sizeof(WaveRec)) + sizeof(Ctl) + GOAL_MAX_CTAS * 12 + (64 << 10);
  TCK(cudaMalloc((void**)&t->arena, need));
  t->arena_bytes = need;
  TreeView& T = t->T;
  TCK(talloc(t, &T.pos, cap + 1));
  TCK(talloc(t, &T.vel, cap + 1));
  TCK(talloc(t, &T.ctc, cap + 1));
  TCK(talloc(t, &T.tim, cap + 1));
  TCK(talloc(t, &T.ecost, cap + 1));
  TCK(talloc(t, &T.etime, cap + 1));
  TCK(talloc(t, &T.parent, cap + 1));
  TCK(talloc(t, &T.link, cap));
  TCK(talloc(t, &T.prev, cap));
  TCK(talloc(t, &T.queue, cap));
  T.cap = c.capacity;
  TCK(talloc(t, &t->d_rho2, cap + 1));
  TCK(talloc(t, &t->d_obs, (size_t)KRRT_MAX_OBS));
  TCK(talloc(t, &t->d_ctl, (size_t)1));
  TCK(talloc(t, &t->goal_scr.part_c, (size_t)GOAL_MAX_CTAS));
  TCK(talloc(t, &t->goal_scr.part_i, (size_t)GOAL_MAX_CTAS));
  TCK(talloc(t, &t->goal_scr.ticket, (size_t)1));
  TCK(cudaMemsetAsync(t->goal_scr.ticket, 0, sizeof(unsigned), t->st));
  TCK(cudaMemsetAsync(t->d_ctl, 0, sizeof(Ctl), t->st));
  TCK(cudaMemsetAsync(t->d_obs, 0, sizeof(double4) * KRRT_MAX_OBS, t->st));
  Batch& B = t->Bt;
  B.B = c.batch;
  B.near_cap = c.near_cap;
  TCK(talloc(t, &B.samples, (size_t)B.B));
  TCK(talloc(t, &B.reject, (size_t)B.B));
  TCK(talloc(t, &B.dup, (size_t)B.B));
  TCK(talloc(t, &B.map, (size_t)B.B));
  TCK(talloc(t, &B.count, (size_t)B.B));
  TCK(talloc(t, &B.ref, bcn));
  TCK(talloc(t, &B.d2, bcn));
  TCK(talloc(t, &B.tf, bcn));
  TCK(talloc(t, &B.cf, bcn));
  TCK(talloc(t, &B.tb, bcn));
  TCK(talloc(t, &B.cb, bcn));
  Wave& W = t->W;
  TCK(talloc(t, &W.rec, (size_t)B.B));
  TCK(talloc(t, &W.wmin, cap));
  TCK(talloc(t, &W.amin, cap));
  TCK(talloc(t, &W.rmin, cap));
  TCK(talloc(t, &W.lock, cap));
  TCK(talloc(t, &W.wlc, (size_t)1024 * GROW_QCAP));
  TCK(talloc(t, &W.done, (size_t)1024));
  TCK(talloc(t, &W.ok, (size_t)1024));
  TCK(talloc(t, &W.nn_d2, (size_t)1024));
  TCK(talloc(t, &W.nn_idx, (size_t)1024));
  W.dbg_log = nullptr;
  TCK(cudaMemsetAsync(W.done, 0, sizeof(unsigned) * 1024, t->st));
  TCK(cudaMemsetAsync(W.ok, 0, sizeof(unsigned) * 1024, t->st));
  TCK(cudaMemsetAsync(W.wmin, 0, sizeof(unsigned) * cap, t->st));
  TCK(cudaMemsetAsync(W.amin, 0, sizeof(unsigned) * cap, t->st));
  TCK(cudaMemsetAsync(W.rmin, 0, sizeof(unsigned long long) * cap, t->st));
  TCK(cudaMemsetAsync(W.lock, 0, sizeof(int) * cap, t->st));
  TCK(talloc(t, &t->gi.pos, cap));
  TCK(talloc(t, &t->gi.vel, cap));
  TCK(talloc(t, &t->gi.id, cap));
  TCK(talloc(t, &t->gi.off, cap));
  TCK(talloc(t, &t->gi.cell, (size_t)GRID_CELLS_MAX + 1));
  t->gi.enabled = 1;
  t->gi.min_n = 8192;
  t->gi.tail_max = 8192;
  if (const char* e = std::getenv("KRRT_GRID_INDEX")) t->gi.enabled = std::atoi(e) != 0;
  if (const char* e = std::getenv("KRRT_WAVE_MIN_N")) t->wave_min_n = std::max(2, std::atoi(e));
  TCK(talloc(t, &t->d_bar, (size_t)2));
  {
    const int dyn = (int)(grow_shared_bytes() + commit_smem_bytes(B.near_cap));
    TCK(cudaFuncSetAttribute(k_grow, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn));
    TCK(cudaFuncSetAttribute(k_grow_logged, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn));
    int per_sm = 0, coop = 0;
    TCK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_grow, GROW_THREADS, dyn));
    TCK(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, c.device));
    TREQ(per_sm >= 1 && coop, KRRT_ERR_UNSUPPORTED, "k_grow (binary32) does not fit one CTA per SM / no cooperative launch");
    t->grow_grid = t->sms;
  }
  TCK(cudaHostAlloc((void**)&t->h_ctl, sizeof(Ctl), cudaHostAllocDefault));
  int rc = upload_radius_table(t);
  if (rc != KRRT_OK) return rc;
  fill_params(t);
  t->P.m = 0;
  t->P.dynamic = 0;
  return KRRT_OK;
}

int tree_create(const krrt_config* cfg, Tree** out, char* err, size_t err_len) {
  Tree* t = new Tree();
  t->cfg = *cfg;
  const int rc = create_body(t);
  if (rc != KRRT_OK) {
    if (err && err_len) snprintf(err, err_len, "%s", t->err.c_str());
    tree_destroy(t);
    return rc;
  }
  *out = t;
  return KRRT_OK;
}

void tree_destroy(Tree* t) {
  if (!t) return;
  cudaSetDevice(t->cfg.device);
  if (t->st) cudaStreamSynchronize(t->st);
  if (t->arena) cudaFree(t->arena);
  if (t->d_user) cudaFree(t->d_user);
  if (t->d_scr) cudaFree(t->d_scr);
  if (t->d_near) cudaFree(t->d_near);
  if (t->d_iterlog) cudaFree(t->d_iterlog);
  if (t->h_ctl) cudaFreeHost(t->h_ctl);
  if (t->ev0) cudaEventDestroy(t->ev0);
  if (t->ev1) cudaEventDestroy(t->ev1);
  if (t->st) cudaStreamDestroy(t->st);
  delete t;
}

const char* tree_error(const Tree* t) { return t ? t->err.c_str() : ""; }
int64_t tree_launches(const Tree* t) { return t ? t->launches : 0; }

int tree_set_params(Tree* t, const krrt_config* cfg) {
  TCK(cudaSetDevice(t->cfg.device));
  const bool radius = cfg->gamma != t->cfg.gamma || cfg->rho_max != t->cfg.rho_max;
  t->cfg = *cfg;
  if (radius) {
    const int rc = upload_radius_table(t);
    if (rc != KRRT_OK) return rc;
  }
  const int m = t->P.m, dyn = t->P.dynamic;
  fill_params(t);
  t->P.m = m;
  t->P.dynamic = dyn;
  return KRRT_OK;
}

int tree_set_obstacles(Tree* t, const krrt_io_real* o, int m, int dynamic) {
  TCK(cudaSetDevice(t->cfg.device));
  if (m > 0) {
    std::vector<double> h(4 * (size_t)m);
    for (size_t k = 0; k < h.size(); ++k) h[k] = (double)o[k];
    TCK(cudaMemcpyAsync(t->d_obs, h.data(), sizeof(double) * h.size(), cudaMemcpyHostToDevice, t->st));
    TCK(cudaStreamSynchronize(t->st));
  }
  t->P.m = m;
  t->P.dynamic = dynamic ? 1 : 0;
  return KRRT_OK;
}

int tree_set_start(Tree* t, const krrt_io_real x0[4]) {
  TCK(cudaSetDevice(t->cfg.device));
  const State4 s = {(double)x0[0], (double)x0[1], (double)x0[2], (double)x0[3]};
  k_set_start<<<1, 32, 0, t->st>>>(t->d_ctl, t->T, s);
  t->launches++;
  TCK(cudaGetLastError());
  t->started = true;
  t->ctl_valid = false;
  t->n_upper = 1;
  return KRRT_OK;
}

static int fetch(Tree* t, Ctl* h) {
  if (!t->ctl_valid) {
    TCK(cudaMemcpyAsync(t->h_ctl, t->d_ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, t->st));
    TCK(cudaStreamSynchronize(t->st));
    t->ctl_valid = true;
  }
  *h = *t->h_ctl;
  t->n_upper = h->n;
  return KRRT_OK;
}

static int grow_body(Tree* t, const krrt_io_real* goal, int n_iter, const krrt_io_real* samples, int* solved, int* goal_index,
                     krrt_iter_record* records);

int tree_grow(Tree* t, const krrt_io_real* goal, int n_iter, const krrt_io_real* samples, int* solved, int* goal_index) {
  return grow_body(t, goal, n_iter, samples, solved, goal_index, nullptr);
}

constexpr int F32_LOG_MAX_ITERS = 256, F32_LOG_STRIDE = 1024;

int tree_grow_logged(Tree* t, const krrt_io_real* goal, int n_iter, const krrt_io_real* samples, krrt_iter_record* records,
                     int* solved, int* goal_index) {
  TREQ(records != nullptr, KRRT_ERR_INVALID, "null records");
  TREQ(n_iter >= 1 && n_iter <= F32_LOG_MAX_ITERS, KRRT_ERR_UNSUPPORTED, "krrt_grow_logged: 1 <= n_iter <= 256");
  TREQ(t->started, KRRT_ERR_STATE, "krrt_grow_logged before krrt_set_start");
  TREQ(t->n_upper + n_iter <= F32_LOG_STRIDE, KRRT_ERR_UNSUPPORTED,
       "krrt_grow_logged: the tree would outgrow the 1024 nodes a logged call can record; use krrt_grow");
  return grow_body(t, goal, n_iter, samples, solved, goal_index, records);
}

static int grow_body(Tree* t, const krrt_io_real* goal, int n_iter, const krrt_io_real* samples, int* solved, int* goal_index,
                     krrt_iter_record* records) {
  TREQ(t->started, KRRT_ERR_STATE, "krrt_grow before krrt_set_start");
  TREQ(n_iter >= 0, KRRT_ERR_INVALID, "n_iter < 0");
  TCK(cudaSetDevice(t->cfg.device));
  cudaStream_t st = t->st;
  if (samples && n_iter > 0) {  // explicit rows, rounded to binary32 (a NaN row stays one: "draw it yourself")
    const size_t cnt = 4 * (size_t)n_iter, need = sizeof(double) * cnt;
    if (t->user_cap < need) {
      if (t->d_user) TCK(cudaFree(t->d_user));
      t->d_user = nullptr;
      t->user_cap = 0;
      TCK(cudaMalloc((void**)&t->d_user, need));
      t->user_cap = need;
    }
    std::vector<double> h(cnt);
    for (size_t k = 0; k < cnt; ++k) h[k] = (double)samples[k];
    TCK(cudaMemcpyAsync(t->d_user, h.data(), need, cudaMemcpyHostToDevice, st));
    TCK(cudaStreamSynchronize(st));
  }
  State4 g = {0, 0, 0, 0};
  if (goal) g = State4{(double)goal[0], (double)goal[1], (double)goal[2], (double)goal[3]};
  TCK(cudaEventRecord(t->ev0, st));
  GrowCfg gcfg;
  gcfg.wave_enabled = 1;
  if (const char* e = std::getenv("KRRT_WAVE")) gcfg.wave_enabled = std::atoi(e) != 0;
  gcfg.wave_min_n = t->wave_min_n;
  gcfg.user_samples = samples ? t->d_user : nullptr;
  gcfg.bar = t->d_bar;
  gcfg.gi = t->gi;
  gcfg.log = IterLog{nullptr, nullptr, nullptr, nullptr, nullptr, 0, 0};
  const int n_before = t->n_upper;
  if (records) {
    if (!t->d_iterlog) {
      const size_t rows = F32_LOG_MAX_ITERS + 1, it = F32_LOG_MAX_ITERS;
      const size_t b_snap = rows * F32_LOG_STRIDE * sizeof(double), b_smp = it * 4 * sizeof(double), b_int = it * sizeof(int),
                   b_rec = (it * sizeof(IterRec) + 15) / 16 * 16;
      TCK(cudaMalloc((void**)&t->d_iterlog, 2 * b_snap + b_smp + b_rec + 2 * b_int));
      TCK(cudaMemsetAsync(t->d_iterlog, 0, 2 * b_snap + b_smp + b_rec + 2 * b_int, st));  // no log entry is ever an uninitialised index
      unsigned char* q = t->d_iterlog;
      t->ilog.snap_ctc = (double*)q; q += b_snap;
      t->ilog.snap_tim = (double*)q; q += b_snap;
      t->ilog.sample = (double*)q; q += b_smp;
      t->d_iterrec = (IterRec*)q; q += b_rec;
      t->ilog.n_after = (int*)q; q += b_int;
      t->ilog.snap_of = (int*)q;
      t->ilog.stride = F32_LOG_STRIDE;
      t->ilog.rows = (int)rows;
    }
    gcfg.log = t->ilog;
    gcfg.wave_enabled = 0;  // every iteration goes through the in-order commit, which keeps the records
  }
  {
    auto envi = [](const char* k, int d) { const char* e = std::getenv(k); return e ? std::atoi(e) : d; };
    gcfg.pend_resolve = envi("KRRT_PEND", 1);
    gcfg.raw_forward = envi("KRRT_RAW_FORWARD", 1);
    gcfg.tail_seq = envi("KRRT_TAIL_SEQ", 1);
    gcfg.intra_fix = envi("KRRT_INTRA_FIX", 1);
    gcfg.prop_forward = envi("KRRT_PROP_FORWARD", 1);
    gcfg.join_test = envi("KRRT_JOIN_TEST", 1);
    gcfg.flow = envi("KRRT_FLOW", 0);
    gcfg.nb_mul = envi("KRRT_NB_MUL", 2); gcfg.nb_add = envi("KRRT_NB_ADD", 8); gcfg.nb_min = envi("KRRT_NB_MIN", 16);
    gcfg.pn_mul = envi("KRRT_PN_MUL", 3); gcfg.pn_add = envi("KRRT_PN_ADD", 16); gcfg.pn_min = envi("KRRT_PN_MIN", 48);
  }
  const int dyn = (int)(grow_shared_bytes() + commit_smem_bytes(t->Bt.near_cap));
  Ctl h;
  for (int launch = 0;; ++launch) {
    const bool wrap = launch > 0;
    if (wrap) {
      const size_t cap = (size_t)t->cfg.capacity;
      TCK(cudaMemsetAsync(t->W.wmin, 0, sizeof(unsigned) * cap, st));
      TCK(cudaMemsetAsync(t->W.amin, 0, sizeof(unsigned) * cap, st));
      TCK(cudaMemsetAsync(t->W.rmin, 0, sizeof(unsigned long long) * cap, st));
      TCK(cudaMemsetAsync(t->W.done, 0, sizeof(unsigned) * 1024, st));
      TCK(cudaMemsetAsync(t->W.ok, 0, sizeof(unsigned) * 1024, st));
    }
    t->ctl_valid = false;
    k_begin_grow<<<1, 32, 0, st>>>(t->d_ctl, launch == 0, goal ? 1 : 0, g, n_iter, t->Bt, t->d_bar, wrap ? 1 : 0, t->T, gcfg.log);
    void* args[] = {(void*)&t->d_ctl, (void*)&t->P, (void*)&t->T, (void*)&t->Bt, (void*)&t->W, (void*)&gcfg};
    if (records && t->n_upper + n_iter <= 512)  // (as in binary64: a small logged call is one CTA, plain launch)
      k_grow_logged<<<1, GROW_THREADS, (size_t)dyn, st>>>(t->d_ctl, t->P, t->T, t->Bt, t->W, gcfg);
    else
      TCK(cudaLaunchCooperativeKernel(records ? (const void*)k_grow_logged : (const void*)k_grow, dim3(t->grow_grid),
                                      dim3(GROW_THREADS), args, (size_t)dyn, st));
    if (t->P.goal_mode == KRRT_GOAL_TOLERANCE) {
      k_find_goal<<<1, 1024, 0, st>>>(t->d_ctl, t->T, t->P.goal_tol2);
    } else {
      const long long n_max = (long long)t->n_upper + n_iter;
      const int ctas = (int)std::max<long long>(1, std::min<long long>(std::min(GOAL_MAX_CTAS, 4 * t->sms),
                                                                       (n_max + GOAL_THREADS - 1) / GOAL_THREADS));
      k_goal_connect<<<ctas, GOAL_THREADS, 0, st>>>(t->d_ctl, t->P, t->T, t->goal_scr);
    }
    if (records) {
      k_goal_log<<<n_iter, GLOG_THREADS, 0, st>>>(t->d_ctl, t->P, t->T, gcfg.log, n_iter, n_before, t->d_iterrec);
      t->launches += 1;
    }
    TCK(cudaEventRecord(t->ev1, st));
    t->launches += 3;
    TCK(cudaGetLastError());
    const int rc = fetch(t, &h);
    if (rc != KRRT_OK) return rc;
    if (h.full || h.error || h.iter >= h.iter_target) break;
    TREQ(h.round >= (1u << 23) - 2, KRRT_ERR_CUDA, "k_grow (binary32) returned before its target");
  }
  TCK(cudaEventElapsedTime(&t->last_ms, t->ev0, t->ev1));
  t->last_iters = h.iter - h.iter_begin;
  t->last_full = h.full ? 1 : 0;
  if (h.error) {
    t->err = h.error == KRRT_ERR_NEAR_OVERFLOW ? "near set exceeded near_cap; raise krrt_config.near_cap" : "device error";
    return h.error;
  }
  if (records) {  // the device records are float: widen into the C ABI's layout
    std::vector<IterRec> hr((size_t)n_iter);
    TCK(cudaMemcpyAsync(hr.data(), t->d_iterrec, sizeof(IterRec) * (size_t)n_iter, cudaMemcpyDeviceToHost, st));
    TCK(cudaStreamSynchronize(st));
    for (int i = 0; i < n_iter; ++i) {
      records[i].n_nodes = hr[i].n_nodes;
      records[i].solved = hr[i].solved;
      records[i].goal_index = hr[i].goal_index;
      records[i].inserted = hr[i].inserted;
      for (int k = 0; k < 4; ++k) records[i].sample[k] = hr[i].sample[k];
      records[i].goal_cost = hr[i].goal_cost;
    }
  }
  if (solved) *solved = h.goal_index >= 0;
  if (goal_index) *goal_index = h.goal_index;
  return KRRT_OK;
}

int tree_grow_done(Tree* t, int64_t* iterations_done, int* full, krrt_io_real* ms) {
  if (iterations_done) *iterations_done = t->last_iters;
  if (full) *full = t->last_full;
  if (ms) *ms = t->last_ms;
  return KRRT_OK;
}

int tree_near(Tree* t, const krrt_io_real x[4], krrt_io_real radius2, int32_t* idx, int cap, int* count, int32_t* nearest,
              krrt_io_real* nearest_d2) {
  TREQ(t->started, KRRT_ERR_STATE, "no tree: call krrt_set_start");
  TREQ(cap >= 0 && (cap == 0 || idx), KRRT_ERR_INVALID, "null buffer");
  TCK(cudaSetDevice(t->cfg.device));
  Ctl h;
  int rc = fetch(t, &h);
  if (rc != KRRT_OK) return rc;
  const size_t need = 4 + (size_t)std::max(cap, 1);  // [count, nearest, d2 (float), pad] + the index list
  if (t->near_cap < need) {
    if (t->d_near) TCK(cudaFree(t->d_near));
    t->d_near = nullptr;
    t->near_cap = 0;
    TCK(cudaMalloc((void**)&t->d_near, sizeof(int) * need));
    t->near_cap = need;
  }
  int* d_buf = t->d_near;
  const State4 s = {(double)x[0], (double)x[1], (double)x[2], (double)x[3]};
  k_near_query<<<1, 1024, 0, t->st>>>(t->T, h.n, s, (double)radius2, d_buf + 4, cap, d_buf, d_buf + 1, (double*)(d_buf + 2));
  t->launches++;
  struct { int cnt, nearest; double nd2; int pad; } r;
  cudaError_t e = cudaMemcpyAsync(&r, d_buf, sizeof(r), cudaMemcpyDeviceToHost, t->st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(t->st);
  const int got = e == cudaSuccess ? std::min(r.cnt, cap) : 0;
  if (e == cudaSuccess && got > 0) e = cudaMemcpy(idx, d_buf + 4, sizeof(int) * (size_t)got, cudaMemcpyDeviceToHost);
  TCK(e);
  if (count) *count = r.cnt;
  if (nearest) *nearest = r.nearest;
  if (nearest_d2) *nearest_d2 = r.nd2;
  return KRRT_OK;
}

int tree_info(Tree* t, int* n, int64_t* iteration, krrt_io_real last_sample[4], krrt_stats* stats, int* goal_index) {
  TREQ(t->started, KRRT_ERR_STATE, "no tree: call krrt_set_start");
  TCK(cudaSetDevice(t->cfg.device));
  Ctl h;
  const int rc = fetch(t, &h);
  if (rc != KRRT_OK) return rc;
  if (n) *n = h.n;
  if (iteration) *iteration = h.iter;
  if (last_sample)
    for (int k = 0; k < 4; ++k) last_sample[k] = h.last_sample[k];
  if (stats) {
    static_assert(sizeof(krrt_stats) == 10 * sizeof(long long), "stats layout");
    memcpy(stats, h.stats, sizeof(krrt_stats));
  }
  if (goal_index) *goal_index = h.goal_index;
  return KRRT_OK;
}

int tree_read(Tree* t, int first, int count, krrt_io_real* states, int32_t* parent, krrt_io_real* cost, krrt_io_real* time) {
  TREQ(t->started, KRRT_ERR_STATE, "no tree: call krrt_set_start");
  TCK(cudaSetDevice(t->cfg.device));
  Ctl h;
  int rc = fetch(t, &h);
  if (rc != KRRT_OK) return rc;
  const int ng = h.n + (h.goal_index == h.n ? 1 : 0);
  TREQ(first >= 0 && count >= 0 && first + count <= ng, KRRT_ERR_INVALID, "node range out of bounds");
  if (count == 0) return KRRT_OK;
  const size_t c = (size_t)count;
  std::vector<double2> p(states ? c : 0), v(states ? c : 0);
  std::vector<double> cc(cost ? c : 0), tt(time ? c : 0);
  if (states) {
    TCK(cudaMemcpyAsync(p.data(), t->T.pos + first, sizeof(double2) * c, cudaMemcpyDeviceToHost, t->st));
    TCK(cudaMemcpyAsync(v.data(), t->T.vel + first, sizeof(double2) * c, cudaMemcpyDeviceToHost, t->st));
  }
  if (parent) TCK(cudaMemcpyAsync(parent, t->T.parent + first, sizeof(int) * c, cudaMemcpyDeviceToHost, t->st));
  if (cost) TCK(cudaMemcpyAsync(cc.data(), t->T.ctc + first, sizeof(double) * c, cudaMemcpyDeviceToHost, t->st));
  if (time) TCK(cudaMemcpyAsync(tt.data(), t->T.tim + first, sizeof(double) * c, cudaMemcpyDeviceToHost, t->st));
  TCK(cudaStreamSynchronize(t->st));
  for (size_t i = 0; i < c; ++i) {
    if (states) {
      states[4 * i + 0] = p[i].x; states[4 * i + 1] = p[i].y; states[4 * i + 2] = v[i].x; states[4 * i + 3] = v[i].y;
    }
    if (cost) cost[i] = cc[i];
    if (time) time[i] = tt[i];
  }
  return KRRT_OK;
}

int tree_path(Tree* t, int idx, krrt_io_real* states, krrt_io_real* times, int cap, int* len) {
  TREQ(t->started, KRRT_ERR_STATE, "no tree: call krrt_set_start");
  TCK(cudaSetDevice(t->cfg.device));
  Ctl h;
  int rc = fetch(t, &h);
  if (rc != KRRT_OK) return rc;
  const int ng = h.n + (h.goal_index == h.n ? 1 : 0);
  TREQ(idx >= 0 && idx < ng, KRRT_ERR_INVALID, "node index out of range");
  const size_t need = sizeof(double) * 5 * (size_t)ng + 16;
  if (t->scr_cap < need) {
    if (t->d_scr) TCK(cudaFree(t->d_scr));
    t->d_scr = nullptr;
    t->scr_cap = 0;
    TCK(cudaMalloc((void**)&t->d_scr, need));
    t->scr_cap = need;
  }
  double* d_states = t->d_scr;
  double* d_times = d_states + 4 * (size_t)ng;
  int* d_len = reinterpret_cast<int*>(d_times + ng);
  k_path_extract<<<1, 32, 0, t->st>>>(t->T, ng, idx, d_states, d_times, ng, d_len);
  t->launches++;
  int hl = 0;
  TCK(cudaMemcpyAsync(&hl, d_len, sizeof(int), cudaMemcpyDeviceToHost, t->st));
  TCK(cudaStreamSynchronize(t->st));
  *len = hl;
  const int got = std::min(hl, cap);
  if (got > 0 && (states || times)) {
    std::vector<double> hs(4 * (size_t)got), ht((size_t)got);
    TCK(cudaMemcpyAsync(hs.data(), d_states, sizeof(double) * 4 * got, cudaMemcpyDeviceToHost, t->st));
    TCK(cudaMemcpyAsync(ht.data(), d_times, sizeof(double) * got, cudaMemcpyDeviceToHost, t->st));
    TCK(cudaStreamSynchronize(t->st));
    if (states)
      for (size_t k = 0; k < hs.size(); ++k) states[k] = hs[k];
    if (times)
      for (size_t k = 0; k < ht.size(); ++k) times[k] = ht[k];
  }
  return KRRT_OK;
}

}  // namespace krrt_f32_api
