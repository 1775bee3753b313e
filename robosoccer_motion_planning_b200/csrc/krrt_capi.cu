// krrt_capi.cu — host runtime behind include/krrt.h: owns device memory, one stream per context, launches the
// sm_100a kernels.  No CPU fallback: when no Blackwell device is present krrt_create fails with
// KRRT_ERR_NO_DEVICE and nothing else works.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/krrt.h"
#include "krrt_batch.cuh"
#include "krrt_batch2.cuh"
#include "krrt_kernels.cuh"
#include "krrt_grow.cuh"
#include "krrt_ops.cuh"
#include "krrt_yaw.cuh"
#include "krrt_f32_api.h"

using namespace krrt;

static_assert(sizeof(ResultRec) == sizeof(krrt_result), "result layout");
static_assert(sizeof(Stats) == sizeof(krrt_stats), "stats layout");
static_assert(KRRT_MAX_OBS_DEV == KRRT_MAX_OBS, "obstacle cap");

static thread_local std::string g_create_error;

struct krrt_ctx {
  krrt_config cfg;
  cudaStream_t stream = nullptr;
  cudaDeviceProp prop;
  std::string err;
  int64_t launches = 0;
  bool started = false;
  int n_upper = 0;  // host-side upper bound of the tree size (sizes the scan grid without a sync)

  TreeView T{};
  double* d_rho2 = nullptr;
  double4* d_obs = nullptr;
  Ctl* d_ctl = nullptr;
  Batch Bt{};
  Wave W{};
  Params P{};
  int wave_enabled = 1;       // KRRT_WAVE=0 in the environment forces the in-order commit (debugging / A-B timing)
  int wave_min_n = 32;        // trees smaller than this commit in order (conflicts everywhere; profiles/r1_param_sweep.txt)
  int small_max_n = 256;      // few-expansion calls on trees below this: one CTA, plain launch, in-order commit
  int grow_grid = 0;          // CTAs of the persistent k_grow (KRRT_GROW_MINB per SM, all co-resident)
  GridBar* d_bar = nullptr;    // its two grid barriers
  GridIndex gi{};              // spatial index of its neighbour scan
  GoalScratch goal_scr{};      // partials of k_goal_connect
  double phase_ms[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  long long phase_n[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  double* d_user = nullptr;
  size_t user_cap = 0;
  void* d_flush = nullptr;     // krrt_flush_l2
  size_t flush_bytes = 0;
  unsigned char* arena = nullptr;
  size_t arena_bytes = 0, arena_used = 0;

  // scratch for stateless operators / batched planner
  void* d_scratch[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  size_t scratch_cap[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  double* d_rho2_batch = nullptr;
  int* d_redo = nullptr;      // k_plan_batch2: [0] work counter of the redo pass, [1] redo count, [2..] redo list
  size_t redo_cap = 0;
  Ctl* h_ctl = nullptr;       // pinned host copy of the control block ...
  bool ctl_valid = false;     // ... current until the next call that changes the tree (krrt_set_start, krrt_grow)
  int batch_v1 = 0;           // KRRT_BATCH_V1=1 in the environment: the shared-memory-list kernel only (A-B timing)
  int b2_wave_fit = 1;        // KRRT_B2_WAVE_FIT=0: k_plan_batch2 with the most warps per CTA that fit even for a small batch (A-B timing)
  int rho2_batch_cap = -1, rho2_batch_alloc = 0;
  std::vector<double> h_rho2_batch;
  int* d_counter = nullptr;
  long long* d_batch_stats = nullptr;

  // timing
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  cudaEvent_t tm0 = nullptr, tm1 = nullptr;
  bool profile = false;
  double last_total_ms = 0.0, last_scan_ms = 0.0;
  int64_t last_scan_launches = 0;
  double last_scan_node_samples = 0.0;
  unsigned char* d_iterlog = nullptr;  // krrt_grow_logged: one allocation behind the IterLog views + the records
  IterLog ilog = {nullptr, nullptr, nullptr, nullptr, nullptr, 0, 0};
  IterRec* d_iterrec = nullptr;
  krrt_f32_api::Tree* f32 = nullptr;  // KRRT_F32 contexts: the single-tree planner in binary32 (krrt_f32.cu)
  int64_t last_grow_iters = 0;  // iterations the last krrt_grow call really ran (fewer than asked when the tree filled up)
  int last_grow_full = 0;
};

#define CK(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      ctx->err = std::string(#call) + ": " + cudaGetErrorString(e_);                               \
      return KRRT_ERR_CUDA;                                                                        \
    }                                                                                              \
  } while (0)

#define REQUIRE(cond, code, msg) \
  do {                           \
    if (!(cond)) {               \
      ctx->err = (msg);          \
      return (code);             \
    }                            \
  } while (0)
// entry points that exist in binary64 only: the development diagnostics (phase timers, sub-round statistics, scan statistics).
// A KRRT_F32 context serves everything else — the planner proper incl. the logged call, the tree read-outs, the batched
// planner and the stateless operators (krrt_f32.cu)
#define F64_ONLY()                                                                                                    \
  REQUIRE(ctx->cfg.precision == KRRT_F64, KRRT_ERR_UNSUPPORTED,                                                       \
          "not available on a KRRT_F32 context (the development diagnostics are binary64 only)")

// KRRT_F32 contexts: the call goes to the binary32 planner; its error text becomes the context's
#define F32_FORWARD(expr)                                                \
  do {                                                                   \
    if (ctx->cfg.precision == KRRT_F32) {                                \
      const int rc_ = (expr);                                            \
      if (rc_ != KRRT_OK) ctx->err = krrt_f32_api::tree_error(ctx->f32); \
      return rc_;                                                        \
    }                                                                    \
  } while (0)

// Everything a context owns on the device comes out of ONE allocation: the sub-rounds of k_grow are chains of
// dependent gathers over ~25 arrays, and one arena keeps them inside a handful of 2 MB pages (TLB reach).
template <typename Tp>
static cudaError_t dev_alloc(krrt_ctx* ctx, Tp** p, size_t count) {
  const size_t bytes = (sizeof(Tp) * std::max<size_t>(count, 1) + 255) / 256 * 256;
  if (ctx->arena_used + bytes > ctx->arena_bytes) return cudaErrorMemoryAllocation;
  *p = reinterpret_cast<Tp*>(ctx->arena + ctx->arena_used);
  ctx->arena_used += bytes;
  return cudaSuccess;
}

static int ensure_scratch(krrt_ctx* ctx, int slot, size_t bytes) {
  if (ctx->scratch_cap[slot] >= bytes) return KRRT_OK;
  if (ctx->d_scratch[slot]) CK(cudaFree(ctx->d_scratch[slot]));
  ctx->d_scratch[slot] = nullptr;
  ctx->scratch_cap[slot] = 0;
  size_t want = std::max<size_t>(bytes, 4096);
  CK(cudaMalloc(&ctx->d_scratch[slot], want));
  ctx->scratch_cap[slot] = want;
  return KRRT_OK;
}

// squared near radius rho(n)^2, rho(n) = min(rho_max, gamma (ln n / n)^(1/4)); 0 for n < 2 (NeighborRadius slot)
static double near_radius2(const krrt_config& c, int n) {
  if (n < 2) return 0.0;
  double q = std::log((double)n) / (double)n;
  double rho = c.gamma * std::sqrt(std::sqrt(q));
  if (rho > c.rho_max) rho = c.rho_max;
  return rho * rho;
}

extern "C" {

void krrt_default_config(krrt_config* c) {
  std::memset(c, 0, sizeof(*c));
  c->device = 0;
  c->precision = KRRT_F64;
  c->capacity = 1 << 16;
  c->S = 10;
  c->r = 0.5;    // config/config.yaml:1
  c->R = 1.35;   // config/config.yaml:5
  c->lo[0] = -9.0; c->hi[0] = 9.0;   // field 18 x 12 m (global_config.yaml:17-18)
  c->lo[1] = -6.0; c->hi[1] = 6.0;
  c->lo[2] = -2.0; c->hi[2] = 2.0;
  c->lo[3] = -2.0; c->hi[3] = 2.0;
  c->gamma = 7.0;
  c->rho_max = 4.0;
  c->p_direct = 0.0;
  c->goal_tol = 0.0;
  c->seed = 0;
  c->query_id = 0;
  c->batch = 0;
  c->near_cap = 0;
  c->goal_mode = KRRT_GOAL_CONNECT;
}

int krrt_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
  return n;
}

const char* krrt_last_error(const krrt_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

// (a KRRT_F32 context's single-tree kernels are counted where they are launched: krrt_f32.cu's Tree)
int64_t krrt_launch_count(const krrt_ctx* ctx) {
  return ctx ? ctx->launches + (ctx->f32 ? krrt_f32_api::tree_launches(ctx->f32) : 0) : 0;
}

void krrt_destroy(krrt_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->cfg.device);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  if (ctx->arena) cudaFree(ctx->arena);
  for (void* p : ctx->d_scratch)
    if (p) cudaFree(p);
  if (ctx->d_user) cudaFree(ctx->d_user);
  if (ctx->d_iterlog) cudaFree(ctx->d_iterlog);
  if (ctx->f32) krrt_f32_api::tree_destroy(ctx->f32);
  if (ctx->d_flush) cudaFree(ctx->d_flush);
  if (ctx->d_rho2_batch) cudaFree(ctx->d_rho2_batch);
  if (ctx->d_redo) cudaFree(ctx->d_redo);
  if (ctx->h_ctl) cudaFreeHost(ctx->h_ctl);
  if (ctx->ev0) cudaEventDestroy(ctx->ev0);
  if (ctx->tm0) cudaEventDestroy(ctx->tm0);
  if (ctx->tm1) cudaEventDestroy(ctx->tm1);
  if (ctx->ev1) cudaEventDestroy(ctx->ev1);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
}

static int create_impl(krrt_ctx* ctx) {
  const krrt_config& c = ctx->cfg;
  REQUIRE(c.precision == KRRT_F64 || c.precision == KRRT_F32, KRRT_ERR_INVALID, "precision is KRRT_F64 or KRRT_F32");
  REQUIRE(c.capacity >= 2 && c.S >= 1 && c.r > 0.0 && c.R >= 0.0, KRRT_ERR_INVALID, "bad capacity / S / r / R");
  REQUIRE(c.batch >= 1 && c.batch <= GROW_WMAX, KRRT_ERR_INVALID, "batch must be in [1,1024]");
  REQUIRE(c.near_cap >= 32 && c.near_cap <= 4096, KRRT_ERR_INVALID, "near_cap must be in [32,4096]");
  REQUIRE(c.goal_mode == KRRT_GOAL_CONNECT || c.goal_mode == KRRT_GOAL_TOLERANCE, KRRT_ERR_INVALID, "bad goal_mode");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0 || c.device < 0 || c.device >= ndev) {
    ctx->err = std::string("no usable CUDA device (") + (e != cudaSuccess ? cudaGetErrorString(e) : "ordinal out of range") +
               "); libkrrt has no CPU fallback";
    return KRRT_ERR_NO_DEVICE;
  }
  CK(cudaSetDevice(c.device));
  CK(cudaGetDeviceProperties(&ctx->prop, c.device));
  if (ctx->prop.major != 10) {
    ctx->err = std::string("device ") + ctx->prop.name + " is not sm_100; libkrrt is built for sm_100a only";
    return KRRT_ERR_NO_DEVICE;
  }
  CK(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
  CK(cudaEventCreate(&ctx->ev0));
  CK(cudaEventCreate(&ctx->ev1));
  CK(cudaEventCreate(&ctx->tm0));
  CK(cudaEventCreate(&ctx->tm1));
  const size_t cap = (size_t)c.capacity;
  {
    const size_t bcn = (size_t)c.batch * c.near_cap;
    size_t need = cap * (16 * 2 + 8 * 4 + 8 + 4 * 3 + 4 * 3) + cap * (16 * 2 + 4 * 2) + cap * 8 + (GRID_CELLS_MAX + 1) * 4 +
                  (cap + 1) * 8 + bcn * (4 + 8 * 5) + (size_t)1024 * GROW_QCAP * 8 + 2 * 1024 * 4 + 512 +
                  (size_t)c.batch * (32 + 1 + 1 + 4 + 4 + sizeof(WaveRec)) + sizeof(Ctl) + GOAL_MAX_CTAS * 12 + (64 << 10);
#ifdef KRRT_DEBUG_CLK
    need += (size_t)32 * ((1 << 20) + 1) + 4096;
#endif
    CK(cudaMalloc((void**)&ctx->arena, need));
    ctx->arena_bytes = need;
    ctx->arena_used = 0;
  }
  CK(dev_alloc(ctx, &ctx->T.pos, cap + 1));  // + the goal node's slot (k_goal_connect)
  CK(dev_alloc(ctx, &ctx->T.vel, cap + 1));
  CK(dev_alloc(ctx, &ctx->T.ctc, cap + 1));
  CK(dev_alloc(ctx, &ctx->T.tim, cap + 1));
  CK(dev_alloc(ctx, &ctx->T.ecost, cap + 1));
  CK(dev_alloc(ctx, &ctx->T.etime, cap + 1));
  CK(dev_alloc(ctx, &ctx->T.parent, cap + 1));
  CK(dev_alloc(ctx, &ctx->T.link, cap));
  CK(dev_alloc(ctx, &ctx->T.prev, cap));
  CK(dev_alloc(ctx, &ctx->T.queue, cap));
  ctx->T.cap = c.capacity;
  CK(dev_alloc(ctx, &ctx->d_rho2, cap + 1));
  CK(dev_alloc(ctx, &ctx->d_obs, (size_t)KRRT_MAX_OBS));
  CK(dev_alloc(ctx, &ctx->d_ctl, (size_t)1));
  CK(dev_alloc(ctx, &ctx->d_counter, (size_t)1));
  CK(dev_alloc(ctx, &ctx->d_batch_stats, (size_t)10));
  CK(dev_alloc(ctx, &ctx->goal_scr.part_c, (size_t)GOAL_MAX_CTAS));
  CK(dev_alloc(ctx, &ctx->goal_scr.part_i, (size_t)GOAL_MAX_CTAS));
  CK(dev_alloc(ctx, &ctx->goal_scr.ticket, (size_t)1));
  CK(cudaMemsetAsync(ctx->goal_scr.ticket, 0, sizeof(unsigned), ctx->stream));
  CK(cudaMemsetAsync(ctx->d_ctl, 0, sizeof(Ctl), ctx->stream));
  CK(cudaMemsetAsync(ctx->d_obs, 0, sizeof(double4) * KRRT_MAX_OBS, ctx->stream));
  {
    std::vector<double> rho2(cap + 1);
    for (size_t n = 0; n <= cap; ++n) rho2[n] = near_radius2(c, (int)n);
    CK(cudaMemcpyAsync(ctx->d_rho2, rho2.data(), sizeof(double) * (cap + 1), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
  }
  Batch& B = ctx->Bt;
  B.B = c.batch;
  B.near_cap = c.near_cap;
  const size_t bc = (size_t)B.B * B.near_cap;
  CK(dev_alloc(ctx, &B.samples, (size_t)B.B));
  CK(dev_alloc(ctx, &B.reject, (size_t)B.B));
  CK(dev_alloc(ctx, &B.dup, (size_t)B.B));
  CK(dev_alloc(ctx, &B.map, (size_t)B.B));
  CK(dev_alloc(ctx, &B.count, (size_t)B.B));
  CK(dev_alloc(ctx, &B.ref, bc));
  CK(dev_alloc(ctx, &B.d2, bc));
  CK(dev_alloc(ctx, &B.tf, bc));
  CK(dev_alloc(ctx, &B.cf, bc));
  CK(dev_alloc(ctx, &B.tb, bc));
  CK(dev_alloc(ctx, &B.cb, bc));
  Wave& W = ctx->W;
  CK(dev_alloc(ctx, &W.rec, (size_t)B.B));
  CK(dev_alloc(ctx, &W.wmin, cap));
  CK(dev_alloc(ctx, &W.amin, cap));
  CK(dev_alloc(ctx, &W.rmin, cap));
  CK(dev_alloc(ctx, &W.lock, cap));
  CK(dev_alloc(ctx, &W.wlc, (size_t)1024 * GROW_QCAP));
  CK(dev_alloc(ctx, &W.done, (size_t)1024));
  CK(dev_alloc(ctx, &W.ok, (size_t)1024));
  CK(cudaMemsetAsync(W.done, 0, sizeof(unsigned) * 1024, ctx->stream));
  CK(cudaMemsetAsync(W.ok, 0, sizeof(unsigned) * 1024, ctx->stream));
  CK(dev_alloc(ctx, &ctx->gi.pos, cap));
  CK(dev_alloc(ctx, &ctx->gi.vel, cap));
  CK(dev_alloc(ctx, &ctx->gi.id, cap));
  CK(dev_alloc(ctx, &ctx->gi.off, cap));
  CK(dev_alloc(ctx, &ctx->gi.cell, (size_t)GRID_CELLS_MAX + 1));
  ctx->gi.enabled = 1;
  ctx->gi.min_n = 8192;
  if (const char* e = std::getenv("KRRT_GRID_INDEX")) ctx->gi.enabled = std::atoi(e) != 0;
  if (const char* e = std::getenv("KRRT_GRID_MIN_N")) ctx->gi.min_n = std::max(2, std::atoi(e));
  ctx->gi.tail_max = 8192;  // measured 2048 .. unbounded on the 1 M-iteration tree: 392 / 383 / 381 / 384 / 392 / 393 ms (profiles/r2_grid_tail_s9.txt)
  if (const char* e = std::getenv("KRRT_GRID_TAIL")) ctx->gi.tail_max = std::max(1024, std::atoi(e));
  CK(dev_alloc(ctx, &W.nn_d2, (size_t)1024));
  CK(dev_alloc(ctx, &W.nn_idx, (size_t)1024));
#ifdef KRRT_DEBUG_CLK
  CK(dev_alloc(ctx, &W.dbg_log, (size_t)8 * ((1 << 20) + 1)));
  CK(cudaMemsetAsync(W.dbg_log, 0, sizeof(int) * 8, ctx->stream));
#endif
  CK(cudaMemsetAsync(W.wmin, 0, sizeof(unsigned) * cap, ctx->stream));
  CK(cudaMemsetAsync(W.amin, 0, sizeof(unsigned) * cap, ctx->stream));
  CK(cudaMemsetAsync(W.rmin, 0, sizeof(unsigned long long) * cap, ctx->stream));
  CK(cudaMemsetAsync(W.lock, 0, sizeof(int) * cap, ctx->stream));
  CK(dev_alloc(ctx, &ctx->d_bar, (size_t)2));
  if (const char* e = std::getenv("KRRT_WAVE")) ctx->wave_enabled = std::atoi(e) != 0;
  if (const char* e = std::getenv("KRRT_BATCH_V1")) ctx->batch_v1 = std::atoi(e) != 0;
  if (const char* e = std::getenv("KRRT_B2_WAVE_FIT")) ctx->b2_wave_fit = std::atoi(e) != 0;
  if (const char* e = std::getenv("KRRT_WAVE_MIN_N")) ctx->wave_min_n = std::max(2, std::atoi(e));
  {
    const int dyn = (int)(grow_shared_bytes() + commit_smem_bytes(B.near_cap));
    CK(cudaFuncSetAttribute(k_grow, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn));
    CK(cudaFuncSetAttribute(k_grow_logged, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn));
    int per_sm = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_grow, GROW_THREADS, dyn));
    int coop = 0;
    CK(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, c.device));
    REQUIRE(per_sm >= 1 && coop, KRRT_ERR_UNSUPPORTED, "k_grow does not fit one CTA per SM / no cooperative launch");
    // every co-resident CTA the kernel's footprint allows (KRRT_GROW_MINB per SM: 128 registers, ~110 KB of shared memory):
    // a sub-round decides one sample per CTA, so twice the CTAs is twice the prefix a sub-round can commit
    int per = std::min(per_sm, KRRT_GROW_MINB);
    if (const char* e = std::getenv("KRRT_GROW_PER_SM")) per = std::max(1, std::min(per, std::atoi(e)));
    ctx->grow_grid = per * ctx->prop.multiProcessorCount;
    if (const char* e = std::getenv("KRRT_GRID")) ctx->grow_grid = std::max(2, std::min(ctx->grow_grid, std::atoi(e)));
  }
  Params& P = ctx->P;
  P.r = c.r;
  P.R2 = c.R * c.R;
  P.S = c.S;
  P.capacity = c.capacity;
  for (int k = 0; k < 4; ++k) {
    P.sp.lo[k] = c.lo[k];
    P.sp.hi[k] = c.hi[k];
  }
  P.sp.p_direct = c.p_direct;
  P.sp.seed = c.seed;
  P.query_id = c.query_id;
  P.m = 0;
  P.dynamic = 0;
  P.goal_tol2 = c.goal_tol * c.goal_tol;
  P.goal_mode = c.goal_mode;
  P.rho2 = ctx->d_rho2;
  P.obs = ctx->d_obs;
  if (c.precision == KRRT_F32) {
    char msg[512] = {0};
    const int rc = krrt_f32_api::tree_create(&ctx->cfg, &ctx->f32, msg, sizeof(msg));
    if (rc != KRRT_OK) {
      ctx->err = msg;
      return rc;
    }
  }
  return KRRT_OK;
}

int krrt_create(const krrt_config* c, krrt_ctx** out) {
  if (!c || !out) {
    g_create_error = "null argument";
    return KRRT_ERR_INVALID;
  }
  *out = nullptr;
  krrt_ctx* ctx = new krrt_ctx();
  ctx->cfg = *c;
  if (ctx->cfg.batch == 0) ctx->cfg.batch = 1024;   // profiles/r1_param_sweep.txt
  if (const char* e = std::getenv("KRRT_WINDOW")) ctx->cfg.batch = std::max(1, std::min(GROW_WMAX, std::atoi(e)));
  if (ctx->cfg.near_cap == 0) ctx->cfg.near_cap = 1024;
  int rc = create_impl(ctx);
  if (rc != KRRT_OK) {
    g_create_error = ctx->err;
    krrt_destroy(ctx);
    return rc;
  }
  *out = ctx;
  return KRRT_OK;
}

int krrt_host_alloc(void** p, uint64_t bytes) {
  if (!p) return KRRT_ERR_INVALID;
  *p = nullptr;
  return cudaHostAlloc(p, bytes > 0 ? (size_t)bytes : 1, cudaHostAllocDefault) == cudaSuccess ? KRRT_OK : KRRT_ERR_CUDA;
}

int krrt_host_free(void* p) { return (!p || cudaFreeHost(p) == cudaSuccess) ? KRRT_OK : KRRT_ERR_CUDA; }

int krrt_flush_l2(krrt_ctx* ctx) {
  if (!ctx) return KRRT_ERR_INVALID;
  CK(cudaSetDevice(ctx->cfg.device));
  if (!ctx->d_flush) {
    ctx->flush_bytes = std::max<size_t>(2 * (size_t)ctx->prop.l2CacheSize, (size_t)256 << 20);
    CK(cudaMalloc(&ctx->d_flush, ctx->flush_bytes));
  }
  CK(cudaMemsetAsync(ctx->d_flush, 0, ctx->flush_bytes, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return KRRT_OK;
}

int krrt_synchronize(krrt_ctx* ctx) {
  if (!ctx) return KRRT_ERR_INVALID;
  CK(cudaSetDevice(ctx->cfg.device));
  CK(cudaStreamSynchronize(ctx->stream));
  return KRRT_OK;
}

int krrt_timer_begin(krrt_ctx* ctx) {
  if (!ctx) return KRRT_ERR_INVALID;
  CK(cudaSetDevice(ctx->cfg.device));
  CK(cudaEventRecord(ctx->tm0, ctx->stream));
  return KRRT_OK;
}

int krrt_timer_end(krrt_ctx* ctx, double* ms) {
  if (!ctx || !ms) return KRRT_ERR_INVALID;
  CK(cudaSetDevice(ctx->cfg.device));
  CK(cudaEventRecord(ctx->tm1, ctx->stream));
  CK(cudaEventSynchronize(ctx->tm1));
  float t = 0.f;
  CK(cudaEventElapsedTime(&t, ctx->tm0, ctx->tm1));
  *ms = t;
  return KRRT_OK;
}

int krrt_set_profile(krrt_ctx* ctx, int on) {
  if (!ctx) return KRRT_ERR_INVALID;
  ctx->profile = on != 0;
  return KRRT_OK;
}

int krrt_set_obstacles(krrt_ctx* ctx, const double* o, int m, int dynamic) {
  if (!ctx) return KRRT_ERR_INVALID;
  REQUIRE(m >= 0 && m <= KRRT_MAX_OBS && (m == 0 || o), KRRT_ERR_INVALID, "0 <= m <= KRRT_MAX_OBS");
  CK(cudaSetDevice(ctx->cfg.device));
  if (m > 0) {
    CK(cudaMemcpyAsync(ctx->d_obs, o, sizeof(double) * 4 * m, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));  // the caller's buffer may be reused right after return
  }
  ctx->P.m = m;
  ctx->P.dynamic = dynamic ? 1 : 0;
  if (ctx->f32) {  // (the binary64 copy above serves krrt_edge_collide's float kernel, which rounds on load)
    const int rc = krrt_f32_api::tree_set_obstacles(ctx->f32, o, m, dynamic);
    if (rc != KRRT_OK) ctx->err = krrt_f32_api::tree_error(ctx->f32);
    return rc;
  }
  return KRRT_OK;
}

int krrt_set_start(krrt_ctx* ctx, const double x0[4]) {
  if (!ctx || !x0) return KRRT_ERR_INVALID;
  F32_FORWARD(krrt_f32_api::tree_set_start(ctx->f32, x0));
  CK(cudaSetDevice(ctx->cfg.device));
  State4 s = {x0[0], x0[1], x0[2], x0[3]};
  ctx->ctl_valid = false;
  k_set_start<<<1, 32, 0, ctx->stream>>>(ctx->d_ctl, ctx->T, s);
  ctx->launches++;
  CK(cudaGetLastError());
  ctx->started = true;
  ctx->n_upper = 1;
  return KRRT_OK;
}

static int fetch_ctl(krrt_ctx* ctx, Ctl* h) {
  // One read-back serves every getter until the tree changes again: the node-style caller asks for the tree size and
  // the last sample after every single grow(xg) (robosoccer_motion_planning_node.cpp:175-179).
  if (!ctx->ctl_valid) {
    if (!ctx->h_ctl) CK(cudaHostAlloc((void**)&ctx->h_ctl, sizeof(Ctl), cudaHostAllocDefault));
    CK(cudaMemcpyAsync(ctx->h_ctl, ctx->d_ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->ctl_valid = true;
  }
  *h = *ctx->h_ctl;
  ctx->n_upper = h->n;
  return KRRT_OK;
}

constexpr int ITERLOG_MAX_ITERS = 256;   // iterations one logged call can record
constexpr int ITERLOG_STRIDE = 1024;     // nodes a snapshot row holds

static int grow_impl(krrt_ctx* ctx, const double* goal, int n_iter, const double* samples, int* solved, int* goal_index,
                     krrt_iter_record* records);

int krrt_grow(krrt_ctx* ctx, const double* goal, int n_iter, const double* samples, int* solved, int* goal_index) {
  return grow_impl(ctx, goal, n_iter, samples, solved, goal_index, nullptr);
}

// The caller's own loop `for(i<iteration) solved = rrt.grow(xg)` (robosoccer_motion_planning_node.cpp:176-185) in ONE device
// call: the same tree as n_iter single calls, plus what each of those calls would have returned (include/krrt.h)
int krrt_grow_logged(krrt_ctx* ctx, const double* goal, int n_iter, const double* samples, krrt_iter_record* records,
                     int* solved, int* goal_index) {
  if (!ctx || !records) return KRRT_ERR_INVALID;
  if (ctx->cfg.precision == KRRT_F32) {
    const int rc32 = krrt_f32_api::tree_grow_logged(ctx->f32, goal, n_iter, samples, records, solved, goal_index);
    double ms32 = 0.0;
    krrt_f32_api::tree_grow_done(ctx->f32, nullptr, nullptr, &ms32);
    ctx->last_total_ms = ms32;
    if (rc32 != KRRT_OK) ctx->err = krrt_f32_api::tree_error(ctx->f32);
    return rc32;
  }
  REQUIRE(n_iter >= 1 && n_iter <= ITERLOG_MAX_ITERS, KRRT_ERR_UNSUPPORTED, "krrt_grow_logged: 1 <= n_iter <= 256");
  REQUIRE(ctx->started, KRRT_ERR_STATE, "krrt_grow_logged before krrt_set_start");
  REQUIRE(ctx->n_upper + n_iter <= ITERLOG_STRIDE, KRRT_ERR_UNSUPPORTED,
          "krrt_grow_logged: the tree would outgrow the 1024 nodes a logged call can record; use krrt_grow");
  return grow_impl(ctx, goal, n_iter, samples, solved, goal_index, records);
}

static int grow_impl(krrt_ctx* ctx, const double* goal, int n_iter, const double* samples, int* solved, int* goal_index,
                     krrt_iter_record* records) {
  if (!ctx) return KRRT_ERR_INVALID;
  if (!records && ctx->cfg.precision == KRRT_F32) {  // the binary32 planner (krrt_f32.cu); krrt_last_timing gives its device time
    const int rc32 = krrt_f32_api::tree_grow(ctx->f32, goal, n_iter, samples, solved, goal_index);
    double ms32 = 0.0;
    krrt_f32_api::tree_grow_done(ctx->f32, nullptr, nullptr, &ms32);
    ctx->last_total_ms = ms32;
    if (rc32 != KRRT_OK) ctx->err = krrt_f32_api::tree_error(ctx->f32);
    return rc32;
  }
  F64_ONLY();
  REQUIRE(ctx->started, KRRT_ERR_STATE, "krrt_grow before krrt_set_start");
  REQUIRE(n_iter >= 0, KRRT_ERR_INVALID, "n_iter < 0");
  CK(cudaSetDevice(ctx->cfg.device));
  cudaStream_t st = ctx->stream;
  if (samples && n_iter > 0) {
    size_t need = sizeof(double) * 4 * (size_t)n_iter;
    if (ctx->user_cap < need) {
      if (ctx->d_user) CK(cudaFree(ctx->d_user));
      ctx->d_user = nullptr;
      ctx->user_cap = 0;
      CK(cudaMalloc((void**)&ctx->d_user, need));
      ctx->user_cap = need;
    }
    CK(cudaMemcpyAsync(ctx->d_user, samples, need, cudaMemcpyHostToDevice, st));
  }
  State4 g = {0, 0, 0, 0};
  if (goal) g = State4{goal[0], goal[1], goal[2], goal[3]};
  CK(cudaEventRecord(ctx->ev0, st));
  GrowCfg gcfg;
  gcfg.wave_enabled = ctx->wave_enabled;
  gcfg.wave_min_n = ctx->wave_min_n;
  gcfg.user_samples = samples ? ctx->d_user : nullptr;
  gcfg.bar = ctx->d_bar;
  gcfg.gi = ctx->gi;
  gcfg.log = IterLog{nullptr, nullptr, nullptr, nullptr, nullptr, 0, 0};
  const int n_before = ctx->n_upper;  // (exact: every grow call ends with a read-back of the control block)
  if (records) {
    if (!ctx->d_iterlog) {
      const size_t rows = ITERLOG_MAX_ITERS + 1, it = ITERLOG_MAX_ITERS;
      const size_t b_snap = rows * ITERLOG_STRIDE * sizeof(double), b_smp = it * 4 * sizeof(double), b_int = it * sizeof(int),
                   b_rec = it * sizeof(IterRec);
      CK(cudaMalloc((void**)&ctx->d_iterlog, 2 * b_snap + b_smp + b_rec + 2 * b_int));
      CK(cudaMemsetAsync(ctx->d_iterlog, 0, 2 * b_snap + b_smp + b_rec + 2 * b_int, st));  // no log entry is ever an uninitialised index
      unsigned char* q = ctx->d_iterlog;
      ctx->ilog.snap_ctc = (double*)q; q += b_snap;
      ctx->ilog.snap_tim = (double*)q; q += b_snap;
      ctx->ilog.sample = (double*)q; q += b_smp;
      ctx->d_iterrec = (IterRec*)q; q += b_rec;
      ctx->ilog.n_after = (int*)q; q += b_int;
      ctx->ilog.snap_of = (int*)q;
      ctx->ilog.stride = ITERLOG_STRIDE;
      ctx->ilog.rows = (int)rows;
    }
    gcfg.log = ctx->ilog;
    gcfg.wave_enabled = 0;  // every iteration goes through the in-order commit, which keeps the records
  }
  {  // sub-round sizing (profiles/r1_param_sweep.txt); the environment overrides are for sweeps
    auto envi = [](const char* k, int d) { const char* e = std::getenv(k); return e ? std::atoi(e) : d; };
    gcfg.pend_resolve = envi("KRRT_PEND", 1);
    gcfg.raw_forward = envi("KRRT_RAW_FORWARD", 1);
    gcfg.tail_seq = envi("KRRT_TAIL_SEQ", 1);
    gcfg.intra_fix = envi("KRRT_INTRA_FIX", 1);
    gcfg.prop_forward = envi("KRRT_PROP_FORWARD", 1);
    gcfg.join_test = envi("KRRT_JOIN_TEST", 1);
    gcfg.flow = envi("KRRT_FLOW", 0);  // measured: no gain
    gcfg.nb_mul = envi("KRRT_NB_MUL", 2); gcfg.nb_add = envi("KRRT_NB_ADD", 8); gcfg.nb_min = envi("KRRT_NB_MIN", 16);
    gcfg.pn_mul = envi("KRRT_PN_MUL", 3); gcfg.pn_add = envi("KRRT_PN_ADD", 16); gcfg.pn_min = envi("KRRT_PN_MIN", 48);
  }
  const int dyn = (int)(grow_shared_bytes() + commit_smem_bytes(ctx->Bt.near_cap));
  Ctl h;
  // ONE persistent kernel does the whole call; it only comes back early when the 23-bit epoch of the conflict marks
  // is used up (8 M sub-rounds): clear the marks and go on
  for (int launch = 0;; ++launch) {
    const bool wrap = launch > 0;
    if (wrap) {
      const size_t cap = (size_t)ctx->cfg.capacity;
      CK(cudaMemsetAsync(ctx->W.wmin, 0, sizeof(unsigned) * cap, st));
      CK(cudaMemsetAsync(ctx->W.amin, 0, sizeof(unsigned) * cap, st));
      CK(cudaMemsetAsync(ctx->W.rmin, 0, sizeof(unsigned long long) * cap, st));
      CK(cudaMemsetAsync(ctx->W.done, 0, sizeof(unsigned) * 1024, st));
      CK(cudaMemsetAsync(ctx->W.ok, 0, sizeof(unsigned) * 1024, st));
    }
    ctx->ctl_valid = false;
    k_begin_grow<<<1, 32, 0, st>>>(ctx->d_ctl, launch == 0, goal ? 1 : 0, g, n_iter, ctx->Bt, ctx->d_bar, wrap ? 1 : 0, ctx->T,
                                   gcfg.log);
    void* args[] = {(void*)&ctx->d_ctl, (void*)&ctx->P, (void*)&ctx->T, (void*)&ctx->Bt, (void*)&ctx->W, (void*)&gcfg};
    // A call of a few expansions that leaves the tree below the size where sub-rounds start (the node's own pattern:
    // one grow(xg) per iteration on 75-200-node trees, config/config.yaml:3) is committed in order by one warp whatever
    // the grid: it gets ONE CTA and a plain launch — no co-residency check, no grid-wide barriers (85 -> 58 us per call).
    // Longer calls keep the full grid: their windows of 32 samples are scanned and steered by all SMs.
    // (a logged call of up to 256 iterations on a tree that stays below 512 nodes likewise: its windows of 64 samples are a
    // few thousand pairs, and the in-order commit is one warp anyway)
    const bool small = ((n_iter <= 4 && ctx->n_upper + n_iter < ctx->small_max_n) || (records && ctx->n_upper + n_iter <= 512)) &&
                       !std::getenv("KRRT_NO_SMALL");
    if (small) {
      GrowCfg one = gcfg;
      one.wave_enabled = 0;  // a single CTA cannot run sub-rounds; in order is what a sequential planner does anyway
      if (records) k_grow_logged<<<1, GROW_THREADS, (size_t)dyn, st>>>(ctx->d_ctl, ctx->P, ctx->T, ctx->Bt, ctx->W, one);
      else k_grow<<<1, GROW_THREADS, (size_t)dyn, st>>>(ctx->d_ctl, ctx->P, ctx->T, ctx->Bt, ctx->W, one);
    } else {
      CK(cudaLaunchCooperativeKernel(records ? (const void*)k_grow_logged : (const void*)k_grow, dim3(ctx->grow_grid),
                                     dim3(GROW_THREADS), args, (size_t)dyn, st));
    }
    // the goal lookup and the end-of-call event follow at once: ONE read-back per launch (when the kernel came back early
    // to have its conflict marks cleared, the lookup and the event are simply repeated after the next launch)
    if (ctx->P.goal_mode == KRRT_GOAL_TOLERANCE) {
      k_find_goal<<<1, 1024, 0, st>>>(ctx->d_ctl, ctx->T, ctx->P.goal_tol2);
    } else {
      const long long n_max = (long long)ctx->n_upper + n_iter;
      const int ctas = (int)std::max<long long>(1, std::min<long long>(std::min(GOAL_MAX_CTAS, 4 * ctx->prop.multiProcessorCount),
                                                                       (n_max + GOAL_THREADS - 1) / GOAL_THREADS));
      k_goal_connect<<<ctas, GOAL_THREADS, 0, st>>>(ctx->d_ctl, ctx->P, ctx->T, ctx->goal_scr);
    }
    if (records) {
      static_assert(sizeof(IterRec) == sizeof(krrt_iter_record), "IterRec mirrors krrt_iter_record");
      k_goal_log<<<n_iter, GLOG_THREADS, 0, st>>>(ctx->d_ctl, ctx->P, ctx->T, gcfg.log, n_iter, n_before, ctx->d_iterrec);
      CK(cudaMemcpyAsync(records, ctx->d_iterrec, sizeof(IterRec) * (size_t)n_iter, cudaMemcpyDeviceToHost, st));
      ctx->launches += 1;
    }
    CK(cudaEventRecord(ctx->ev1, st));
    ctx->launches += 3;
    CK(cudaGetLastError());
    int rc = fetch_ctl(ctx, &h);
    if (rc != KRRT_OK) return rc;
    if (h.full || h.error || h.iter >= h.iter_target) break;
    REQUIRE(h.round >= (1u << 23) - 2, KRRT_ERR_CUDA, "k_grow returned before its target");
  }
  float ms = 0.f;
  CK(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
  ctx->last_total_ms = ms;
  ctx->last_grow_iters = h.iter - h.iter_begin;
  ctx->last_grow_full = h.full ? 1 : 0;
  ctx->last_scan_node_samples = h.scan_pairs;
  {
    // phase clocks of the kernel's last CTA -> shares of the call's device time
    double tot = 0.0;
    for (int k = 0; k < 8; ++k) tot += (double)h.phase_clk[k];
    const long long subs = h.rounds_wave + h.rounds_seq - h.tail_commits;
    for (int k = 0; k < 8; ++k) {
      ctx->phase_ms[k] = tot > 0.0 ? ms * (double)h.phase_clk[k] / tot : 0.0;
      ctx->phase_n[k] = k < 2 ? h.windows : (k < 5 ? subs : 0);
    }
    ctx->last_scan_ms = ctx->phase_ms[0];
    ctx->last_scan_launches = h.windows;
  }
  if (h.error) {
    ctx->err = h.error == KRRT_ERR_NEAR_OVERFLOW ? "near set exceeded near_cap; raise krrt_config.near_cap" : "device error";
    return h.error;
  }
  if (solved) *solved = h.goal_index >= 0;
  if (goal_index) *goal_index = h.goal_index;
  return KRRT_OK;
}

// how the last krrt_grow call ended: RRTStar::grow advances its iteration counter by n_iter whatever happens
// (include/rrtstar.hpp mirrors that from here); a tree that reached `capacity` stops early and says so
int krrt_grow_done(krrt_ctx* ctx, int64_t* iterations_done, int* full) {
  if (!ctx) return KRRT_ERR_INVALID;
  F32_FORWARD(krrt_f32_api::tree_grow_done(ctx->f32, iterations_done, full, nullptr));
  if (iterations_done) *iterations_done = ctx->last_grow_iters;
  if (full) *full = ctx->last_grow_full;
  return KRRT_OK;
}

int krrt_tree_size(krrt_ctx* ctx, int* n) {
  if (!ctx || !n) return KRRT_ERR_INVALID;
  F32_FORWARD(krrt_f32_api::tree_info(ctx->f32, n, nullptr, nullptr, nullptr, nullptr));
  REQUIRE(ctx->started, KRRT_ERR_STATE, "no tree: call krrt_set_start");
  CK(cudaSetDevice(ctx->cfg.device));
  Ctl h;
  int rc = fetch_ctl(ctx, &h);
  if (rc != KRRT_OK) return rc;
  *n = h.n;
  return KRRT_OK;
}

int krrt_iteration(krrt_ctx* ctx, int64_t* iter) {
  if (!ctx || !iter) return KRRT_ERR_INVALID;
  F32_FORWARD(krrt_f32_api::tree_info(ctx->f32, nullptr, iter, nullptr, nullptr, nullptr));
  REQUIRE(ctx->started, KRRT_ERR_STATE, "no tree: call krrt_set_start");
  CK(cudaSetDevice(ctx->cfg.device));
  Ctl h;
  int rc = fetch_ctl(ctx, &h);
  if (rc != KRRT_OK) return rc;
  *iter = h.iter;
  return KRRT_OK;
}

int krrt_last_sample(krrt_ctx* ctx, double out[4]) {
  if (!ctx || !out) return KRRT_ERR_INVALID;
  F32_FORWARD(krrt_f32_api::tree_info(ctx->f32, nullptr, nullptr, out, nullptr, nullptr));
  REQUIRE(ctx->started, KRRT_ERR_STATE, "no tree: call krrt_set_start");
  CK(cudaSetDevice(ctx->cfg.device));
  Ctl h;
  int rc = fetch_ctl(ctx, &h);
  if (rc != KRRT_OK) return rc;
  std::memcpy(out, h.last_sample, sizeof(double) * 4);
  return KRRT_OK;
}

int krrt_get_stats(krrt_ctx* ctx, krrt_stats* out) {
  if (!ctx || !out) return KRRT_ERR_INVALID;
  F32_FORWARD(krrt_f32_api::tree_info(ctx->f32, nullptr, nullptr, nullptr, out, nullptr));
  REQUIRE(ctx->started, KRRT_ERR_STATE, "no tree: call krrt_set_start");
  CK(cudaSetDevice(ctx->cfg.device));
  Ctl h;
  int rc = fetch_ctl(ctx, &h);
  if (rc != KRRT_OK) return rc;
  std::memcpy(out, h.stats, sizeof(krrt_stats));
  return KRRT_OK;
}

int krrt_kernel_times(krrt_ctx* ctx, double ms[8], int64_t launches[8]) {
  if (!ctx || !ms || !launches) return KRRT_ERR_INVALID;
  for (int k = 0; k < 8; ++k) {
    ms[k] = ctx->phase_ms[k];
    launches[k] = ctx->phase_n[k];
  }
  return KRRT_OK;
}

int krrt_debug_clocks(krrt_ctx* ctx, int64_t out[16]) {
  if (!ctx || !out) return KRRT_ERR_INVALID;
  CK(cudaSetDevice(ctx->cfg.device));
  Ctl h;
  int rc = fetch_ctl(ctx, &h);
  if (rc != KRRT_OK) return rc;
  for (int k = 0; k < 16; ++k) out[k] = h.dbg_clk[k];
  if (const char* path = std::getenv("KRRT_DEBUG_LOG")) {
    if (ctx->W.dbg_log) {
      int n = 0;
      cudaMemcpy(&n, ctx->W.dbg_log, sizeof(int), cudaMemcpyDeviceToHost);
      n = std::min(n, 1 << 20);
      std::vector<int> buf((size_t)8 * n);
      cudaMemcpy(buf.data(), ctx->W.dbg_log + 8, sizeof(int) * 8 * (size_t)n, cudaMemcpyDeviceToHost);
      if (FILE* f = std::fopen(path, "wb")) {
        std::fwrite(buf.data(), sizeof(int), buf.size(), f);
        std::fclose(f);
      }
    }
  }
  if (std::getenv("KRRT_DEBUG_DUMP")) {
    for (int c = 0; c < 160; ++c)
      if (h.dbg_cta[0][c] || h.dbg_cta[2][c])
        std::printf("cta %3d decide %10lld validate %10lld apply %10lld\n", c, h.dbg_cta[0][c], h.dbg_cta[1][c], h.dbg_cta[2][c]);
  }
  return KRRT_OK;
}

int krrt_wave_stats(krrt_ctx* ctx, int64_t out[8]) {
  if (!ctx || !out) return KRRT_ERR_INVALID;
  REQUIRE(ctx->started, KRRT_ERR_STATE, "no tree: call krrt_set_start");
  CK(cudaSetDevice(ctx->cfg.device));
  Ctl h;
  int rc = fetch_ctl(ctx, &h);
  if (rc != KRRT_OK) return rc;
  out[0] = h.rounds_wave;
  out[1] = h.rounds_seq;
  out[2] = h.wave_committed;
  for (int k = 0; k < 5; ++k) out[3 + k] = h.why[k];
  return KRRT_OK;
}

int krrt_tail_commits(krrt_ctx* ctx, int64_t* out) {
  if (!ctx || !out) return KRRT_ERR_INVALID;
  REQUIRE(ctx->started, KRRT_ERR_STATE, "no tree: call krrt_set_start");
  CK(cudaSetDevice(ctx->cfg.device));
  Ctl h;
  int rc = fetch_ctl(ctx, &h);
  if (rc != KRRT_OK) return rc;
  *out = h.tail_commits;
  return KRRT_OK;
}

int krrt_last_timing(krrt_ctx* ctx, double* total_ms, double* scan_ms, int64_t* scan_launches,
                     double* scan_node_samples) {
  if (!ctx) return KRRT_ERR_INVALID;
  if (total_ms) *total_ms = ctx->last_total_ms;
  if (scan_ms) *scan_ms = ctx->last_scan_ms;
  if (scan_launches) *scan_launches = ctx->last_scan_launches;
  if (scan_node_samples) *scan_node_samples = ctx->last_scan_node_samples;
  return KRRT_OK;
}

int krrt_scan_stats(krrt_ctx* ctx, double* visited_pairs, int64_t* rebuilds, int32_t* indexed_nodes) {
  if (!ctx) return KRRT_ERR_INVALID;
  CK(cudaSetDevice(ctx->cfg.device));
  Ctl h;
  int rc = fetch_ctl(ctx, &h);
  if (rc != KRRT_OK) return rc;
  if (visited_pairs) *visited_pairs = h.scan_visited;
  if (rebuilds) *rebuilds = h.grid_rebuilds;
  if (indexed_nodes) *indexed_nodes = h.grid_n;
  return KRRT_OK;
}

}  // extern "C"

#include "krrt_capi_ops.inl"
#include "krrt_capi_static.inl"
