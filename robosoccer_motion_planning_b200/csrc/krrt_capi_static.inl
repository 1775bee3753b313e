// krrt_capi_static.inl — krrt_plan_batch_static (include/krrt.h): the batched planner for callers that must not allocate.
// Included at the end of krrt_capi.cu, so the scratch array below, the cudaGetSymbolAddress that resolves it and the
// kernel launches that use it are all in ONE translation unit (a __device__ symbol has no meaning in another one).
//
// Who this is for.  Every other planning entry point goes through a krrt_ctx, and a context sizes its device buffers
// with cudaMalloc / cudaFree (at krrt_create, and again when a batch is larger than any before it).  Two kinds of caller
// cannot live with that:
//   * a planner that sits inside a robot's fixed-rate control cycle.  cudaMalloc and cudaFree take a driver-wide lock and
//     may wait for the whole device to go idle, so one of them in the cycle turns a 1 ms plan into a stall of unknown
//     length for every stream of the process;
//   * a process whose device memory already belongs to somebody else's pool (a framework's caching allocator that took
//     the whole GPU at start-up): a raw cudaMalloc from this library then fails, or eats into memory the pool counts on.
// Such a caller brings its own input and output buffers (it has them already) and this entry point brings the rest — the
// planner's scratch — out of the library image instead of the heap.
//
// What the scratch is.  k_plan_batch2 and its redo pass k_plan_batch need, beyond inputs and outputs: the table of squared
// near radii rho(n)^2 (one double per possible tree size), the work counter the warps pull queries from, ten 64-bit
// counters, and the redo list (one int per query at most, + its count and its own work counter).  Sized for
// KRRT_STATIC_MAX_QUERIES queries of KRRT_STATIC_MAX_ITER iterations that is 264 KB, which every GPU that loads libkrrt.so
// holds from module load until the process exits, used or not.  That is why the limits are modest.
//
// What follows from ONE array.  All calls on a device share it.  It keeps whatever the last call left: this call zeroes the
// counters and the redo header on the caller's stream and overwrites the part of the radius table it reads; nothing else
// is read before it is written.  Two calls that run on the GPU at the same time (two streams, two threads) write the same
// counters and corrupt each other: the caller orders them (same stream, or an event between them).
//
// g_static_scratch is a DEVICE symbol.  In host code its name stands for a host-side shadow the runtime keeps for
// look-ups — passing `g_static_scratch` (or &g_static_scratch[k]) to a kernel or to cudaMemsetAsync from host code would
// hand over a host address.  The device address comes from cudaGetSymbolAddress, once per device with that device
// current (each device has its own copy of the array at its own address), and every pointer below is derived from it.

#include <mutex>

namespace {

struct StaticScratch {
  double rho2[KRRT_STATIC_MAX_ITER + 2];      // tree sizes 0 .. cap, cap <= KRRT_STATIC_MAX_ITER + 1
  // --- zeroed by every call, as one range: [stats, redo + 2)
  long long stats[10];
  int counter;                                // k_plan_batch2's work counter
  int redo[2 + KRRT_STATIC_MAX_QUERIES];      // [0] work counter of the redo pass, [1] redo count, [2..] redo list
};
constexpr size_t kStaticZeroBegin = offsetof(StaticScratch, stats);
constexpr size_t kStaticZeroEnd = offsetof(StaticScratch, redo) + 2 * sizeof(int);
constexpr size_t kStaticBytes = (sizeof(StaticScratch) + 255) / 256 * 256;
constexpr int kStaticMaxDevices = 16;

struct StaticDevice {
  bool ready = false;
  unsigned char* base = nullptr;  // device address of g_static_scratch on this device
  int sms = 0;
  size_t smem_optin = 0;          // per CTA, opt-in (227 KB on sm_100)
  size_t smem_sm = 0;             // per SM (228 KB)
  double h_rho2[KRRT_STATIC_MAX_ITER + 2];  // host side of the radius table (the upload's source outlives the call)
};
StaticDevice g_static_dev[kStaticMaxDevices];
std::mutex g_static_setup_mu;     // two threads may make their first call at once; the calls themselves are the caller's to order

}  // namespace

static __device__ __align__(256) unsigned char g_static_scratch[kStaticBytes];

// no context to carry the text: it goes where krrt_create's goes (krrt_last_error(NULL), per thread)
#define SREQUIRE(cond, code, msg) \
  do {                            \
    if (!(cond)) {                \
      g_create_error = (msg);     \
      return (code);              \
    }                             \
  } while (0)
#define SCK(call)                                                              \
  do {                                                                         \
    cudaError_t e_ = (call);                                                   \
    if (e_ != cudaSuccess) {                                                   \
      g_create_error = std::string(#call) + ": " + cudaGetErrorString(e_);     \
      return KRRT_ERR_CUDA;                                                    \
    }                                                                          \
  } while (0)

// once per device, with that device current: where this device keeps the array, and what the launch maths needs to know
static int static_device_setup(int device, StaticDevice** out) {
  StaticDevice& d = g_static_dev[device];
  std::lock_guard<std::mutex> lock(g_static_setup_mu);
  if (!d.ready) {
    int major = 0, sms = 0, optin = 0, per_sm = 0;
    SCK(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
    if (major != 10) {
      g_create_error = "device is not sm_100; libkrrt is built for sm_100a only and has no CPU fallback";
      return KRRT_ERR_NO_DEVICE;
    }
    SCK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    SCK(cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
    SCK(cudaDeviceGetAttribute(&per_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, device));
    void* p = nullptr;
    SCK(cudaGetSymbolAddress(&p, g_static_scratch));
    SREQUIRE(p != nullptr && (reinterpret_cast<uintptr_t>(p) & 255) == 0, KRRT_ERR_CUDA,
             "cudaGetSymbolAddress gave no usable address for the static scratch");
    d.base = static_cast<unsigned char*>(p);
    d.sms = sms;
    d.smem_optin = (size_t)optin;
    d.smem_sm = (size_t)per_sm;
    d.ready = true;
  }
  *out = &d;
  return KRRT_OK;
}

extern "C" int krrt_plan_batch_static(const krrt_config* cfg, const double* d_starts, const double* d_goals,
                                      const double* d_obstacles, int q, int m, int dynamic, int n_iter,
                                      uint32_t query_base, krrt_result* d_out, void* cuda_stream) {
  // ---- arguments first: nothing below touches CUDA until they hold
  if (!cfg) {
    g_create_error = "krrt_plan_batch_static: cfg is NULL";
    return KRRT_ERR_INVALID;
  }
  const krrt_config& c = *cfg;
  SREQUIRE(q >= 0 && m >= 0 && m <= KRRT_MAX_OBS && n_iter >= 0, KRRT_ERR_INVALID, "bad q / m / n_iter");
  if (q > KRRT_STATIC_MAX_QUERIES || n_iter > KRRT_STATIC_MAX_ITER) {
    char msg[256];
    std::snprintf(msg, sizeof(msg),
                  "krrt_plan_batch_static: q = %d / n_iter = %d is beyond the static scratch (at most %d queries of %d "
                  "iterations per call); split the batch, or use krrt_plan_batch_device on a context",
                  q, n_iter, KRRT_STATIC_MAX_QUERIES, KRRT_STATIC_MAX_ITER);
    g_create_error = msg;
    return KRRT_ERR_INVALID;
  }
  SREQUIRE(c.precision == KRRT_F64, KRRT_ERR_UNSUPPORTED, "krrt_plan_batch_static plans in binary64 only");
  SREQUIRE(c.S >= 1 && c.r > 0.0 && c.R >= 0.0, KRRT_ERR_INVALID, "bad S / r / R");
  SREQUIRE(c.goal_mode == KRRT_GOAL_CONNECT || c.goal_mode == KRRT_GOAL_TOLERANCE, KRRT_ERR_INVALID, "bad goal_mode");
  SREQUIRE(c.device >= 0 && c.device < kStaticMaxDevices, KRRT_ERR_INVALID, "device ordinal out of range (0 .. 15)");
  if (q == 0) return KRRT_OK;
  SREQUIRE(d_starts && d_goals && d_out && (m == 0 || d_obstacles), KRRT_ERR_INVALID, "null buffer");
  const int cap = std::min(c.capacity, std::max(n_iter + 1, 2));  // as on a context: the root + one node per iteration at most
  SREQUIRE(cap >= 2, KRRT_ERR_INVALID, "n_iter / capacity too small");

  SCK(cudaSetDevice(c.device));
  StaticDevice* dev = nullptr;
  int rc = static_device_setup(c.device, &dev);
  if (rc != KRRT_OK) return rc;
  cudaStream_t st = (cudaStream_t)cuda_stream;  // NULL: the default stream (there is no context stream to fall back to)

  // ---- the same two kernels, the same launch shapes as launch_plan_batch (krrt_capi_ops.inl) picks for a binary64 context
  const size_t smem_max = dev->smem_optin;
  const size_t pw2 = batch2_warp_smem_bytes(cap, m);
  const int w2 = (int)std::min<size_t>(B2_MAX_WARPS, (smem_max - 1024) / pw2);
  const size_t pw = batch_warp_smem_bytes(cap, m);
  SREQUIRE(w2 >= 1 && pw <= smem_max, KRRT_ERR_UNSUPPORTED, "n_iter too large for the shared-memory batched planner");

  // pointers into the array: all arithmetic on the address cudaGetSymbolAddress returned, never on the symbol's name
  unsigned char* base = dev->base;
  double* s_rho2 = reinterpret_cast<double*>(base + offsetof(StaticScratch, rho2));
  long long* s_stats = reinterpret_cast<long long*>(base + offsetof(StaticScratch, stats));
  int* s_counter = reinterpret_cast<int*>(base + offsetof(StaticScratch, counter));
  int* s_redo = reinterpret_cast<int*>(base + offsetof(StaticScratch, redo));

  // what must start at zero does: both work counters, the ten counters, the redo count (the list itself is written before
  // it is read, and only its first `redo count` entries are)
  SCK(cudaMemsetAsync(base + kStaticZeroBegin, 0, kStaticZeroEnd - kStaticZeroBegin, st));
  // the radius table depends on gamma / rho_max, so every call uploads the cap + 1 entries it will read
  for (int n = 0; n <= cap; ++n) dev->h_rho2[n] = near_radius2(c, n);
  SCK(cudaMemcpyAsync(s_rho2, dev->h_rho2, sizeof(double) * (cap + 1), cudaMemcpyHostToDevice, st));

  BatchJob J;
  J.starts = d_starts; J.goals = d_goals; J.obstacles = d_obstacles;
  J.out = reinterpret_cast<ResultRec*>(d_out);
  J.next_query = s_counter;
  J.list = nullptr; J.list_count = nullptr;
  J.stats = s_stats;
  J.rho2 = s_rho2;
  J.q = q; J.m = m; J.dynamic = dynamic ? 1 : 0; J.n_iter = n_iter; J.cap = cap;
  J.query_base = query_base;
  J.r = c.r; J.R2 = c.R * c.R; J.goal_tol2 = c.goal_tol * c.goal_tol; J.S = c.S;
  J.goal_mode = c.goal_mode;
  J.chunk_done = nullptr;
  J.chunk_queries = 1;
  for (int k = 0; k < 4; ++k) { J.sp.lo[k] = c.lo[k]; J.sp.hi[k] = c.hi[k]; }
  J.sp.p_direct = c.p_direct;
  J.sp.seed = c.seed;

  {  // k_plan_batch2: near lists in registers; a query whose near set outgrows them goes on the redo list
    Batch2Job J2;
    J2.b = J;
    J2.redo_count = s_redo + 1;
    J2.redo_list = s_redo + 2;
    int w_run = w2;
    if ((long long)q <= (long long)dev->sms * w2) w_run = (int)std::max<long long>(1, ((long long)q + dev->sms - 1) / dev->sms);
    const size_t smem2 = (size_t)w_run * pw2;
    const int grid2 = (int)std::max<long long>(1, std::min<long long>(dev->sms, ((long long)q + w_run - 1) / w_run));
    SCK(cudaFuncSetAttribute(k_plan_batch2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
    k_plan_batch2<<<grid2, w_run * 32, smem2, st>>>(J2);
    SCK(cudaGetLastError());
  }
  {  // k_plan_batch over the redo list (every warp leaves after one atomic when the list is empty)
    J.next_query = s_redo;
    J.list = s_redo + 2;
    J.list_count = s_redo + 1;
    const size_t smem_sm = dev->smem_sm;
    int best_w = 1, best_total = 0, best_ctas = 1;
    for (int w = 1; w <= 8; ++w) {
      const size_t per_cta = (size_t)w * pw;
      if (per_cta > smem_max) break;
      int ctas = (int)(smem_sm / (per_cta + 1024));
      ctas = std::min(ctas, 32);
      if (ctas < 1) continue;
      if (w * ctas >= best_total) {
        best_total = w * ctas;
        best_w = w;
        best_ctas = ctas;
      }
    }
    const size_t smem = (size_t)best_w * pw;
    const int grid = (int)std::max<long long>(1, std::min<long long>((long long)dev->sms * best_ctas, ((long long)q + best_w - 1) / best_w));
    SCK(cudaFuncSetAttribute(k_plan_batch, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_plan_batch<<<grid, best_w * 32, smem, st>>>(J);
    SCK(cudaGetLastError());
  }
  return KRRT_OK;
}

#undef SREQUIRE
#undef SCK
