// krrt_mgpu.cu — libkrrt_mgpu.so: the planning sweep over the GPUs of one box (include/krrt_mgpu.h).
//
// Host code only: the kernels are libkrrt.so's (krrt_plan_batch_device_chunks), the exchange is NCCL's.  One calling
// thread drives every GPU of the handle: all work is enqueued asynchronously on two streams per GPU —
//   s_plan  the batched planner (one launch over the GPU's whole block: warps pull queries from a counter, so there is
//           no per-chunk tail)
//   s_xfer  for every chunk of the block: wait until the planner's completion counter of the chunk has reached the chunk's
//           size (cuStreamWaitValue32 — a front-end memory operation, no SM), then all-gather the chunk over NCCL
// so only the last chunk's gather is exposed.  With NCCL >= 2.28 the record buffers come from ncclMemAlloc, are registered
// as symmetric windows and the communicator asks for NCCL_CTA_POLICY_ZERO: the all-gathers run on the copy engines and
// take no SM from the planner, which holds every SM (one 15-warp CTA with 227 KB of shared memory per SM).
// Replaces solve_rrt() (robosoccer_motion_planning_node.cpp:167-196) x Q; SURVEY.md 8(b)/8(e).
#include <cuda.h>
#include <cuda_runtime.h>
#include <nccl.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/krrt_mgpu.h"

namespace {

thread_local std::string g_create_error;

typedef CUresult (*WaitValue32Fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);

struct Dev {
  int device = 0, rank = 0;
  krrt_ctx* ctx = nullptr;
  cudaStream_t s_plan = nullptr, s_xfer = nullptr;
  cudaEvent_t e0 = nullptr, e1 = nullptr, e_plan = nullptr, e_xfer = nullptr, e_load = nullptr, e_res = nullptr;
  ncclComm_t comm = nullptr;
  double *d_starts = nullptr, *d_goals = nullptr, *d_obs = nullptr;
  unsigned char *d_out = nullptr, *d_all = nullptr;  // this GPU's (padded) block of records / the gathered array
  bool nccl_mem = false;
#if NCCL_VERSION_CODE >= 22800
  ncclWindow_t w_out = nullptr, w_all = nullptr;
#endif
  uint32_t* d_done = nullptr;  // per-chunk completion counters
  double* d_red = nullptr;     // 16 doubles for the small reductions
};

}  // namespace

struct krrt_mgpu {
  krrt_config cfg;
  int world = 1, first = 0;
  std::vector<Dev> devs;
  std::string err;
  WaitValue32Fn wait_value = nullptr;
  int chunk = 1024;          // queries per gather chunk (KRRT_SWEEP_CHUNK)
  // the loaded sweep
  int q = 0, m = 0, pad = 0, n_chunks = 0;
  size_t in_cap = 0, obs_cap = 0, rec_cap = 0, done_cap = 0;
  bool ran = false, overlapped = false, have_results = false;
  double last_run_ms = 0.0;
  int nccl_version = 0;
};

#define MCK(call)                                                                      \
  do {                                                                                 \
    cudaError_t e_ = (call);                                                           \
    if (e_ != cudaSuccess) {                                                           \
      mg->err = std::string(#call) + ": " + cudaGetErrorString(e_);                    \
      return KRRT_ERR_CUDA;                                                            \
    }                                                                                  \
  } while (0)
#define NCK(call)                                                                      \
  do {                                                                                 \
    ncclResult_t r_ = (call);                                                          \
    if (r_ != ncclSuccess) {                                                           \
      mg->err = std::string(#call) + ": " + ncclGetErrorString(r_);                    \
      return KRRT_ERR_CUDA;                                                            \
    }                                                                                  \
  } while (0)
#define KCK(d, call)                                                                   \
  do {                                                                                 \
    int rc_ = (call);                                                                  \
    if (rc_ != KRRT_OK) {                                                              \
      mg->err = std::string(#call) + ": " + krrt_last_error((d).ctx);                  \
      return rc_;                                                                      \
    }                                                                                  \
  } while (0)
#define MREQUIRE(cond, code, msg) \
  do {                            \
    if (!(cond)) {                \
      mg->err = (msg);            \
      return (code);              \
    }                             \
  } while (0)

// One NCCL group over the local GPUs.  The group is closed on EVERY path: an error inside it is reported after
// ncclGroupEnd, so no later NCCL call of this thread (krrt_mgpu_destroy's ncclCommDestroy included) runs inside a group
// that was left open.  per_dev(Dev&) -> KRRT_OK or a status with mg->err set (MCK / NCK inside it return from the lambda).
template <typename F>
static int nccl_group(krrt_mgpu* mg, F&& per_dev) {
  NCK(ncclGroupStart());
  int rc = KRRT_OK;
  for (Dev& d : mg->devs)
    if ((rc = per_dev(d)) != KRRT_OK) break;
  const ncclResult_t e = ncclGroupEnd();
  if (rc != KRRT_OK) return rc;  // (mg->err is the first failure's)
  if (e != ncclSuccess) {
    mg->err = std::string("ncclGroupEnd: ") + ncclGetErrorString(e);
    return KRRT_ERR_CUDA;
  }
  return KRRT_OK;
}

static const size_t ITEM = sizeof(krrt_result);

extern "C" {

void krrt_mgpu_shard(int q, int rank, int world, int* lo, int* hi) {
  if (world < 1) world = 1;
  if (lo) *lo = (int)(((long long)q * rank) / world);
  if (hi) *hi = (int)(((long long)q * (rank + 1)) / world);
}

int krrt_mgpu_chunk_layout(int q, int world, int chunk, int rank, int c, int64_t* offset, int* count, int* global) {
  if (q < 0 || world < 1 || chunk < 1 || rank < 0 || rank >= world || c < 0) return KRRT_ERR_INVALID;
  const int pad = (q + world - 1) / world;
  const int n_chunks = std::max(1, (pad + chunk - 1) / chunk);
  if (c >= n_chunks) return KRRT_ERR_INVALID;
  int lo, hi;
  krrt_mgpu_shard(q, rank, world, &lo, &hi);
  const int c0 = c * chunk;                           // first local query of the chunk
  const int cc = std::min(chunk, pad - c0);           // records every rank contributes to it (padded)
  if (offset) *offset = (int64_t)c0 * world + (int64_t)rank * cc;   // all earlier chunks are full: chunk * world records each
  if (count) *count = std::max(0, std::min(cc, (hi - lo) - c0));
  if (global) *global = lo + c0;
  return KRRT_OK;
}

const char* krrt_mgpu_last_error(const krrt_mgpu* mg) { return mg ? mg->err.c_str() : g_create_error.c_str(); }

int krrt_mgpu_unique_id(void* id) {
  if (!id) return KRRT_ERR_INVALID;
  static_assert(sizeof(ncclUniqueId) == KRRT_MGPU_ID_BYTES, "id size");
  ncclUniqueId u;
  if (ncclGetUniqueId(&u) != ncclSuccess) return KRRT_ERR_CUDA;
  std::memcpy(id, &u, sizeof(u));
  return KRRT_OK;
}

void krrt_mgpu_destroy(krrt_mgpu* mg) {
  if (!mg) return;
  for (Dev& d : mg->devs) {
    cudaSetDevice(d.device);
    if (d.s_plan) cudaStreamSynchronize(d.s_plan);
    if (d.s_xfer) cudaStreamSynchronize(d.s_xfer);
  }
  for (Dev& d : mg->devs) {
    cudaSetDevice(d.device);
#if NCCL_VERSION_CODE >= 22800
    if (d.comm && d.w_out) ncclCommWindowDeregister(d.comm, d.w_out);
    if (d.comm && d.w_all) ncclCommWindowDeregister(d.comm, d.w_all);
#endif
    if (d.nccl_mem) {
      if (d.d_out) ncclMemFree(d.d_out);
      if (d.d_all) ncclMemFree(d.d_all);
    } else {
      if (d.d_out) cudaFree(d.d_out);
      if (d.d_all) cudaFree(d.d_all);
    }
    if (d.d_starts) cudaFree(d.d_starts);
    if (d.d_goals) cudaFree(d.d_goals);
    if (d.d_obs) cudaFree(d.d_obs);
    if (d.d_done) cudaFree(d.d_done);
    if (d.d_red) cudaFree(d.d_red);
    if (d.ctx) krrt_destroy(d.ctx);
  }
  for (Dev& d : mg->devs) {
    cudaSetDevice(d.device);
    if (d.comm) ncclCommDestroy(d.comm);
    if (d.e0) cudaEventDestroy(d.e0);
    if (d.e1) cudaEventDestroy(d.e1);
    if (d.e_plan) cudaEventDestroy(d.e_plan);
    if (d.e_xfer) cudaEventDestroy(d.e_xfer);
    if (d.e_load) cudaEventDestroy(d.e_load);
    if (d.e_res) cudaEventDestroy(d.e_res);
    if (d.s_plan) cudaStreamDestroy(d.s_plan);
    if (d.s_xfer) cudaStreamDestroy(d.s_xfer);
  }
  delete mg;
}

static int init_devices(krrt_mgpu* mg, const ncclUniqueId* id) {
  if (const char* e = std::getenv("KRRT_SWEEP_CHUNK")) mg->chunk = std::max(1, std::atoi(e));
  // cuStreamWaitValue32 through the runtime's driver entry point (no link against libcuda)
  {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &fn, cudaEnableDefault, &qr) == cudaSuccess &&
        qr == cudaDriverEntryPointSuccess && !std::getenv("KRRT_SWEEP_NO_OVERLAP"))
      mg->wait_value = (WaitValue32Fn)fn;
  }
  for (Dev& d : mg->devs) {
    MCK(cudaSetDevice(d.device));
    krrt_config c = mg->cfg;
    c.device = d.device;
    c.batch = 32;       // the contexts only run the batched planner: keep the single-tree window buffers small
    c.near_cap = 32;
    int rc = krrt_create(&c, &d.ctx);
    if (rc != KRRT_OK) {
      mg->err = std::string("krrt_create: ") + krrt_last_error(nullptr);
      return rc;
    }
    int lo_p = 0, hi_p = 0;
    MCK(cudaDeviceGetStreamPriorityRange(&lo_p, &hi_p));
    MCK(cudaStreamCreateWithPriority(&d.s_plan, cudaStreamNonBlocking, lo_p));
    MCK(cudaStreamCreateWithPriority(&d.s_xfer, cudaStreamNonBlocking, hi_p));
    MCK(cudaEventCreate(&d.e0));
    MCK(cudaEventCreate(&d.e1));
    MCK(cudaEventCreate(&d.e_load));
    MCK(cudaEventCreate(&d.e_res));
    MCK(cudaEventCreateWithFlags(&d.e_plan, cudaEventDisableTiming));
    MCK(cudaEventCreateWithFlags(&d.e_xfer, cudaEventDisableTiming));
    MCK(cudaMalloc((void**)&d.d_red, sizeof(double) * 16));
  }
  if (mg->world > 1) {
    int v = 0;
    ncclGetVersion(&v);
    mg->nccl_version = v;
    ncclConfig_t conf = NCCL_CONFIG_INITIALIZER;
#if NCCL_VERSION_CODE >= 22800
    if (v >= 22800 && !std::getenv("KRRT_SWEEP_NO_CE")) conf.CTAPolicy = NCCL_CTA_POLICY_ZERO;
#endif
    const int grc = nccl_group(mg, [&](Dev& d) -> int {
      MCK(cudaSetDevice(d.device));
      NCK(ncclCommInitRankConfig(&d.comm, mg->world, *id, d.rank, &conf));
      return KRRT_OK;
    });
    if (grc != KRRT_OK) return grc;
  }
  return KRRT_OK;
}

int krrt_mgpu_create(const krrt_config* cfg, int n_gpus, const int* devices, krrt_mgpu** out) {
  if (!cfg || !out || n_gpus < 1) {
    g_create_error = "null argument / n_gpus < 1";
    return KRRT_ERR_INVALID;
  }
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < n_gpus) {
    g_create_error = "fewer CUDA devices than n_gpus; libkrrt_mgpu has no CPU fallback";
    return KRRT_ERR_NO_DEVICE;
  }
  krrt_mgpu* mg = new krrt_mgpu();
  mg->cfg = *cfg;
  mg->world = n_gpus;
  mg->first = 0;
  mg->devs.resize(n_gpus);
  for (int g = 0; g < n_gpus; ++g) {
    mg->devs[g].device = devices ? devices[g] : g;
    mg->devs[g].rank = g;
  }
  ncclUniqueId id;
  std::memset(&id, 0, sizeof(id));
  int rc = KRRT_OK;
  if (n_gpus > 1 && ncclGetUniqueId(&id) != ncclSuccess) {
    mg->err = "ncclGetUniqueId failed";
    rc = KRRT_ERR_CUDA;
  }
  if (rc == KRRT_OK) rc = init_devices(mg, &id);
  if (rc != KRRT_OK) {
    g_create_error = mg->err;
    krrt_mgpu_destroy(mg);
    return rc;
  }
  *out = mg;
  return KRRT_OK;
}

int krrt_mgpu_create_rank(const krrt_config* cfg, int rank, int world, const void* id, krrt_mgpu** out) {
  if (!cfg || !out || world < 1 || rank < 0 || rank >= world || (world > 1 && !id)) {
    g_create_error = "bad rank / world / id";
    return KRRT_ERR_INVALID;
  }
  *out = nullptr;
  krrt_mgpu* mg = new krrt_mgpu();
  mg->cfg = *cfg;
  mg->world = world;
  mg->first = rank;
  mg->devs.resize(1);
  mg->devs[0].device = cfg->device;
  mg->devs[0].rank = rank;
  ncclUniqueId u;
  std::memset(&u, 0, sizeof(u));
  if (id) std::memcpy(&u, id, sizeof(u));
  int rc = init_devices(mg, &u);
  if (rc != KRRT_OK) {
    g_create_error = mg->err;
    krrt_mgpu_destroy(mg);
    return rc;
  }
  *out = mg;
  return KRRT_OK;
}

int krrt_mgpu_info(const krrt_mgpu* mg, int* world, int* local, int* first, int* nccl, int* nccl_version) {
  if (!mg) return KRRT_ERR_INVALID;
  if (world) *world = mg->world;
  if (local) *local = (int)mg->devs.size();
  if (first) *first = mg->first;
  if (nccl) *nccl = mg->world > 1 ? 1 : 0;
  if (nccl_version) *nccl_version = mg->nccl_version;
  return KRRT_OK;
}

int64_t krrt_mgpu_launch_count(const krrt_mgpu* mg) {
  int64_t n = 0;
  if (mg)
    for (const Dev& d : mg->devs) n += krrt_launch_count(d.ctx);
  return n;
}

int krrt_sweep_overlap(const krrt_mgpu* mg, int* overlapped, int* chunks) {
  if (!mg) return KRRT_ERR_INVALID;
  if (overlapped) *overlapped = mg->overlapped ? 1 : 0;
  if (chunks) *chunks = mg->n_chunks;
  return KRRT_OK;
}

// record buffers: NCCL's allocator + symmetric windows when the library offers them (copy-engine collectives), else cudaMalloc
static int alloc_records(krrt_mgpu* mg, size_t pad_records) {
  for (Dev& d : mg->devs) {
    MCK(cudaSetDevice(d.device));
    MCK(cudaStreamSynchronize(d.s_plan));
    MCK(cudaStreamSynchronize(d.s_xfer));
  }
  const size_t out_bytes = pad_records * ITEM, all_bytes = pad_records * ITEM * mg->world;
  bool use_nccl_mem = false;
#if NCCL_VERSION_CODE >= 22800
  use_nccl_mem = mg->world > 1 && mg->nccl_version >= 22800 && !std::getenv("KRRT_SWEEP_NO_CE");
#endif
  for (Dev& d : mg->devs) {
    MCK(cudaSetDevice(d.device));
#if NCCL_VERSION_CODE >= 22800
    if (d.w_out) NCK(ncclCommWindowDeregister(d.comm, d.w_out));
    if (d.w_all) NCK(ncclCommWindowDeregister(d.comm, d.w_all));
    d.w_out = d.w_all = nullptr;
#endif
    if (d.nccl_mem) {
      if (d.d_out) NCK(ncclMemFree(d.d_out));
      if (d.d_all) NCK(ncclMemFree(d.d_all));
    } else {
      if (d.d_out) MCK(cudaFree(d.d_out));
      if (d.d_all) MCK(cudaFree(d.d_all));
    }
    d.d_out = d.d_all = nullptr;
    d.nccl_mem = false;
  }
  mg->rec_cap = 0;
  if (use_nccl_mem) {
    bool ok = true;
    for (Dev& d : mg->devs) {
      MCK(cudaSetDevice(d.device));
      ok = ok && ncclMemAlloc((void**)&d.d_out, out_bytes) == ncclSuccess;
      ok = ok && ncclMemAlloc((void**)&d.d_all, all_bytes) == ncclSuccess;
      d.nccl_mem = true;
    }
    MREQUIRE(ok, KRRT_ERR_CUDA, "ncclMemAlloc failed");
#if NCCL_VERSION_CODE >= 22800
    // collective: every rank registers its buffers as symmetric windows (a failure here only loses the copy-engine path)
    bool wok = ncclGroupStart() == ncclSuccess;
    for (Dev& d : mg->devs) {
      cudaSetDevice(d.device);
      wok = wok && ncclCommWindowRegister(d.comm, d.d_out, out_bytes, &d.w_out, NCCL_WIN_COLL_SYMMETRIC) == ncclSuccess;
      wok = wok && ncclCommWindowRegister(d.comm, d.d_all, all_bytes, &d.w_all, NCCL_WIN_COLL_SYMMETRIC) == ncclSuccess;
    }
    wok = (ncclGroupEnd() == ncclSuccess) && wok;
    if (!wok && std::getenv("KRRT_SWEEP_DEBUG")) std::fprintf(stderr, "krrt_mgpu: symmetric window registration failed\n");
#endif
  } else {
    for (Dev& d : mg->devs) {
      MCK(cudaSetDevice(d.device));
      MCK(cudaMalloc((void**)&d.d_out, out_bytes));
      if (mg->world > 1) MCK(cudaMalloc((void**)&d.d_all, all_bytes));
    }
  }
  for (Dev& d : mg->devs) {
    MCK(cudaSetDevice(d.device));
    MCK(cudaMemsetAsync(d.d_out, 0, out_bytes, d.s_plan));
  }
  mg->rec_cap = pad_records;
  return KRRT_OK;
}

int krrt_sweep_load(krrt_mgpu* mg, const double* starts, const double* goals, const double* obstacles, int q, int m) {
  if (!mg) return KRRT_ERR_INVALID;
  MREQUIRE(q >= 0 && m >= 0 && m <= KRRT_MAX_OBS, KRRT_ERR_INVALID, "bad q / m");
  MREQUIRE(q == 0 || (starts && goals && (m == 0 || obstacles)), KRRT_ERR_INVALID, "null buffer");
  const int pad = std::max(1, (q + mg->world - 1) / mg->world);
  const int n_chunks = std::max(1, (pad + mg->chunk - 1) / mg->chunk);
  if ((size_t)pad > mg->in_cap || (size_t)pad * std::max(m, 1) > mg->obs_cap) {
    for (Dev& d : mg->devs) {
      MCK(cudaSetDevice(d.device));
      MCK(cudaStreamSynchronize(d.s_plan));
      if (d.d_starts) MCK(cudaFree(d.d_starts));
      if (d.d_goals) MCK(cudaFree(d.d_goals));
      if (d.d_obs) MCK(cudaFree(d.d_obs));
      d.d_starts = d.d_goals = d.d_obs = nullptr;
      MCK(cudaMalloc((void**)&d.d_starts, sizeof(double) * 4 * pad));
      MCK(cudaMalloc((void**)&d.d_goals, sizeof(double) * 4 * pad));
      MCK(cudaMalloc((void**)&d.d_obs, sizeof(double) * 4 * (size_t)pad * std::max(m, 1)));
    }
    mg->in_cap = pad;
    mg->obs_cap = (size_t)pad * std::max(m, 1);
  }
  if ((size_t)pad > mg->rec_cap) {
    int rc = alloc_records(mg, pad);
    if (rc != KRRT_OK) return rc;
  }
  if ((size_t)n_chunks > mg->done_cap) {
    for (Dev& d : mg->devs) {
      MCK(cudaSetDevice(d.device));
      MCK(cudaStreamSynchronize(d.s_plan));
      MCK(cudaStreamSynchronize(d.s_xfer));
      if (d.d_done) MCK(cudaFree(d.d_done));
      d.d_done = nullptr;
      MCK(cudaMalloc((void**)&d.d_done, sizeof(uint32_t) * n_chunks));
    }
    mg->done_cap = n_chunks;
  }
  for (Dev& d : mg->devs) {
    MCK(cudaSetDevice(d.device));
    MCK(cudaEventRecord(d.e_load, d.s_plan));
    int lo, hi;
    krrt_mgpu_shard(q, d.rank, mg->world, &lo, &hi);
    const size_t n = (size_t)(hi - lo);
    if (n == 0) continue;
    MCK(cudaMemcpyAsync(d.d_starts, starts + 4 * (size_t)lo, sizeof(double) * 4 * n, cudaMemcpyHostToDevice, d.s_plan));
    MCK(cudaMemcpyAsync(d.d_goals, goals + 4 * (size_t)lo, sizeof(double) * 4 * n, cudaMemcpyHostToDevice, d.s_plan));
    if (m > 0)
      MCK(cudaMemcpyAsync(d.d_obs, obstacles + 4 * (size_t)lo * m, sizeof(double) * 4 * n * m, cudaMemcpyHostToDevice, d.s_plan));
  }
  for (Dev& d : mg->devs) {  // the caller's buffers may be reused right after return
    MCK(cudaSetDevice(d.device));
    MCK(cudaStreamSynchronize(d.s_plan));
  }
  mg->q = q;
  mg->m = m;
  mg->pad = pad;
  mg->n_chunks = n_chunks;
  mg->ran = false;
  mg->have_results = false;
  return KRRT_OK;
}

int krrt_sweep_run(krrt_mgpu* mg, int dynamic, int n_iter, uint32_t query_base, double* device_ms) {
  if (!mg) return KRRT_ERR_INVALID;
  MREQUIRE(mg->pad > 0, KRRT_ERR_STATE, "krrt_sweep_run before krrt_sweep_load");
  MREQUIRE(n_iter >= 0, KRRT_ERR_INVALID, "n_iter < 0");
  const int C = mg->chunk, nch = mg->n_chunks, W = mg->world;
  const bool gather = W > 1;
  const bool overlap = gather && mg->wait_value != nullptr && nch > 1;
  for (Dev& d : mg->devs) {
    MCK(cudaSetDevice(d.device));
    int lo, hi;
    krrt_mgpu_shard(mg->q, d.rank, W, &lo, &hi);
    MCK(cudaMemsetAsync(d.d_done, 0, sizeof(uint32_t) * nch, d.s_plan));
    MCK(cudaEventRecord(d.e0, d.s_plan));
    if (gather) MCK(cudaStreamWaitEvent(d.s_xfer, d.e0, 0));  // the counters are armed before anyone waits on them
    if (hi > lo)
      KCK(d, krrt_plan_batch_device_chunks(d.ctx, d.d_starts, d.d_goals, d.d_obs, hi - lo, mg->m, dynamic, n_iter,
                                           query_base + (uint32_t)lo, (krrt_result*)d.d_out, C, d.d_done, d.s_plan));
    if (gather && !overlap) {
      MCK(cudaEventRecord(d.e_plan, d.s_plan));
      MCK(cudaStreamWaitEvent(d.s_xfer, d.e_plan, 0));
    }
  }
  if (gather) {
    for (int c = 0; c < nch; ++c) {
      const int c0 = c * C, cc = std::min(C, mg->pad - c0);
      if (overlap) {
        for (Dev& d : mg->devs) {
          MCK(cudaSetDevice(d.device));
          int cnt = 0;
          krrt_mgpu_chunk_layout(mg->q, W, C, d.rank, c, nullptr, &cnt, nullptr);
          if (cnt > 0) {
            CUresult r = mg->wait_value((CUstream)d.s_xfer, (CUdeviceptr)(uintptr_t)(d.d_done + c), (cuuint32_t)cnt,
                                        CU_STREAM_WAIT_VALUE_GEQ);
            MREQUIRE(r == CUDA_SUCCESS, KRRT_ERR_CUDA, "cuStreamWaitValue32 failed");
          }
        }
      }
      const int grc = nccl_group(mg, [&](Dev& d) -> int {
        MCK(cudaSetDevice(d.device));
        NCK(ncclAllGather(d.d_out + (size_t)c0 * ITEM, d.d_all + (size_t)c0 * W * ITEM, (size_t)cc * ITEM, ncclChar, d.comm,
                          d.s_xfer));
        return KRRT_OK;
      });
      if (grc != KRRT_OK) return grc;
    }
    for (Dev& d : mg->devs) {
      MCK(cudaSetDevice(d.device));
      MCK(cudaEventRecord(d.e_xfer, d.s_xfer));
      MCK(cudaStreamWaitEvent(d.s_plan, d.e_xfer, 0));
    }
  }
  for (Dev& d : mg->devs) {
    MCK(cudaSetDevice(d.device));
    MCK(cudaEventRecord(d.e1, d.s_plan));
  }
  double ms_max = 0.0;
  for (Dev& d : mg->devs) {
    MCK(cudaSetDevice(d.device));
    MCK(cudaEventSynchronize(d.e1));
    float ms = 0.f;
    MCK(cudaEventElapsedTime(&ms, d.e0, d.e1));
    ms_max = std::max(ms_max, (double)ms);
  }
  if (device_ms) *device_ms = ms_max;
  mg->last_run_ms = ms_max;
  mg->ran = true;
  mg->overlapped = overlap;
  return KRRT_OK;
}

int krrt_sweep_results(krrt_mgpu* mg, krrt_result* out) {
  if (!mg || !out) return KRRT_ERR_INVALID;
  MREQUIRE(mg->ran, KRRT_ERR_STATE, "krrt_sweep_results before krrt_sweep_run");
  Dev& d = mg->devs[0];
  MCK(cudaSetDevice(d.device));
  if (mg->world == 1) {
    if (mg->q > 0) MCK(cudaMemcpyAsync(out, d.d_out, (size_t)mg->q * ITEM, cudaMemcpyDeviceToHost, d.s_plan));
  } else {
    for (int r = 0; r < mg->world; ++r)
      for (int c = 0; c < mg->n_chunks; ++c) {
        int64_t off = 0;
        int cnt = 0, glob = 0;
        krrt_mgpu_chunk_layout(mg->q, mg->world, mg->chunk, r, c, &off, &cnt, &glob);
        if (cnt > 0)
          MCK(cudaMemcpyAsync(out + glob, d.d_all + (size_t)off * ITEM, (size_t)cnt * ITEM, cudaMemcpyDeviceToHost, d.s_plan));
      }
  }
  MCK(cudaEventRecord(d.e_res, d.s_plan));
  MCK(cudaStreamSynchronize(d.s_plan));
  mg->have_results = true;
  return KRRT_OK;
}

int krrt_sweep_last_timing(krrt_mgpu* mg, double* run_ms, double* e2e_ms) {
  if (!mg) return KRRT_ERR_INVALID;
  if (run_ms) *run_ms = mg->last_run_ms;
  if (e2e_ms) {
    *e2e_ms = 0.0;
    if (mg->have_results) {
      Dev& d = mg->devs[0];
      MCK(cudaSetDevice(d.device));
      float ms = 0.f;
      MCK(cudaEventElapsedTime(&ms, d.e_load, d.e_res));
      *e2e_ms = ms;
    }
  }
  return KRRT_OK;
}

// all-reduce of n <= 16 doubles over every rank of the job; the result is read back from the first local GPU
static int reduce_small(krrt_mgpu* mg, double* x, int n, ncclRedOp_t op) {
  if (mg->world == 1) return KRRT_OK;
  // each local GPU contributes: the first one the values, the others the identity — except that every local GPU of mode A
  // carries its own partial, supplied by the caller through x[g * n .. ) when local > 1
  const int L = (int)mg->devs.size();
  for (int g = 0; g < L; ++g) {
    Dev& d = mg->devs[g];
    MCK(cudaSetDevice(d.device));
    MCK(cudaMemcpyAsync(d.d_red, x + (size_t)g * n, sizeof(double) * n, cudaMemcpyHostToDevice, d.s_plan));
  }
  const int grc = nccl_group(mg, [&](Dev& d) -> int {
    MCK(cudaSetDevice(d.device));
    NCK(ncclAllReduce(d.d_red, d.d_red, n, ncclDouble, op, d.comm, d.s_plan));
    return KRRT_OK;
  });
  if (grc != KRRT_OK) return grc;
  Dev& d0 = mg->devs[0];
  MCK(cudaSetDevice(d0.device));
  MCK(cudaMemcpyAsync(x, d0.d_red, sizeof(double) * n, cudaMemcpyDeviceToHost, d0.s_plan));
  for (Dev& d : mg->devs) {
    MCK(cudaSetDevice(d.device));
    MCK(cudaStreamSynchronize(d.s_plan));
  }
  return KRRT_OK;
}

int krrt_sweep_stats(krrt_mgpu* mg, krrt_stats* out) {
  if (!mg || !out) return KRRT_ERR_INVALID;
  MREQUIRE(mg->ran, KRRT_ERR_STATE, "krrt_sweep_stats before krrt_sweep_run");
  const int L = (int)mg->devs.size();
  std::vector<double> v((size_t)L * 10, 0.0);   // counters of a sweep are far below 2^53: exact in a double
  for (int g = 0; g < L; ++g) {
    Dev& d = mg->devs[g];
    int lo, hi;
    krrt_mgpu_shard(mg->q, d.rank, mg->world, &lo, &hi);
    if (hi == lo) continue;
    krrt_stats s;
    KCK(d, krrt_batch_stats(d.ctx, &s));
    const int64_t* p = &s.iters;
    for (int k = 0; k < 10; ++k) v[(size_t)g * 10 + k] = (double)p[k];
  }
  if (mg->world > 1) {
    int rc = reduce_small(mg, v.data(), 10, ncclSum);
    if (rc != KRRT_OK) return rc;
  }
  int64_t* o = &out->iters;
  for (int k = 0; k < 10; ++k) o[k] = (int64_t)v[k];
  return KRRT_OK;
}

int krrt_mgpu_max(krrt_mgpu* mg, double* x) {
  if (!mg || !x) return KRRT_ERR_INVALID;
  std::vector<double> v(mg->devs.size(), *x);
  int rc = reduce_small(mg, v.data(), 1, ncclMax);
  if (rc != KRRT_OK) return rc;
  *x = v[0];
  return KRRT_OK;
}

int krrt_mgpu_barrier(krrt_mgpu* mg) {
  if (!mg) return KRRT_ERR_INVALID;
  for (Dev& d : mg->devs) {
    MCK(cudaSetDevice(d.device));
    MCK(cudaStreamSynchronize(d.s_plan));
    MCK(cudaStreamSynchronize(d.s_xfer));
  }
  double one = 1.0;
  return krrt_mgpu_max(mg, &one);
}

int krrt_sweep_multi_gpu(krrt_mgpu* mg, const double* starts, const double* goals, const double* obstacles, int q, int m,
                         int dynamic, int n_iter, uint32_t query_base, krrt_result* out) {
  if (!mg) return KRRT_ERR_INVALID;
  MREQUIRE(q == 0 || out, KRRT_ERR_INVALID, "null result buffer");
  int rc = krrt_sweep_load(mg, starts, goals, obstacles, q, m);
  if (rc != KRRT_OK) return rc;
  if (q == 0) return KRRT_OK;
  if ((rc = krrt_sweep_run(mg, dynamic, n_iter, query_base, nullptr)) != KRRT_OK) return rc;
  return krrt_sweep_results(mg, out);
}

}  // extern "C"
