// krrt_kernels.cuh — sm_100a kernels of the kinodynamic RRT* hot path.
//
// Single big tree (configs C1-C3): samples are counter-based (Philox), so the tree-independent work of a
// whole batch of B future expansions is evaluated in parallel against the tree as it stands ("frozen"),
// and a single warp then commits the B expansions strictly in order:
//   k_batch_prepare  draw the B samples, static rejection flags, in-batch (sample-vs-earlier-sample) pairs
//   k_near_scan      every frozen node vs every batch sample: 4-D distance, radius test, candidate append
//   k_steer_pairs    tau*/c* of both directions for every candidate pair
//   k_commit         in order: exact radius filter, parent choice, collision, insert, rewire, propagate
// Every arithmetic operation a sequential planner would do is done with the same operands; only the point
// in time differs, so trees are bit-identical with the CPU oracle whatever B is.
//
// Batched small trees (configs C4/C5): k_plan_batch, one warp per query, tree in shared memory.
#pragma once
#include "krrt_warp.cuh"

namespace krrt {

constexpr int IN_BATCH = 0x40000000;  // candidate ref flag: refers to an earlier sample of the same batch

struct Ctl {  // device-resident control block of one single-tree context
  int n;
  int full;
  int error;
  int goal_index;
  long long iter;
  int has_goal;
  int n0;       // tree size the current batch was evaluated against
  int nb;       // samples in the current batch
  int pad;
  double rho2max;
  double scan_pairs;  // sum over scan launches of (frozen nodes x batch samples) since the last grow began
  double scan_nodes;  // sum over scan launches of frozen nodes
  double goal[4];
  double last_sample[4];
  long long stats[10];
  // round control (all decided on the device: the host only enqueues rounds until iter == iter_target)
  long long iter_begin;   // iteration counter when the current krrt_grow call began (index base of user samples)
  long long iter_target;  // iteration counter at which the current krrt_grow call is complete
  unsigned round;         // round id: epoch of the conflict marks (krrt_wave.cuh)
  int wave;               // this round commits a validated prefix in parallel (tree large enough)
  int wave_F;             // length of that prefix
  int ema16;              // 16 x running mean of the prefix length: sizes the next round's batch
  long long rounds_wave, rounds_seq, wave_committed;  // diagnostics since set_start
  long long why[5];  // what ended the prefix: batch end, in-batch neighbour, needs sequential path, no free parent, read/write conflict
};

struct Params {
  double r;
  double R2;
  int S;
  int capacity;
  SamplerParams sp;
  uint32_t query_id;
  int m;
  int dynamic;
  double goal_tol2;
  const double* rho2;  // [capacity + 1], host-computed table of squared near radii
  const double4* obs;
};

struct Batch {
  State4* samples;        // [B]
  unsigned char* reject;  // [B] sample inside a static obstacle
  int* map;               // [B] node index the sample became, or -1
  unsigned char* dup;     // [B] a frozen node coincides with the sample (distance exactly 0)
  int* count;             // [B] candidates found (may exceed near_cap: overflow)
  int* ref;               // [B][near_cap]
  double* d2;             // [B][near_cap]
  double* tf;
  double* cf;
  double* tb;
  double* cb;
  int B;
  int near_cap;
};

// ---------------------------------------------------------------------------------------------------
__global__ void k_batch_prepare(Ctl* ctl, Params P, Batch Bt, int wave_enabled, int wave_min_n,
                                const double* user_samples) {
  __shared__ State4 ss[256];
  __shared__ double s_rho2max;
  const int b = threadIdx.x;
  const int n0 = ctl->n;
  const long long iter0 = ctl->iter;
  const long long left = ctl->iter_target - iter0;
  const int wave = wave_enabled && n0 >= wave_min_n;
  int nb_req = Bt.B;
  if (wave) nb_req = min(Bt.B, max(16, (2 * ctl->ema16) / 16 + 8));  // ~2x the prefix the last rounds committed
  const int nb = (ctl->full || ctl->error || left <= 0) ? 0 : (int)min((long long)nb_req, left);
  const int user_offset = (int)(iter0 - ctl->iter_begin);
  __syncthreads();  // every thread has read the control block before thread 0 rewrites it
  if (b == 0) {
    ctl->n0 = n0;
    ctl->nb = nb;
    ctl->wave = wave;
    ctl->wave_F = nb;
    ctl->round += 1;
    double mx = 0.0;
    int hi = min(n0 + nb, P.capacity);
    for (int k = n0; k <= hi; ++k) mx = fmax(mx, P.rho2[k]);
    s_rho2max = mx;
    ctl->rho2max = mx;
    if (nb > 0) {
      ctl->scan_pairs += (double)n0 * (double)nb;
      ctl->scan_nodes += (double)n0;
    }
  }
  State4 s = {0, 0, 0, 0};
  if (b < nb) {
    bool drawn = true;
    if (user_samples) {
      const double* u = user_samples + 4 * (size_t)(user_offset + b);
      if (!(u[0] != u[0])) {
        s.x = u[0]; s.y = u[1]; s.vx = u[2]; s.vy = u[3];
        drawn = false;
      }
    }
    if (drawn) {
      State4 g = {ctl->goal[0], ctl->goal[1], ctl->goal[2], ctl->goal[3]};
      s = draw_sample(P.sp, P.query_id, (uint64_t)(iter0 + b), ctl->has_goal != 0, g);
    }
    bool hit = false;
    if (!P.dynamic) {
      for (int j = 0; j < P.m; ++j) {
        double dx = s.x - P.obs[j].x, dy = s.y - P.obs[j].y;
        hit |= fma(dy, dy, dx * dx) < P.R2;
      }
    }
    Bt.samples[b] = s;
    Bt.reject[b] = hit ? 1 : 0;
    Bt.map[b] = -1;
    Bt.dup[b] = 0;
    ss[b] = s;
  }
  __syncthreads();
  if (b < nb) {
    const double rmax = s_rho2max;
    int cnt = 0;
    for (int e = 0; e < b; ++e) {
      double d = dist2(ss[e], s);
      if (d <= rmax) {
        if (cnt < Bt.near_cap) {
          Bt.ref[(size_t)b * Bt.near_cap + cnt] = IN_BATCH | e;
          Bt.d2[(size_t)b * Bt.near_cap + cnt] = d;
        }
        ++cnt;
      }
    }
    Bt.count[b] = cnt;
  }
}

// ---------------------------------------------------------------------------------------------------
// Neighbour scan: HBM/L2-bound read of the SoA states (32 B per node), B distance evaluations per node.
// Each thread keeps two nodes in registers and walks the batch samples broadcast from shared memory.
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_NODES_PER_CTA = 2 * SCAN_THREADS;

__global__ void __launch_bounds__(SCAN_THREADS) k_near_scan(const Ctl* ctl, const double2* __restrict__ pos,
                                                              const double2* __restrict__ vel, Batch Bt) {
  __shared__ State4 ss[256];
  const int n0 = ctl->n0;
  const int nb = ctl->nb;
  if (nb == 0 || (long long)blockIdx.x * SCAN_NODES_PER_CTA >= n0) return;
  const double rmax = ctl->rho2max;
  for (int b = threadIdx.x; b < nb; b += SCAN_THREADS) ss[b] = Bt.samples[b];
  __syncthreads();
  for (long long base = (long long)blockIdx.x * SCAN_NODES_PER_CTA; base < n0;
       base += (long long)gridDim.x * SCAN_NODES_PER_CTA) {
    const int i0 = (int)base + threadIdx.x, i1 = i0 + SCAN_THREADS;
    const bool v0 = i0 < n0, v1 = i1 < n0;
    State4 a = {0, 0, 0, 0}, c = {0, 0, 0, 0};
    if (v0) {
      double2 p = pos[i0], v = vel[i0];
      a.x = p.x; a.y = p.y; a.vx = v.x; a.vy = v.y;
    }
    if (v1) {
      double2 p = pos[i1], v = vel[i1];
      c.x = p.x; c.y = p.y; c.vx = v.x; c.vy = v.y;
    }
#pragma unroll 4
    for (int b = 0; b < nb; ++b) {
      const State4 s = ss[b];
      double da = dist2(a, s), dc = dist2(c, s);
      if (v0 && da <= rmax) {
        if (da == 0.0) Bt.dup[b] = 1;
        int slot = atomicAdd(&Bt.count[b], 1);
        if (slot < Bt.near_cap) {
          Bt.ref[(size_t)b * Bt.near_cap + slot] = i0;
          Bt.d2[(size_t)b * Bt.near_cap + slot] = da;
        }
      }
      if (v1 && dc <= rmax) {
        if (dc == 0.0) Bt.dup[b] = 1;
        int slot = atomicAdd(&Bt.count[b], 1);
        if (slot < Bt.near_cap) {
          Bt.ref[(size_t)b * Bt.near_cap + slot] = i1;
          Bt.d2[(size_t)b * Bt.near_cap + slot] = dc;
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------
constexpr int STEER_THREADS = 128;
__global__ void __launch_bounds__(STEER_THREADS) k_steer_pairs(const Ctl* ctl, const double2* __restrict__ pos,
                                                                 const double2* __restrict__ vel, Batch Bt,
                                                                 double r) {
  const int b = blockIdx.y;
  if (b >= ctl->nb) return;
  const int cnt = min(Bt.count[b], Bt.near_cap);
  const int e = blockIdx.x * STEER_THREADS + threadIdx.x;
  if (e >= cnt) return;
  const size_t o = (size_t)b * Bt.near_cap + e;
  const int ref = Bt.ref[o];
  State4 x;
  if (ref & IN_BATCH) {
    x = Bt.samples[ref & ~IN_BATCH];
  } else {
    double2 p = pos[ref], v = vel[ref];
    x.x = p.x; x.y = p.y; x.vx = v.x; x.vy = v.y;
  }
  const State4 s = Bt.samples[b];
  double t, c;
  steer(x, s, r, t, c);
  Bt.tf[o] = t;
  Bt.cf[o] = c;
  steer(s, x, r, t, c);
  Bt.tb[o] = t;
  Bt.cb[o] = c;
}

// ---------------------------------------------------------------------------------------------------
// Per-sample decision of a wave round (krrt_wave.cuh) and the conflict marks it is validated with.
enum { OUT_REJECT = 0, OUT_DUP = 1, OUT_NOFREE = 2, OUT_INSERT = 3, OUT_COMPLEX = 4, OUT_CONFLICT = 5 };
constexpr int WAVE_QCAP = 2048;  // write-list capacity per sample (descendants of its rewired nodes)

struct WaveRec {
  int outcome;
  int n_b;    // tree size this sample expects to see (== index of the node it inserts)
  int par;    // chosen parent (node index)
  int park;   // its slot in the sample's candidate list
  int K;      // near-set size
  int nrew;   // rewired nodes
  int nwl;    // strict descendants of the rewired nodes (== cost/time updates to propagate)
  int pad;
  double ctc_new, tim_new;
  long long collide_calls;
  double d2_in_max;   // largest squared distance inside the near set ...
  double d2_out_min;  // ... and smallest one outside it: the decision holds for any radius in between
};

struct Wave {
  WaveRec* rec;     // [B]
  int* near_idx;    // [B][near_cap] near set (node indices)
  int* rew_node;    // [B][near_cap] rewired nodes ...
  int* rew_slot;    // [B][near_cap] ... and their slots in the candidate list
  int* wl;          // [B][WAVE_QCAP] strict descendants of the rewired nodes / BFS scratch of the apply pass
  unsigned* wmin;   // [capacity] mark: lowest sample of this round that WRITES the node's cost/time
  unsigned* amin;   // [capacity] mark: lowest sample of this round that reads or writes it
  unsigned* dsc;    // [capacity] mark: lowest sample for which the node is a strict descendant of a rewired node
  int* lock;        // [capacity] child-list lock of a parent
};

// marks carry the round id, so they never need clearing: (round << 9) | (511 - sample), combined with atomicMax
__device__ __forceinline__ unsigned mark_key(unsigned round, int b) { return (round << 9) | (unsigned)(511 - b); }
__device__ __forceinline__ int mark_sample(unsigned m, unsigned round) {
  return (m >> 9) == round ? 511 - (int)(m & 511u) : 0x7fffffff;
}

// ---------------------------------------------------------------------------------------------------
// In-order commit of one batch by a single warp.  Dynamic shared memory: near list (near_cap entries) + obstacles.
__global__ void __launch_bounds__(32) k_commit(Ctl* ctl, Params P, TreeView T, Batch Bt, Wave W) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int lane = lane_id();
  const int cap = Bt.near_cap;
  double4* s_obs = reinterpret_cast<double4*>(smem);
  NearView N;
  N.cf = reinterpret_cast<double*>(s_obs + KRRT_MAX_OBS_DEV);
  N.tf = N.cf + cap;
  N.cb = N.tf + cap;
  N.tb = N.cb + cap;
  N.idx = reinterpret_cast<int*>(N.tb + cap);
  N.tried = reinterpret_cast<unsigned char*>(N.idx + cap);
  int nb = ctl->nb;
  if (nb == 0 || ctl->error) return;
  if (ctl->wave) {
    const int F = ctl->wave_F;
    if (F > 0) {  // the first F samples were committed in parallel (k_wave_apply): fold their counters in
      long long v[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
      for (int b = lane; b < F; b += 32) {
        const WaveRec r = W.rec[b];
        v[0] += 1;
        if (r.outcome == OUT_INSERT) {
          v[1] += 1; v[5] += r.nrew; v[6] += r.nwl; v[7] += r.K; v[8] += 2 * r.K - 1; v[9] += r.collide_calls;
        } else if (r.outcome == OUT_REJECT) {
          v[2] += 1;
        } else if (r.outcome == OUT_DUP) {
          v[3] += 1;
        } else {  // OUT_NOFREE
          v[4] += 1; v[7] += r.K; v[8] += r.K; v[9] += r.collide_calls;
        }
      }
#pragma unroll
      for (int k = 0; k < 10; ++k)
        for (int off = 16; off > 0; off >>= 1) v[k] += __shfl_xor_sync(FULL, v[k], off);
      if (lane == 0) {
        ctl->n += (int)v[1];
        ctl->iter += F;
        for (int k = 0; k < 10; ++k) ctl->stats[k] += v[k];
        const State4 ls = Bt.samples[F - 1];
        ctl->last_sample[0] = ls.x; ctl->last_sample[1] = ls.y; ctl->last_sample[2] = ls.vx; ctl->last_sample[3] = ls.vy;
        ctl->ema16 = (3 * ctl->ema16 + 16 * F) / 4;
        ctl->rounds_wave += 1;
        ctl->wave_committed += F;
        int why = 0;
        if (F < nb) {
          const int o = W.rec[F].outcome;
          why = o == OUT_CONFLICT ? 1 : o == OUT_COMPLEX ? 2 : (o == OUT_NOFREE ? 3 : 4);
        }
        ctl->why[why] += 1;
      }
      return;
    }
    nb = 1;  // the first sample needs the sequential path (empty near set, huge subtree, ...): commit it alone
    if (lane == 0) ctl->ema16 = (3 * ctl->ema16 + 16) / 4;
  }
  if (lane == 0) ctl->rounds_seq += 1;
  for (int j = lane; j < P.m; j += 32) s_obs[j] = P.obs[j];
  __syncwarp();
  EnvView E = {s_obs, P.m, P.dynamic, P.R2, P.S};

  Stats st = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  int n = ctl->n;
  long long iter = ctl->iter;
  int full = 0, error = 0;
  State4 s = {0, 0, 0, 0};
  for (int b = 0; b < nb; ++b) {
    s = Bt.samples[b];
    ++iter;
    st.iters++;
    if (!P.dynamic && Bt.reject[b]) {
      st.fail_reject++;
      continue;
    }
    if (n >= P.capacity) {
      full = 1;
      break;
    }
    const int cnt = Bt.count[b];
    if (cnt > cap) {
      error = 4;  // KRRT_ERR_NEAR_OVERFLOW
      --iter;
      st.iters--;
      break;
    }
    const double rho2n = P.rho2[n];
    int K = 0;
    bool dup = false;
    for (int base = 0; base < cnt; base += 32) {
      const int e = base + lane;
      bool valid = false;
      int node = -1;
      size_t o = (size_t)b * cap + e;
      if (e < cnt) {
        int ref = Bt.ref[o];
        double d = Bt.d2[o];
        node = (ref & IN_BATCH) ? Bt.map[ref & ~IN_BATCH] : ref;
        valid = node >= 0 && d <= rho2n;
        dup |= valid && d == 0.0;
      }
      unsigned mk = __ballot_sync(FULL, valid);
      if (valid) {
        int k = K + __popc(mk & ((1u << lane) - 1));
        N.idx[k] = node;
        N.cf[k] = Bt.cf[o];
        N.tf[k] = Bt.tf[o];
        N.cb[k] = Bt.cb[o];
        N.tb[k] = Bt.tb[o];
      }
      K += __popc(mk);
    }
    if (__any_sync(FULL, dup)) {
      st.fail_dup++;
      continue;
    }
    if (K == 0) {  // empty near set: exact nearest over the whole tree (rare once the tree is dense)
      int ni;
      double nd2;
      warp_nearest(T, n, s, ni, nd2);
      if (nd2 == 0.0) {
        st.fail_dup++;
        continue;
      }
      if (lane == 0) {
        double t, c;
        steer(load_state(T, ni), s, P.r, t, c);
        N.idx[0] = ni;
        N.cf[0] = c;
        N.tf[0] = t;
        N.cb[0] = 0.0;
        N.tb[0] = 0.0;
      }
      K = 1;
    }
    __syncwarp();
    st.sum_near += K;
    int nw = warp_commit(T, N, K, n, s, E, st);
    if (nw >= 0) {
      if (lane == 0) Bt.map[b] = nw;
      n = nw + 1;
    }
    __syncwarp();
  }
  if (lane == 0) {
    ctl->n = n;
    ctl->iter = iter;
    ctl->full = full;
    if (error) ctl->error = error;
    ctl->last_sample[0] = s.x; ctl->last_sample[1] = s.y; ctl->last_sample[2] = s.vx; ctl->last_sample[3] = s.vy;
    const long long* sv = &st.iters;
    for (int k = 0; k < 10; ++k) ctl->stats[k] += sv[k];
  }
}

// goal lookup after a grow call: cheapest node within goal_tol of the goal (rrt.goalIndex(), robosoccer_visual.hpp:19)
__global__ void __launch_bounds__(1024) k_find_goal(Ctl* ctl, TreeView T, double tol2) {
  __shared__ double s_c[32];
  __shared__ int s_i[32];
  if (!ctl->has_goal) {
    if (threadIdx.x == 0) ctl->goal_index = -1;
    return;
  }
  const int n = ctl->n;
  State4 g = {ctl->goal[0], ctl->goal[1], ctl->goal[2], ctl->goal[3]};
  double bc = INFINITY;
  int bi = -1, aux = 0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    if (dist2(load_state(T, i), g) <= tol2) {
      double c = T.ctc[i];
      if (c < bc) {
        bc = c;
        bi = i;
      }
    }
  }
  warp_argmin(bc, bi, aux);
  if ((threadIdx.x & 31) == 0) {
    s_c[threadIdx.x >> 5] = bc;
    s_i[threadIdx.x >> 5] = bi;
  }
  __syncthreads();
  if (threadIdx.x < 32) {
    bc = s_c[threadIdx.x];
    bi = s_i[threadIdx.x];
    warp_argmin(bc, bi, aux);
    if (threadIdx.x == 0) ctl->goal_index = bi;
  }
}

}  // namespace krrt
