// krrt_kernels.cuh — sm_100a kernels of the kinodynamic RRT* hot path.
//
// Single big tree (configs C1-C3): samples are counter-based (Philox), so the tree-independent work of a whole
// batch of future expansions is evaluated in parallel against the tree as it stands ("frozen"); a validated prefix of
// the batch is then committed in parallel (krrt_wave.cuh) or, while the tree is small, the batch is committed strictly
// in order by one warp (round_finish below).  Every arithmetic operation a sequential planner would do is done with
// the same operands; only the point in time differs, so trees are bit-identical with a sequential run.
//
// Batched small trees (configs C4/C5): k_plan_batch (krrt_batch.cuh), one warp per query, tree in shared memory.
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
// End of a round, ONE warp: fold in the counters of the prefix the wave committed in parallel, or commit in order
// (the whole batch while the tree is small; only the first sample when that sample needs the sequential path).
// smem: obstacles + near list (commit_smem_bytes).
__host__ __device__ inline size_t commit_smem_bytes(int near_cap) {
  return sizeof(double4) * KRRT_MAX_OBS_DEV + (size_t)near_cap * (4 * 8 + 4 + 1) + 16;
}

__device__ __noinline__ void round_finish_body(Ctl* ctl, const Params& P, const TreeView& T, const Batch& Bt,
                                               const Wave& W, unsigned char* smem) {
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

__device__ __forceinline__ void round_finish(Ctl* ctl, const Params& P, const TreeView& T, const Batch& Bt,
                                             const Wave& W, unsigned char* smem) {
  round_finish_body(ctl, P, T, Bt, W, smem);
  __syncwarp();
  for (int b = lane_id(); b < Bt.B; b += 32) {  // candidate lists of the next round start empty
    Bt.count[b] = 0;
    Bt.dup[b] = 0;
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
