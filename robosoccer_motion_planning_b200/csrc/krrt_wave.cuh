// krrt_wave.cuh — one ROUND of the single-tree planner: four kernels, all decisions taken on the device.
//
//   k_round_scan      draw the batch (Philox), static rejection, in-batch pairs; every frozen node x every batch
//                     sample: 4-D distance, radius test, candidate append            (HBM/L2 bytes, fp64 at large B)
//   k_round_decide    one CTA per sample: tau*/c* of both directions for every candidate pair, then the sample's
//                     whole expansion against the frozen tree (near set at the exact radius, parent = arg-min over
//                     collision-free candidates, rewire set), and conflict marks on what it read / would write
//   k_round_validate  one warp per sample: longest prefix [0, F) of the batch whose decisions are exactly what a
//                     sequential planner would take
//   k_round_apply     one CTA per sample of the prefix: insert, rewire, level-order cost/time propagation; one
//                     extra CTA finishes the round (counters, control block, sequential fall-back)
//
// Why the prefix is exact.  RRT* is sequential: expansion k+1 sees the tree after k.  But an expansion only READS
// the cost/time of its near set and only WRITES the cost/time of the nodes it rewires and of their descendants.
// Each decision marks both sets; sample b stays in the prefix iff no earlier sample writes anything b reads or
// writes, no earlier sample reads anything b writes, no earlier sample's new node can fall inside b's near radius,
// and b's near set is unchanged at the radius of the tree size it really sees (earlier samples that found no
// collision-free parent did not insert).  For such a prefix the sequential outcome IS the speculated one.  The batch
// restarts at sample F in the next round (as sample 0, which has nothing before it).  A sample that cannot be decided
// this way at all (empty near set -> global nearest, a rewired subtree larger than WAVE_QCAP, a near node inside one
// of its own rewired subtrees) is committed alone by the sequential warp; while the tree is small (conflicts
// everywhere) whole batches are.  Trees are bit-identical to the in-order commit whatever the prefix lengths are.
//
// Stands in for the inside of RRTStar::grow (robosoccer_motion_planning_node.cpp:177), like krrt_warp.cuh.
#pragma once
#include "krrt_kernels.cuh"

namespace krrt {

constexpr int ROUND_THREADS = 256;

struct RoundCfg {
  int wave_enabled;
  int wave_min_n;
  const double* user_samples;  // device copy of the caller's explicit samples, or nullptr
};

struct RoundPlan {
  int n0, nb, wave, user_offset;
  long long iter0;
};

// what this round will do: a pure function of the control block's persistent fields, so every CTA of the scan
// derives it on its own (CTA 0 publishes it for the later kernels)
__device__ __forceinline__ RoundPlan round_plan(const Ctl* ctl, int B, const RoundCfg& cfg) {
  RoundPlan p;
  p.n0 = ctl->n;
  p.iter0 = ctl->iter;
  const long long left = ctl->iter_target - p.iter0;
  p.wave = cfg.wave_enabled && p.n0 >= cfg.wave_min_n;
  int nb_req = B;
  if (p.wave) nb_req = min(B, max(16, (2 * ctl->ema16) / 16 + 8));  // ~2x the prefix the last rounds committed
  p.nb = (ctl->full || ctl->error || left <= 0) ? 0 : (int)min((long long)nb_req, left);
  p.user_offset = (int)(p.iter0 - ctl->iter_begin);
  return p;
}

__device__ __forceinline__ void append_candidate(const Batch& Bt, int b, int ref, double d) {
  const int slot = atomicAdd(&Bt.count[b], 1);
  if (slot < Bt.near_cap) {
    Bt.ref[(size_t)b * Bt.near_cap + slot] = ref;
    Bt.d2[(size_t)b * Bt.near_cap + slot] = d;
  }
}

// ---------------------------------------------------------------------------------------------------
// grid (node tiles, sample chunks): thread = one frozen node, loop over this CTA's chunk of the batch samples
// (broadcast from shared memory).  32 B per node read once per chunk, coalesced 16-B loads.
__global__ void __launch_bounds__(ROUND_THREADS) k_round_scan(Ctl* ctl, Params P, Batch Bt, RoundCfg cfg,
                                                                const double2* __restrict__ pos,
                                                                const double2* __restrict__ vel) {
  __shared__ State4 ss[256];
  __shared__ double s_red[ROUND_THREADS / 32];
  const RoundPlan rp = round_plan(ctl, Bt.B, cfg);
  const int t = threadIdx.x;
  const bool lead = blockIdx.x == 0 && blockIdx.y == 0;
  const int nb = rp.nb, n0 = rp.n0;
  if (nb == 0) {
    if (lead && t == 0) {
      ctl->nb = 0;
      ctl->wave = 0;
    }
    return;
  }
  // squared radius that covers every tree size this batch can see
  double rmax = 0.0;
  if (t <= nb) rmax = P.rho2[min(n0 + t, P.capacity)];
  for (int off = 16; off > 0; off >>= 1) rmax = fmax(rmax, __shfl_xor_sync(FULL, rmax, off));
  if ((t & 31) == 0) s_red[t >> 5] = rmax;
  // the batch: sample b belongs to iteration iter0 + b
  bool hit = false;
  if (t < nb) {
    State4 s = {0, 0, 0, 0};
    bool drawn = true;
    if (cfg.user_samples) {
      const double* u = cfg.user_samples + 4 * (size_t)(rp.user_offset + t);
      if (!(u[0] != u[0])) {
        s.x = u[0]; s.y = u[1]; s.vx = u[2]; s.vy = u[3];
        drawn = false;
      }
    }
    if (drawn) {
      const State4 g = {ctl->goal[0], ctl->goal[1], ctl->goal[2], ctl->goal[3]};
      s = draw_sample(P.sp, P.query_id, (uint64_t)(rp.iter0 + t), ctl->has_goal != 0, g);
    }
    if (!P.dynamic) {
      for (int j = 0; j < P.m; ++j) {
        const double dx = s.x - P.obs[j].x, dy = s.y - P.obs[j].y;
        hit |= fma(dy, dy, dx * dx) < P.R2;
      }
    }
    ss[t] = s;
  }
  __syncthreads();
  rmax = s_red[0];
#pragma unroll
  for (int w = 1; w < ROUND_THREADS / 32; ++w) rmax = fmax(rmax, s_red[w]);
  if (lead) {
    if (t < nb) {
      Bt.samples[t] = ss[t];
      Bt.reject[t] = hit ? 1 : 0;
      Bt.map[t] = -1;
    }
    // in-batch pairs: an earlier sample of the batch may become a node inside a later sample's radius
    for (int p = t; p < nb * nb; p += ROUND_THREADS) {
      const int e = p / nb, b = p - e * nb;
      if (e < b) {
        const double d = dist2(ss[e], ss[b]);
        if (d <= rmax) append_candidate(Bt, b, IN_BATCH | e, d);
      }
    }
    if (t == 0) {
      ctl->n0 = n0;
      ctl->nb = nb;
      ctl->wave = rp.wave;
      ctl->wave_F = nb;
      ctl->round += 1;
      ctl->rho2max = rmax;
      ctl->scan_pairs += (double)n0 * (double)nb;
      ctl->scan_nodes += (double)n0;
    }
  }
  const int chunk = (nb + gridDim.y - 1) / gridDim.y;
  const int b0 = blockIdx.y * chunk, b1 = min(nb, b0 + chunk);
  for (int i = blockIdx.x * ROUND_THREADS + t; i < n0; i += gridDim.x * ROUND_THREADS) {
    const double2 pp = pos[i], vv = vel[i];
    const State4 a = {pp.x, pp.y, vv.x, vv.y};
#pragma unroll 4
    for (int b = b0; b < b1; ++b) {
      const double d = dist2(a, ss[b]);
      if (d <= rmax) {
        if (d == 0.0) Bt.dup[b] = 1;
        append_candidate(Bt, b, i, d);
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool wave_insertable(const Batch& Bt, const Params& P, int e) {
  return !(!P.dynamic && Bt.reject[e]) && !Bt.dup[e];
}

// dynamic shared memory of k_round_decide: near list (node, slot, parent-choice key, flags)
__host__ __device__ inline size_t decide_smem_bytes(int near_cap) { return (size_t)near_cap * (8 + 4 + 4 + 1) + 16; }

// block-wide sum / OR over ROUND_THREADS threads (s_w: ROUND_THREADS / 32 ints)
__device__ __forceinline__ int block_sum(int v, int* s_w) {
  for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(FULL, v, off);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = v;
  __syncthreads();
  int tot = 0;
#pragma unroll
  for (int w = 0; w < ROUND_THREADS / 32; ++w) tot += s_w[w];
  return tot;
}

// One CTA per sample.  Phase 1 (always): steering of every candidate pair, both directions, one thread per
// (candidate, direction).  Phase 2 (wave rounds): the sample's expansion against the frozen tree + conflict marks.
__global__ void __launch_bounds__(ROUND_THREADS) k_round_decide(const Ctl* ctl, Params P, TreeView T, Batch Bt,
                                                                  Wave W) {
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ double4 s_obs[KRRT_MAX_OBS_DEV];
  __shared__ int s_w[ROUND_THREADS / 32];
  __shared__ int s_K, s_flags, s_nrew, s_tail, s_par_k;
  __shared__ unsigned long long s_din, s_dout;
  const int b = blockIdx.x;
  const int nb = ctl->nb;
  if (b >= nb) return;
  const int t = threadIdx.x;
  const int cap = Bt.near_cap;
  const int cnt_raw = Bt.count[b];
  const int cnt = min(cnt_raw, cap);
  const State4 s = Bt.samples[b];
  const size_t base = (size_t)b * cap;

  // ---- phase 1: tau*, c* of candidate -> sample and sample -> candidate
  for (int w = t; w < 2 * cnt; w += ROUND_THREADS) {
    const int e = w >> 1;
    const bool fwd = (w & 1) == 0;
    const int ref = Bt.ref[base + e];
    State4 x;
    if (ref & IN_BATCH) {
      x = Bt.samples[ref & ~IN_BATCH];
    } else {
      x = load_state(T, ref);
    }
    double tau, c;
    steer(fwd ? x : s, fwd ? s : x, P.r, tau, c);
    if (fwd) {
      Bt.tf[base + e] = tau;
      Bt.cf[base + e] = c;
    } else {
      Bt.tb[base + e] = tau;
      Bt.cb[base + e] = c;
    }
  }
  if (!ctl->wave) return;

  // ---- phase 2: decision
  double* s_key = reinterpret_cast<double*>(smem);
  int* s_idx = reinterpret_cast<int*>(s_key + cap);
  int* s_slot = s_idx + cap;
  unsigned char* s_flag = reinterpret_cast<unsigned char*>(s_slot + cap);  // bit 0: forward edge collides,
                                                                           // bit 1: rewire candidate, bit 2: its edge collides
  for (int j = t; j < P.m; j += ROUND_THREADS) s_obs[j] = P.obs[j];
  if (t == 0) {
    s_K = 0; s_flags = 0; s_nrew = 0; s_tail = 0; s_par_k = -1;
    s_din = 0ull; s_dout = ~0ull;
  }
  const unsigned round = ctl->round;
  const unsigned key = mark_key(round, b);
  WaveRec r;
  r.outcome = OUT_COMPLEX; r.n_b = 0; r.par = -1; r.park = -1; r.K = 0; r.nrew = 0; r.nwl = 0; r.pad = 0;
  r.ctc_new = 0.0; r.tim_new = 0.0; r.collide_calls = 0; r.d2_in_max = 0.0; r.d2_out_min = INFINITY;
  if (!P.dynamic && Bt.reject[b]) {
    r.outcome = OUT_REJECT;
    if (t == 0) W.rec[b] = r;
    return;
  }
  // tree size this sample sees if every earlier sample that can insert does insert
  int ins = 0;
  for (int e = t; e < b; e += ROUND_THREADS) ins += wave_insertable(Bt, P, e) ? 1 : 0;
  ins = block_sum(ins, s_w);  // (also orders the s_* initialisation and phase-1 stores before what follows)
  const int n_b = ctl->n0 + ins;
  r.n_b = n_b;
  if (n_b >= P.capacity || cnt_raw > cap) {  // tree full / near overflow: the sequential path reports them
    if (t == 0) W.rec[b] = r;
    return;
  }
  // near set at the exact radius of tree size n_b (order within the list is irrelevant: ties go to node indices)
  const double rho2n = P.rho2[n_b];
  for (int e = t; e < cnt; e += ROUND_THREADS) {
    const int ref = Bt.ref[base + e];
    const double d = Bt.d2[base + e];
    if (ref & IN_BATCH) {
      if (wave_insertable(Bt, P, ref & ~IN_BATCH)) atomicOr(&s_flags, 2);
    } else if (d <= rho2n) {
      if (d == 0.0) atomicOr(&s_flags, 1);
      atomicMax(&s_din, (unsigned long long)__double_as_longlong(d));
      const int k = atomicAdd(&s_K, 1);
      s_idx[k] = ref;
      s_slot[k] = e;
    } else {
      atomicMin(&s_dout, (unsigned long long)__double_as_longlong(d));
    }
  }
  __syncthreads();
  const int K = s_K;
  if (s_flags & 1) {
    r.outcome = OUT_DUP;
    if (t == 0) W.rec[b] = r;
    return;
  }
  if (s_flags & 2) {  // an earlier sample of the batch may land inside the radius: wait for the next round
    r.outcome = OUT_CONFLICT;
    if (t == 0) W.rec[b] = r;
    return;
  }
  r.d2_in_max = __longlong_as_double((long long)s_din);
  r.d2_out_min = s_dout == ~0ull ? INFINITY : __longlong_as_double((long long)s_dout);
  if (K == 0) {  // nearest-node fallback scans the whole tree: sequential path
    if (t == 0) W.rec[b] = r;
    return;
  }
  r.K = K;
  int* g_near = W.near_idx + base;
  for (int k = t; k < K; k += ROUND_THREADS) {
    const int i = s_idx[k];
    g_near[k] = i;
    s_key[k] = T.ctc[i] + Bt.cf[base + s_slot[k]];
    s_flag[k] = 0;
    atomicMax(&W.amin[i], key);  // read mark
  }
  __syncthreads();
  // forward edges of ALL near nodes at once: one thread per (edge, sample point)
  const int S = P.S;
  for (int w = t; w < K * S; w += ROUND_THREADS) {
    const int k = w / S, p = w - k * S;
    const int i = s_idx[k];
    const Edge ed = make_edge(load_state(T, i), s, Bt.tf[base + s_slot[k]], T.tim[i]);
    bool h = false;
    for (int j = 0; j < P.m; ++j) h |= edge_point_hits(ed, p, S, s_obs[j], P.dynamic, P.R2);
    if (h) atomicOr(reinterpret_cast<unsigned*>(s_flag + (k & ~3)), 1u << (8 * (k & 3)));
  }
  __syncthreads();
  // parent = arg-min (cost, node index) over collision-free near nodes
  {
    int bk = -1;
    double bc = INFINITY;
    int bi = 0x7fffffff;
    for (int k = t; k < K; k += ROUND_THREADS) {
      if (s_flag[k] & 1) continue;
      const double c = s_key[k];
      const int i = s_idx[k];
      if (c < bc || (c == bc && i < bi)) {
        bc = c;
        bi = i;
        bk = k;
      }
    }
    // warp arg-min, then one candidate per warp through shared memory
    int aux = bk;
    int idx = bk >= 0 ? bi : -1;
    warp_argmin(bc, idx, aux);
    __shared__ double s_wc[ROUND_THREADS / 32];
    __shared__ int s_wi[ROUND_THREADS / 32], s_wk[ROUND_THREADS / 32];
    if ((t & 31) == 0) {
      s_wc[t >> 5] = bc;
      s_wi[t >> 5] = idx;
      s_wk[t >> 5] = aux;
    }
    __syncthreads();
    if (t < 32) {
      double c2 = t < ROUND_THREADS / 32 ? s_wc[t] : INFINITY;
      int i2 = t < ROUND_THREADS / 32 ? s_wi[t] : -1;
      int k2 = t < ROUND_THREADS / 32 ? s_wk[t] : -1;
      warp_argmin(c2, i2, k2);
      if (t == 0) s_par_k = i2 >= 0 ? k2 : -1;
    }
    __syncthreads();
  }
  const int park = s_par_k;
  if (park < 0) {
    r.outcome = OUT_NOFREE;
    r.collide_calls = K;  // a sequential planner tries every candidate before giving up
    if (t == 0) W.rec[b] = r;
    return;
  }
  const int par = s_idx[park];
  const double key_par = s_key[park];
  const double ctc_new = T.ctc[par] + Bt.cf[base + s_slot[park]];
  const double tim_new = T.tim[par] + Bt.tf[base + s_slot[park]];
  r.par = par;
  r.park = s_slot[park];
  r.ctc_new = ctc_new;
  r.tim_new = tim_new;
  // a sequential planner checks candidates in (cost, index) order until the first free one; rewire candidates:
  // near nodes that would get cheaper through the new node
  int tries = 0, ncand = 0;
  for (int k = t; k < K; k += ROUND_THREADS) {
    const int i = s_idx[k];
    const double c = s_key[k];
    if (c < key_par || (c == key_par && i <= par)) ++tries;  // all of these but the parent collide
    if (i != par && (ctc_new + Bt.cb[base + s_slot[k]] < T.ctc[i])) {
      s_flag[k] |= 2;
      ++ncand;
    }
  }
  tries = block_sum(tries, s_w);
  ncand = block_sum(ncand, s_w);
  r.collide_calls = tries + ncand;
  // edges new -> candidate, started at the new node's arrival time
  for (int w = t; w < K * S; w += ROUND_THREADS) {
    const int k = w / S, p = w - k * S;
    if (!(s_flag[k] & 2)) continue;
    const int i = s_idx[k];
    const Edge ed = make_edge(s, load_state(T, i), Bt.tb[base + s_slot[k]], tim_new);
    bool h = false;
    for (int j = 0; j < P.m; ++j) h |= edge_point_hits(ed, p, S, s_obs[j], P.dynamic, P.R2);
    if (h) atomicOr(reinterpret_cast<unsigned*>(s_flag + (k & ~3)), 4u << (8 * (k & 3)));
  }
  __syncthreads();
  int* g_rn = W.rew_node + base;
  int* g_rs = W.rew_slot + base;
  int* wl = W.wl + (size_t)b * WAVE_QCAP;
  for (int k = t; k < K; k += ROUND_THREADS) {
    if ((s_flag[k] & 6) == 2) {  // candidate and free: rewired.  Its children seed the write list
      const int q = atomicAdd(&s_nrew, 1);
      const int j = s_idx[k];
      g_rn[q] = j;
      g_rs[q] = s_slot[k];
      atomicMax(&W.wmin[j], key);
      for (int c = T.head[j], guard = 0; c >= 0 && guard < T.cap; c = T.next[c], ++guard) {
        const int pos = atomicAdd(&s_tail, 1);
        if (pos < WAVE_QCAP) wl[pos] = c;
      }
    }
  }
  __syncthreads();
  r.nrew = s_nrew;
  // level-order walk of everything below the rewired nodes: write marks
  int level0 = 0;
  bool overflow = false;
  for (;;) {
    const int level1 = s_tail;
    if (level1 > WAVE_QCAP) {
      overflow = true;
      break;
    }
    if (level0 == level1) break;
    __syncthreads();  // every thread has read s_tail before the next level pushes
    for (int q = level0 + t; q < level1; q += ROUND_THREADS) {
      const int d = wl[q];
      atomicMax(&W.wmin[d], key);
      atomicMax(&W.amin[d], key);
      atomicMax(&W.dsc[d], key);
      for (int c = T.head[d], guard = 0; c >= 0 && guard < T.cap; c = T.next[c], ++guard) {
        const int pos = atomicAdd(&s_tail, 1);
        if (pos < WAVE_QCAP) wl[pos] = c;
      }
    }
    level0 = level1;
    __syncthreads();
  }
  r.nwl = overflow ? 0 : level0;
  // a near node below one of this sample's own rewired nodes would see a changed cost mid-expansion
  int intra = 0;
  if (!overflow) {
    __threadfence();
    for (int k = t; k < K; k += ROUND_THREADS) intra |= mark_sample(__ldcg(&W.dsc[s_idx[k]]), round) == b ? 1 : 0;
  }
  intra = block_sum(intra, s_w);
  r.outcome = (overflow || intra) ? OUT_COMPLEX : OUT_INSERT;
  if (t == 0) W.rec[b] = r;
}

// ---------------------------------------------------------------------------------------------------
// One warp per sample: is the decision still the sequential one given everything earlier samples do?
__global__ void __launch_bounds__(128) k_round_validate(Ctl* ctl, Params P, Batch Bt, Wave W) {
  if (!ctl->wave) return;
  const int nb = ctl->nb;
  const int lane = lane_id();
  const int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (b >= nb) return;
  const WaveRec r = W.rec[b];
  if (r.outcome == OUT_REJECT || r.outcome == OUT_DUP) return;  // depend on nothing that can change
  if (r.outcome == OUT_COMPLEX || r.outcome == OUT_CONFLICT) {
    if (lane == 0) atomicMin(&ctl->wave_F, b);
    return;
  }
  const unsigned round = ctl->round;
  const size_t base = (size_t)b * Bt.near_cap;
  bool conf = false;
  // the node index / tree size this sample really gets: earlier samples that found no free parent did not insert.
  // The decision was taken at the radius of r.n_b; it stands if no candidate lies between the two radii.
  int ins = 0;
  for (int e = lane; e < b; e += 32) ins += W.rec[e].outcome == OUT_INSERT ? 1 : 0;
  for (int off = 16; off > 0; off >>= 1) ins += __shfl_xor_sync(FULL, ins, off);
  const int n_true = ctl->n0 + ins;
  if (n_true != r.n_b) {
    const double rt = P.rho2[n_true];
    conf = !(r.d2_in_max <= rt && r.d2_out_min > rt);
    if (lane == 0) W.rec[b].n_b = n_true;
  }
  const int* g_near = W.near_idx + base;
  for (int k = lane; k < r.K; k += 32) conf |= mark_sample(__ldcg(&W.wmin[g_near[k]]), round) < b;
  if (r.outcome == OUT_INSERT) {
    const int* g_rn = W.rew_node + base;
    const int* wl = W.wl + (size_t)b * WAVE_QCAP;
    for (int q = lane; q < r.nrew; q += 32) conf |= mark_sample(__ldcg(&W.amin[g_rn[q]]), round) < b;
    for (int q = lane; q < r.nwl; q += 32) conf |= mark_sample(__ldcg(&W.amin[wl[q]]), round) < b;
  }
  if (__any_sync(FULL, conf) && lane == 0) atomicMin(&ctl->wave_F, b);
}

// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ void list_lock(int* lock, int p) {
  while (atomicCAS(&lock[p], 0, 1) != 0) __nanosleep(20);
  __threadfence();
}
__device__ __forceinline__ void list_unlock(int* lock, int p) {
  __threadfence();
  atomicExch(&lock[p], 0);
}

constexpr int APPLY_THREADS = 128;

// CTA b < B: sample b of the validated prefix — insert, rewire, level-order propagation.
// CTA B: the round's finisher (one warp): counters + control block, or the sequential commit.
__global__ void __launch_bounds__(APPLY_THREADS) k_round_apply(Ctl* ctl, Params P, TreeView T, Batch Bt, Wave W) {
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ int s_tail;
  const int t = threadIdx.x;
  if ((int)blockIdx.x == Bt.B) {
    if (t < 32) round_finish(ctl, P, T, Bt, W, smem);
    return;
  }
  if (!ctl->wave) return;
  const int F = ctl->wave_F;
  const int b = blockIdx.x;
  if (b >= F) return;
  const WaveRec r = W.rec[b];
  if (r.outcome != OUT_INSERT) return;
  const size_t base = (size_t)b * Bt.near_cap;
  const int nw = r.n_b;
  const int nrew = r.nrew;
  const int* g_rn = W.rew_node + base;
  const int* g_rs = W.rew_slot + base;
  int* wl = W.wl + (size_t)b * WAVE_QCAP;
  if (t == 0) {
    const State4 s = Bt.samples[b];
    const size_t op = base + r.park;
    T.pos[nw] = make_double2(s.x, s.y);
    T.vel[nw] = make_double2(s.vx, s.vy);
    T.parent[nw] = r.par;
    T.ecost[nw] = Bt.cf[op];
    T.etime[nw] = Bt.tf[op];
    T.ctc[nw] = r.ctc_new;
    T.tim[nw] = r.tim_new;
    T.head[nw] = nrew > 0 ? g_rn[0] : -1;  // the rewired nodes become the new node's children, in list order
    list_lock(W.lock, r.par);
    link_child(T, r.par, nw);
    list_unlock(W.lock, r.par);
    Bt.map[b] = nw;
    s_tail = 0;
  }
  if (nrew == 0) return;
  // detach each rewired node from its old parent (one thread per warp takes the lock: no spinning inside a warp)
  if ((t & 31) == 0) {
    for (int q = t >> 5; q < nrew; q += APPLY_THREADS / 32) {
      const int j = g_rn[q];
      const int old = T.parent[j];
      list_lock(W.lock, old);
      unlink_child(T, j);
      list_unlock(W.lock, old);
    }
  }
  __syncthreads();
  for (int q = t; q < nrew; q += APPLY_THREADS) {
    const int j = g_rn[q];
    const double cb = Bt.cb[base + g_rs[q]], tb = Bt.tb[base + g_rs[q]];
    T.parent[j] = nw;
    T.next[j] = q + 1 < nrew ? g_rn[q + 1] : -1;
    T.prev[j] = q > 0 ? g_rn[q - 1] : -1;
    T.ecost[j] = cb;
    T.etime[j] = tb;
    T.ctc[j] = r.ctc_new + cb;
    T.tim[j] = r.tim_new + tb;
    for (int c = T.head[j], guard = 0; c >= 0 && guard < T.cap; c = T.next[c], ++guard) {
      const int pos = atomicAdd(&s_tail, 1);
      if (pos < WAVE_QCAP) wl[pos] = c;
    }
  }
  __syncthreads();
  // level order: a node is updated one level after its parent (same walk as the decision's, so it fits WAVE_QCAP)
  int level0 = 0;
  for (;;) {
    const int level1 = min(s_tail, WAVE_QCAP);
    if (level0 == level1) break;
    __syncthreads();
    for (int q = level0 + t; q < level1; q += APPLY_THREADS) {
      const int d = wl[q];
      const int par = T.parent[d];
      T.ctc[d] = __ldcg(&T.ctc[par]) + T.ecost[d];
      T.tim[d] = __ldcg(&T.tim[par]) + T.etime[d];
      for (int c = T.head[d], guard = 0; c >= 0 && guard < T.cap; c = T.next[c], ++guard) {
        const int pos = atomicAdd(&s_tail, 1);
        if (pos < WAVE_QCAP) wl[pos] = c;
      }
    }
    level0 = level1;
    __threadfence_block();
    __syncthreads();
  }
}

}  // namespace krrt
