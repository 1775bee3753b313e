// krrt_wave.cuh — parallel, validated commit of a prefix of a speculative batch ("wave round").
//
// RRT* is sequential: expansion k+1 sees the tree after k.  But an expansion only READS the cost/time of its
// near set and only WRITES the cost/time of the nodes it rewires and their descendants.  So every sample of the
// batch is decided in parallel against the tree as it stands (k_wave_decide), each decision marks what it read and
// what it would write, and k_wave_validate finds the longest prefix [0, F) of the batch in which no sample touches
// anything an earlier sample writes (and no earlier sample touches anything it writes), no earlier sample's new
// node falls inside its near radius, and its near set is the same at the radius of the tree size it really sees.  For that
// prefix the sequential outcome IS the speculated one, so k_wave_apply commits it in parallel; the batch restarts
// at sample F in the next round (as sample 0, which by construction has nothing before it).  A sample that cannot
// be decided this way at all (empty near set -> global nearest, a rewired subtree larger than WAVE_QCAP, a near
// node inside one of its own rewired subtrees) is committed alone by the sequential warp (k_commit).
// The result is bit-identical to the in-order commit whatever the prefix lengths are; only throughput depends on
// them (a 20 k-node tree commits ~20-30 expansions per round, a 100 k-node tree more: conflicts ~ K/N per pair).
//
// Stands in for the inside of RRTStar::grow (robosoccer_motion_planning_node.cpp:177), like krrt_warp.cuh.
#pragma once
#include "krrt_kernels.cuh"

namespace krrt {

__device__ __forceinline__ bool wave_insertable(const Batch& Bt, const Params& P, int e) {
  return !(!P.dynamic && Bt.reject[e]) && !Bt.dup[e];
}

// shared memory of one deciding warp: the near list with the parent-choice keys
__host__ __device__ inline size_t wave_warp_smem_bytes(int near_cap) { return (size_t)near_cap * 16; }

// ---------------------------------------------------------------------------------------------------
// One warp per sample: decide against the frozen tree, then mark the read and write sets.
__global__ void __launch_bounds__(128) k_wave_decide(const Ctl* ctl, Params P, TreeView T, Batch Bt, Wave W,
                                                       int warps_per_cta) {
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ double4 s_obs[KRRT_MAX_OBS_DEV];
  if (!ctl->wave) return;
  const int nb = ctl->nb;
  const int lane = lane_id(), warp = threadIdx.x >> 5;
  for (int j = threadIdx.x; j < P.m; j += blockDim.x) s_obs[j] = P.obs[j];
  __syncthreads();
  const int b = blockIdx.x * warps_per_cta + warp;
  if (warp >= warps_per_cta || b >= nb) return;
  const int cap = Bt.near_cap;
  double* s_key = reinterpret_cast<double*>(smem + (size_t)warp * wave_warp_smem_bytes(cap));
  int* s_idx = reinterpret_cast<int*>(s_key + cap);
  int* s_slot = s_idx + cap;
  EnvView E = {s_obs, P.m, P.dynamic, P.R2, P.S};
  const unsigned round = ctl->round;
  const unsigned key = mark_key(round, b);

  WaveRec r;
  r.outcome = OUT_COMPLEX; r.n_b = 0; r.par = -1; r.park = -1; r.K = 0; r.nrew = 0; r.nwl = 0; r.pad = 0;
  r.ctc_new = 0.0; r.tim_new = 0.0; r.collide_calls = 0; r.d2_in_max = 0.0; r.d2_out_min = INFINITY;
  const State4 s = Bt.samples[b];
  if (!P.dynamic && Bt.reject[b]) {
    r.outcome = OUT_REJECT;
    if (lane == 0) W.rec[b] = r;
    return;
  }
  // tree size this sample will see if every earlier sample that can insert does insert
  int ins = 0;
  for (int e = lane; e < b; e += 32) ins += wave_insertable(Bt, P, e) ? 1 : 0;
  for (int off = 16; off > 0; off >>= 1) ins += __shfl_xor_sync(FULL, ins, off);
  const int n_b = ctl->n0 + ins;
  r.n_b = n_b;
  const int cnt = Bt.count[b];
  if (n_b >= P.capacity || cnt > cap) {  // tree full / near overflow: the sequential path reports them
    if (lane == 0) W.rec[b] = r;
    return;
  }
  // near set at the exact radius of tree size n_b; in-batch neighbours make the sample wait for the next round
  const double rho2n = P.rho2[n_b];
  int K = 0;
  bool dup = false, inbatch = false;
  double din = 0.0, dout = INFINITY;
  for (int base = 0; base < cnt; base += 32) {
    const int e = base + lane;
    bool valid = false;
    int ref = -1;
    if (e < cnt) {
      const size_t o = (size_t)b * cap + e;
      ref = Bt.ref[o];
      const double d = Bt.d2[o];
      if (ref & IN_BATCH) {
        inbatch |= wave_insertable(Bt, P, ref & ~IN_BATCH);
      } else {
        valid = d <= rho2n;
        dup |= d == 0.0;
        if (valid) din = fmax(din, d); else dout = fmin(dout, d);
      }
    }
    const unsigned mk = __ballot_sync(FULL, valid);
    if (valid) {
      const int k = K + __popc(mk & ((1u << lane) - 1));
      s_idx[k] = ref;
      s_slot[k] = e;
    }
    K += __popc(mk);
  }
  if (__any_sync(FULL, dup)) {
    r.outcome = OUT_DUP;
    if (lane == 0) W.rec[b] = r;
    return;
  }
  if (__any_sync(FULL, inbatch)) {
    r.outcome = OUT_CONFLICT;
    if (lane == 0) W.rec[b] = r;
    return;
  }
  for (int off = 16; off > 0; off >>= 1) {
    din = fmax(din, __shfl_xor_sync(FULL, din, off));
    dout = fmin(dout, __shfl_xor_sync(FULL, dout, off));
  }
  r.d2_in_max = din;
  r.d2_out_min = dout;
  if (K == 0 || K > 1024) {  // nearest-node fallback scans the whole tree (and 1024 = 32 lanes x 32 tried bits): sequential path
    if (lane == 0) W.rec[b] = r;
    return;
  }
  __syncwarp();
  r.K = K;
  int* g_near = W.near_idx + (size_t)b * cap;
  for (int k = lane; k < K; k += 32) {
    const int i = s_idx[k];
    g_near[k] = i;
    s_key[k] = T.ctc[i] + Bt.cf[(size_t)b * cap + s_slot[k]];
  }
  __syncwarp();
  // parent = arg-min over collision-free candidates, lowest node index on ties (as warp_commit)
  int par = -1, park = -1;
  unsigned tried = 0;  // bit i: candidate lane + 32 i
  for (;;) {
    double bc = INFINITY;
    int bi = -1, bk = -1;
    for (int k = lane, i = 0; k < K; k += 32, ++i) {
      if ((tried >> i) & 1u) continue;
      const int node = s_idx[k];
      const double cand = s_key[k];
      if (cand < bc || (bi >= 0 && cand == bc && node < bi)) {
        bc = cand;
        bi = node;
        bk = k;
      }
    }
    warp_argmin(bc, bi, bk);
    if (bi < 0) break;
    if ((bk & 31) == lane) tried |= 1u << (bk >> 5);
    r.collide_calls++;
    const State4 a = load_state(T, bi);
    if (!warp_edge_collide(a, s, Bt.tf[(size_t)b * cap + s_slot[bk]], T.tim[bi], E)) {
      par = bi;
      park = bk;
      break;
    }
  }
  if (par < 0) {
    r.outcome = OUT_NOFREE;
    if (lane == 0) W.rec[b] = r;
    return;
  }
  const size_t op = (size_t)b * cap + s_slot[park];
  const double ctc_new = T.ctc[par] + Bt.cf[op];
  const double tim_new = T.tim[par] + Bt.tf[op];
  r.par = par;
  r.park = s_slot[park];
  r.ctc_new = ctc_new;
  r.tim_new = tim_new;
  // rewire set (each test against the frozen cost: valid unless one rewired node sits below another, checked below)
  int nrew = 0;
  int* g_rn = W.rew_node + (size_t)b * cap;
  int* g_rs = W.rew_slot + (size_t)b * cap;
  for (int base = 0; base < K; base += 32) {
    const int k = base + lane;
    bool cand = false;
    int j = -1, slot = -1;
    if (k < K) {
      j = s_idx[k];
      slot = s_slot[k];
      cand = j != par && (ctc_new + Bt.cb[(size_t)b * cap + slot] < T.ctc[j]);
    }
    unsigned mk = __ballot_sync(FULL, cand);
    while (mk) {
      const int src = __ffs(mk) - 1;
      mk &= mk - 1;
      const int jj = __shfl_sync(FULL, j, src), ss = __shfl_sync(FULL, slot, src);
      r.collide_calls++;
      const State4 xj = load_state(T, jj);
      if (!warp_edge_collide(s, xj, Bt.tb[(size_t)b * cap + ss], tim_new, E)) {
        if (lane == 0) {
          g_rn[nrew] = jj;
          g_rs[nrew] = ss;
        }
        ++nrew;
      }
    }
  }
  r.nrew = nrew;
  __syncwarp();
  // marks: reads (whole near set), writes (rewired nodes and everything below them)
  for (int k = lane; k < K; k += 32) atomicMax(&W.amin[s_idx[k]], key);
  int* wl = W.wl + (size_t)b * WAVE_QCAP;
  int qtail = 0;
  bool overflow = false;
  for (int q = 0; q < nrew && !overflow; ++q) {
    const int j = g_rn[q];
    if (lane == 0) atomicMax(&W.wmin[j], key);
    int qhead = qtail;
    {  // children of j
      int c = (lane == 0) ? T.head[j] : -1;
      if (lane == 0) {
        while (c >= 0 && qtail < WAVE_QCAP) {
          wl[qtail++] = c;
          c = T.next[c];
        }
        if (c >= 0) qtail = WAVE_QCAP + 1;
      }
      qtail = __shfl_sync(FULL, qtail, 0);
      if (qtail > WAVE_QCAP) {
        overflow = true;
        break;
      }
    }
    __syncwarp();
    while (qhead < qtail) {
      const int n_now = min(32, qtail - qhead);
      int c = -1;
      if (lane < n_now) {
        const int d = wl[qhead + lane];
        atomicMax(&W.wmin[d], key);
        atomicMax(&W.amin[d], key);
        atomicMax(&W.dsc[d], key);
        c = T.head[d];
      }
      qhead += n_now;
      for (;;) {
        const unsigned m = __ballot_sync(FULL, c >= 0);
        if (m == 0) break;
        if (qtail + 32 > WAVE_QCAP) {
          overflow = true;
          break;
        }
        if (c >= 0) {
          wl[qtail + __popc(m & ((1u << lane) - 1))] = c;
          c = T.next[c];
        }
        qtail += __popc(m);
      }
      __syncwarp();
      if (overflow) break;
    }
  }
  r.nwl = qtail;
  bool intra = false;
  if (!overflow) {
    __threadfence();
    for (int k = lane; k < K; k += 32) intra |= mark_sample(__ldcg(&W.dsc[s_idx[k]]), round) == b;
  }
  r.outcome = (overflow || __any_sync(FULL, intra)) ? OUT_COMPLEX : OUT_INSERT;
  if (lane == 0) W.rec[b] = r;
}

// ---------------------------------------------------------------------------------------------------
// One warp per sample: is the decision still the sequential one given everything earlier samples do?
__global__ void __launch_bounds__(128) k_wave_validate(Ctl* ctl, Params P, Batch Bt, Wave W) {
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
  const int cap = Bt.near_cap;
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
  const int* g_near = W.near_idx + (size_t)b * cap;
  for (int k = lane; k < r.K; k += 32) conf |= mark_sample(__ldcg(&W.wmin[g_near[k]]), round) < b;
  if (r.outcome == OUT_INSERT) {
    const int* g_rn = W.rew_node + (size_t)b * cap;
    const int* wl = W.wl + (size_t)b * WAVE_QCAP;
    for (int q = lane; q < r.nrew; q += 32) conf |= mark_sample(__ldcg(&W.amin[g_rn[q]]), round) < b;
    for (int q = lane; q < r.nwl; q += 32) conf |= mark_sample(__ldcg(&W.amin[wl[q]]), round) < b;
  }
  conf = __any_sync(FULL, conf);
  if (lane == 0) {
    if (conf) atomicMin(&ctl->wave_F, b);
  }
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

// One warp per sample of the validated prefix: insert, rewire, propagate.
__global__ void __launch_bounds__(128) k_wave_apply(const Ctl* ctl, TreeView T, Batch Bt, Wave W) {
  if (!ctl->wave) return;
  const int F = ctl->wave_F;
  const int lane = lane_id();
  const int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (b >= F) return;
  const WaveRec r = W.rec[b];
  if (r.outcome != OUT_INSERT) return;
  const int cap = Bt.near_cap;
  const int nw = r.n_b;
  const State4 s = Bt.samples[b];
  if (lane == 0) {
    const size_t op = (size_t)b * cap + r.park;
    T.pos[nw] = make_double2(s.x, s.y);
    T.vel[nw] = make_double2(s.vx, s.vy);
    T.parent[nw] = r.par;
    T.ecost[nw] = Bt.cf[op];
    T.etime[nw] = Bt.tf[op];
    T.ctc[nw] = r.ctc_new;
    T.tim[nw] = r.tim_new;
    T.head[nw] = -1;
    list_lock(W.lock, r.par);
    link_child(T, r.par, nw);
    list_unlock(W.lock, r.par);
    Bt.map[b] = nw;
  }
  __syncwarp();
  if (r.nrew == 0) return;
  TreeView Tb = T;
  Tb.queue = W.wl + (size_t)b * WAVE_QCAP;  // private BFS scratch (the decision's list is no longer needed)
  Tb.cap = WAVE_QCAP;
  const int* g_rn = W.rew_node + (size_t)b * cap;
  const int* g_rs = W.rew_slot + (size_t)b * cap;
  for (int q = 0; q < r.nrew; ++q) {
    const int j = g_rn[q];
    if (lane == 0) {
      const size_t o = (size_t)b * cap + g_rs[q];
      const double cb = Bt.cb[o], tb = Bt.tb[o];
      const int old = T.parent[j];
      list_lock(W.lock, old);
      unlink_child(T, j);
      list_unlock(W.lock, old);
      T.parent[j] = nw;
      link_child(T, nw, j);  // nw's list is private to this warp during the round
      T.ecost[j] = cb;
      T.etime[j] = tb;
      T.ctc[j] = r.ctc_new + cb;
      T.tim[j] = r.tim_new + tb;
    }
    __syncwarp();
    warp_propagate(Tb, j);
  }
}

}  // namespace krrt
