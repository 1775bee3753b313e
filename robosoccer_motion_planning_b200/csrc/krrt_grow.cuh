// krrt_grow.cuh — k_grow: ONE persistent cooperative kernel per krrt_grow call (one CTA per SM, grid barriers
// instead of launches).  Stands in for the inside of `for(i<iteration) rrt.grow(xg)`
// (robosoccer_motion_planning_node.cpp:176-185).
//
// WINDOW (up to Bt.B future samples; samples are counter-based, so they are known in advance)
//   scan    all CTAs: draw the window (Philox), static rejection, every frozen node x every window sample
//           (4-D distance, radius test, candidate append), window-sample pairs            [HBM/L2 bytes, fp64]
//   steer   all CTAs: tau*/c* of both directions of every candidate pair — tree-independent, so it is done once
//           per window at full-machine parallelism, off the commit critical path               [fp64 pipe]
// SUB-ROUNDS until the window is used up (three grid barriers each)
//   decide    one CTA per sample, one thread per candidate: the sample's whole expansion against the tree as it
//             stands (near set at the exact radius, parent = arg-min over collision-free candidates, rewire set,
//             descendants of the rewired nodes) + conflict marks on what it reads / would write
//   validate  same CTA: is that decision what a sequential planner would take, given everything earlier samples of
//             the sub-round do?  -> longest valid prefix F
//   apply     same CTA, for the prefix: insert, rewire, level-order cost/time propagation (lists kept in shared
//             memory since decide); the last CTA folds the counters / runs the sequential fall-back
//
// Why the prefix is exact.  RRT* is sequential: expansion k+1 sees the tree after k.  But an expansion only READS
// the cost/time of its near set and only WRITES the cost/time of the nodes it rewires and of their descendants.
// Each decision marks both sets; sample b stays in the prefix iff no earlier sample writes anything b reads or
// writes, no earlier sample reads anything b writes, no earlier uncommitted sample's new node can fall inside b's
// near radius, and b's near set is unchanged at the radius of the tree size it really sees (earlier samples that
// found no collision-free parent did not insert).  For such a prefix the sequential outcome IS the speculated one.
// A sample that cannot be decided this way (empty near set -> global nearest, a rewired subtree larger than
// WAVE_QCAP, a near node inside one of its own rewired subtrees) is committed alone by one warp; while the tree is
// small (conflicts everywhere) whole windows are.  Trees are bit-identical to the in-order commit whatever the
// prefix lengths are.
#pragma once
#include "krrt_kernels.cuh"

namespace krrt {

constexpr int GROW_THREADS = 256;
constexpr int GROW_WMAX = 512;    // window cap
constexpr int GROW_LEVELS = 62;   // BFS depth below the rewired nodes that the parallel path handles
constexpr int GROW_SEQ_WINDOW = 64;

struct GrowCfg {
  int wave_enabled;
  int wave_min_n;
  const double* user_samples;  // device copy of the caller's explicit samples, or nullptr
  unsigned long long* bar;     // grid barrier counter (zero at launch)
};

__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// all CTAs are co-resident (cooperative launch, one per SM): arrive with a release, spin with an acquire
__device__ __forceinline__ void grid_barrier(unsigned long long* bar, unsigned long long& target) {
  __syncthreads();
  target += gridDim.x;
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(bar, 1ull);
    while (ld_acquire_u64(bar) < target) {
    }
    __threadfence();
  }
  __syncthreads();
}

__device__ __forceinline__ void append_candidate(const Batch& Bt, int b, int ref, double d) {
  const int slot = atomicAdd(&Bt.count[b], 1);
  if (slot < Bt.near_cap) {
    Bt.ref[(size_t)b * Bt.near_cap + slot] = ref;
    Bt.d2[(size_t)b * Bt.near_cap + slot] = d;
  }
}

__device__ __forceinline__ bool sample_insertable(const Batch& Bt, const Params& P, int e) {
  return !(!P.dynamic && __ldcg(&Bt.reject[e])) && !__ldcg(&Bt.dup[e]);
}

__device__ __forceinline__ State4 load_state_cg(const TreeView& T, int i) {
  const double2 p = __ldcg(&T.pos[i]), v = __ldcg(&T.vel[i]);
  State4 s = {p.x, p.y, v.x, v.y};
  return s;
}

// child-list maintenance under the parent's lock, every access at L2 (other SMs edit the same lists)
__device__ __forceinline__ void list_lock(int* lock, int p) {
  while (atomicCAS(&lock[p], 0, 1) != 0) __nanosleep(20);
  __threadfence();
}
__device__ __forceinline__ void list_unlock(int* lock, int p) {
  __threadfence();
  atomicExch(&lock[p], 0);
}
__device__ __forceinline__ void link_child_cg(const TreeView& T, int par, int c) {
  const int h = __ldcg(&T.head[par]);
  __stcg(&T.next[c], h);
  __stcg(&T.prev[c], -1);
  if (h >= 0) __stcg(&T.prev[h], c);
  __stcg(&T.head[par], c);
}
__device__ __forceinline__ void unlink_child_cg(const TreeView& T, int par, int c) {
  const int nx = __ldcg(&T.next[c]), pv = __ldcg(&T.prev[c]);
  if (pv >= 0) __stcg(&T.next[pv], nx); else __stcg(&T.head[par], nx);
  if (nx >= 0) __stcg(&T.prev[nx], pv);
}

struct GrowShared {
  union {
    State4 ss[GROW_WMAX];  // scan: the window's samples
    struct {
      State4 st[GROW_THREADS];   // state of candidate slot k's node
      double tf[GROW_THREADS];   // tau*(node -> sample)
      double tb[GROW_THREADS];   // tau*(sample -> node)
      double tim[GROW_THREADS];  // arrival time of the node
      int flag[GROW_THREADS];    // 1 in near set, 2 forward edge collides, 8 rewire edge collides
      int cand[GROW_THREADS];    // slots of the rewire candidates
      int rewn[GROW_THREADS];    // rewired nodes, in the order they become children of the new node
      int wl[WAVE_QCAP];         // strict descendants of the rewired nodes, level by level
      int lvl[GROW_LEVELS + 2];  // level boundaries in wl
    } d;
  } u;
  double4 obs[KRRT_MAX_OBS_DEV];
  double red_c[GROW_THREADS / 32];
  int red_i[GROW_THREADS / 32], red_k[GROW_THREADS / 32];
  double par_key, par_tim;
  int par_k, par_node;
  int ncand, nrew, tail;
};

// ---------------------------------------------------------------------------------------------------
// In-order commit of window samples [b0, b1) by ONE warp (the whole window while the tree is small; a single sample
// when it needs the sequential path).  smem: obstacles + near list (commit_smem_bytes).
__device__ __noinline__ void seq_commit(Ctl* ctl, const Params& P, const TreeView& T, const Batch& Bt,
                                        unsigned char* smem, int b0, int b1) {
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
  for (int j = lane; j < P.m; j += 32) s_obs[j] = P.obs[j];
  __syncwarp();
  EnvView E = {s_obs, P.m, P.dynamic, P.R2, P.S};

  Stats st = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  int n = __ldcg(&ctl->n);
  long long iter = __ldcg(&ctl->iter);
  int full = 0, error = 0;
  State4 s = {0, 0, 0, 0};
  for (int b = b0; b < b1; ++b) {
    s = Bt.samples[b];
    ++iter;
    st.iters++;
    if (!P.dynamic && __ldcg(&Bt.reject[b])) {
      st.fail_reject++;
      continue;
    }
    if (n >= P.capacity) {
      full = 1;
      break;
    }
    const int cnt = __ldcg(&Bt.count[b]);
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
        const int ref = __ldcg(&Bt.ref[o]);
        const double d = __ldcg(&Bt.d2[o]);
        node = (ref & IN_BATCH) ? __ldcg(&Bt.map[ref & ~IN_BATCH]) : ref;
        valid = node >= 0 && d <= rho2n;
        dup |= valid && d == 0.0;
      }
      unsigned mk = __ballot_sync(FULL, valid);
      if (valid) {
        int k = K + __popc(mk & ((1u << lane) - 1));
        N.idx[k] = node;
        N.cf[k] = __ldcg(&Bt.cf[o]);
        N.tf[k] = __ldcg(&Bt.tf[o]);
        N.cb[k] = __ldcg(&Bt.cb[o]);
        N.tb[k] = __ldcg(&Bt.tb[o]);
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

// candidate lists of the next window start empty
__device__ __forceinline__ void clear_window(const Batch& Bt) {
  for (int b = lane_id(); b < Bt.B; b += 32) {
    Bt.count[b] = 0;
    Bt.dup[b] = 0;
  }
}

// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(GROW_THREADS, 1) k_grow(Ctl* ctl, Params P, TreeView T, Batch Bt, Wave W,
                                                            GrowCfg cfg) {
  extern __shared__ __align__(16) unsigned char dyn_smem[];  // sequential commit scratch (last CTA only)
  __shared__ GrowShared sm;
  const int t = threadIdx.x;
  const int G = gridDim.x;
  const int cta = blockIdx.x;
  const bool last = cta == G - 1;  // never owns a sample: finishes the sub-rounds, commits sequentially
  const int cap = Bt.near_cap;
  const int S = P.S;
  unsigned long long bar_target = 0;
  long long clk_mark = clock64();
  auto phase_done = [&](int which) {  // phase clocks of the last CTA (it waits at every barrier like the others)
    if (last && t == 0) {
      const long long now = clock64();
      ctl->phase_clk[which] += now - clk_mark;
      clk_mark = now;
    }
  };

  for (;;) {
    // ================================================================ window
    const int n0 = __ldcg(&ctl->n);
    const long long iter0 = __ldcg(&ctl->iter);
    const long long left = __ldcg(&ctl->iter_target) - iter0;
    if (__ldcg(&ctl->full) || __ldcg(&ctl->error) || left <= 0 || __ldcg(&ctl->round) >= (1u << 23) - 2) break;
    const bool wave = cfg.wave_enabled && n0 >= cfg.wave_min_n;
    const int Wn = (int)min((long long)(wave ? Bt.B : min(Bt.B, GROW_SEQ_WINDOW)), left);
    const int user_off = (int)(iter0 - __ldcg(&ctl->iter_begin));

    // squared radius that covers every tree size this window can see
    double rmax = 0.0;
    for (int k = t; k <= Wn; k += GROW_THREADS) rmax = fmax(rmax, P.rho2[min(n0 + k, P.capacity)]);
    for (int off = 16; off > 0; off >>= 1) rmax = fmax(rmax, __shfl_xor_sync(FULL, rmax, off));
    if ((t & 31) == 0) sm.red_c[t >> 5] = rmax;
    // the window: sample b belongs to iteration iter0 + b (every CTA draws all of them into its shared memory)
    {
      const bool has_goal = __ldcg(&ctl->has_goal) != 0;
      const State4 g = {ctl->goal[0], ctl->goal[1], ctl->goal[2], ctl->goal[3]};
      for (int b = t; b < Wn; b += GROW_THREADS) {
        State4 s = {0, 0, 0, 0};
        bool drawn = true;
        if (cfg.user_samples) {
          const double* u = cfg.user_samples + 4 * (size_t)(user_off + b);
          if (!(u[0] != u[0])) {
            s.x = u[0]; s.y = u[1]; s.vx = u[2]; s.vy = u[3];
            drawn = false;
          }
        }
        if (drawn) s = draw_sample(P.sp, P.query_id, (uint64_t)(iter0 + b), has_goal, g);
        sm.u.ss[b] = s;
        if (last) {
          bool hit = false;
          if (!P.dynamic) {
            for (int j = 0; j < P.m; ++j) {
              const double dx = s.x - P.obs[j].x, dy = s.y - P.obs[j].y;
              hit |= fma(dy, dy, dx * dx) < P.R2;
            }
          }
          Bt.samples[b] = s;
          Bt.reject[b] = hit ? 1 : 0;
          Bt.map[b] = -1;
        }
      }
    }
    __syncthreads();
    rmax = sm.red_c[0];
#pragma unroll
    for (int w = 1; w < GROW_THREADS / 32; ++w) rmax = fmax(rmax, sm.red_c[w]);
    if (last && t == 0) {
      ctl->n0 = n0;
      ctl->nb = Wn;
      ctl->rho2max = rmax;
      ctl->scan_pairs += (double)n0 * (double)Wn;
      ctl->scan_nodes += (double)n0;
      ctl->windows += 1;
    }
    // ---- scan: item = (tile of GROW_THREADS frozen nodes, chunk of the window's samples)
    {
      const int tiles = (n0 + GROW_THREADS - 1) / GROW_THREADS;
      const int chunks = max(1, min(8, (2 * G + tiles - 1) / tiles));
      const int per = (Wn + chunks - 1) / chunks;
      for (int item = cta; item < tiles * chunks; item += G) {
        const int tile = item / chunks, ch = item - tile * chunks;
        const int i = tile * GROW_THREADS + t;
        const int c0 = ch * per, c1 = min(Wn, c0 + per);
        if (i < n0) {
          const State4 a = load_state_cg(T, i);
#pragma unroll 4
          for (int b = c0; b < c1; ++b) {
            const double d = dist2(a, sm.u.ss[b]);
            if (d <= rmax) {
              if (d == 0.0) Bt.dup[b] = 1;
              append_candidate(Bt, b, i, d);
            }
          }
        }
      }
      // window pairs: an earlier sample of the window may become a node inside a later sample's radius
      for (int p = cta * GROW_THREADS + t; p < Wn * Wn; p += G * GROW_THREADS) {
        const int e = p / Wn, b = p - e * Wn;
        if (e < b) {
          const double d = dist2(sm.u.ss[e], sm.u.ss[b]);
          if (d <= rmax) append_candidate(Bt, b, IN_BATCH | e, d);
        }
      }
    }
    grid_barrier(cfg.bar, bar_target);
    phase_done(0);
    // ---- steer: tau*, c* of candidate -> sample and sample -> candidate, one thread per (candidate, direction)
    for (int b = cta; b < Wn; b += G) {
      const int cnt = min(__ldcg(&Bt.count[b]), cap);
      const State4 s = sm.u.ss[b];
      const size_t base = (size_t)b * cap;
      for (int w = t; w < 2 * cnt; w += GROW_THREADS) {
        const int e = w >> 1;
        const bool fwd = (w & 1) == 0;
        const int ref = __ldcg(&Bt.ref[base + e]);
        const State4 x = (ref & IN_BATCH) ? sm.u.ss[ref & ~IN_BATCH] : load_state_cg(T, ref);
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
    }
    grid_barrier(cfg.bar, bar_target);
    phase_done(1);

    if (!wave) {  // small tree: the whole window in order, one warp
      if (last && t < 32) {
        seq_commit(ctl, P, T, Bt, dyn_smem, 0, Wn);
        __syncwarp();
        clear_window(Bt);
        if (t == 0) ctl->rounds_seq += 1;
      }
      grid_barrier(cfg.bar, bar_target);
      phase_done(5);
      continue;
    }

    // ================================================================ sub-rounds
    for (int j = t; j < P.m; j += GROW_THREADS) sm.obs[j] = P.obs[j];
    int F0 = 0;
    while (F0 < Wn) {
      const int nsub = __ldcg(&ctl->n);
      const unsigned round = __ldcg(&ctl->round);
      if (__ldcg(&ctl->full) || __ldcg(&ctl->error) || round >= (1u << 23) - 2) break;
      const int slot = (int)(round & 1u);
      const int nb = min(min(G - 1, Wn - F0), max(16, (2 * __ldcg(&ctl->ema16)) / 16 + 8));
      const bool active = cta < nb;
      const int b = F0 + cta;
      const size_t base = (size_t)b * cap;
      const unsigned key = mark_key(round, cta);
      // what this CTA carries from decide to validate and apply
      int outcome = OUT_REJECT, n_b = 0, K = 0, par = -1, park = -1, nrew = 0, nwl = 0, nlvl = 0;
      long long collide_calls = 0;
      double ctc_new = 0.0, tim_new = 0.0, rho2n = 0.0;
      int my_node = -1, my_q = -1;
      double my_d = 0.0, my_cf = 0.0, my_cb = 0.0, my_ctc = 0.0;
      bool my_near = false;
      State4 s = {0, 0, 0, 0};

      // -------------------------------------------------------------- decide
      if (active) do {
        s = Bt.samples[b];
        if (t == 0) {
          sm.ncand = 0;
          sm.nrew = 0;
          sm.tail = 0;
          sm.par_k = -1;
        }
        if (!P.dynamic && __ldcg(&Bt.reject[b])) break;  // OUT_REJECT
        outcome = OUT_COMPLEX;
        const int cnt_raw = __ldcg(&Bt.count[b]);
        // tree size this sample sees if every earlier sample of the sub-round that can insert does insert
        const int ins = __syncthreads_count(t < cta && sample_insertable(Bt, P, F0 + t));
        n_b = nsub + ins;
        if (n_b >= P.capacity || cnt_raw > min(cap, GROW_THREADS)) break;  // tree full / overflow / huge near set
        rho2n = P.rho2[n_b];
        bool dup = false, pend = false;
        if (t < cnt_raw) {
          const int ref = __ldcg(&Bt.ref[base + t]);
          my_d = __ldcg(&Bt.d2[base + t]);
          if (ref & IN_BATCH) {
            const int e = ref & ~IN_BATCH;
            if (e < F0) my_node = __ldcg(&Bt.map[e]);        // committed in an earlier sub-round (-1: it did not insert)
            else pend = sample_insertable(Bt, P, e);          // may land inside the radius: wait for the next sub-round
          } else {
            my_node = ref;
          }
          if (my_node >= 0) {
            my_near = my_d <= rho2n;
            dup = my_near && my_d == 0.0;
          }
        }
        const int any_dup = __syncthreads_or(dup);
        const int any_pend = __syncthreads_or(pend);
        K = __syncthreads_count(my_near);
        if (any_dup) {
          outcome = OUT_DUP;
          break;
        }
        if (any_pend) {
          outcome = OUT_CONFLICT;
          break;
        }
        if (K == 0) break;  // nearest-node fallback scans the whole tree: sequential path
        // near set: gathers + read marks
        if (my_near) {
          sm.u.d.st[t] = load_state_cg(T, my_node);
          my_ctc = __ldcg(&T.ctc[my_node]);
          sm.u.d.tim[t] = __ldcg(&T.tim[my_node]);
          my_cf = __ldcg(&Bt.cf[base + t]);
          my_cb = __ldcg(&Bt.cb[base + t]);
          sm.u.d.tf[t] = __ldcg(&Bt.tf[base + t]);
          sm.u.d.tb[t] = __ldcg(&Bt.tb[base + t]);
          atomicMax(&W.amin[my_node], key);
        }
        sm.u.d.flag[t] = my_near ? 1 : 0;
        __syncthreads();
        // forward edges of ALL near nodes at once: one thread per (edge, sample point)
        if (P.m > 0) {
          const int tot = min(cnt_raw, GROW_THREADS) * S;
          for (int w = t; w < tot; w += GROW_THREADS) {
            const int k = w / S, p = w - k * S;
            if (!(sm.u.d.flag[k] & 1)) continue;
            const Edge ed = make_edge(sm.u.d.st[k], s, sm.u.d.tf[k], sm.u.d.tim[k]);
            bool h = false;
            for (int j = 0; j < P.m; ++j) h |= edge_point_hits(ed, p, S, sm.obs[j], P.dynamic, P.R2);
            if (h) atomicOr(&sm.u.d.flag[k], 2);
          }
          __syncthreads();
        }
        // parent = arg-min (cost, node index) over collision-free near nodes
        {
          double bc = INFINITY;
          int bi = -1, bk = -1;
          if (my_near && !(sm.u.d.flag[t] & 2)) {
            bc = my_ctc + my_cf;
            bi = my_node;
            bk = t;
          }
          warp_argmin(bc, bi, bk);
          if ((t & 31) == 0) {
            sm.red_c[t >> 5] = bc;
            sm.red_i[t >> 5] = bi;
            sm.red_k[t >> 5] = bk;
          }
          __syncthreads();
          if (t < 32) {
            double c2 = t < GROW_THREADS / 32 ? sm.red_c[t] : INFINITY;
            int i2 = t < GROW_THREADS / 32 ? sm.red_i[t] : -1;
            int k2 = t < GROW_THREADS / 32 ? sm.red_k[t] : -1;
            warp_argmin(c2, i2, k2);
            if (t == 0 && i2 >= 0) {
              sm.par_k = k2;
              sm.par_node = i2;
              sm.par_key = c2;
              sm.par_tim = sm.u.d.tim[k2] + sm.u.d.tf[k2];
            }
          }
          __syncthreads();
        }
        park = sm.par_k;
        if (park < 0) {
          outcome = OUT_NOFREE;
          collide_calls = K;  // a sequential planner tries every candidate before giving up
          break;
        }
        par = sm.par_node;
        ctc_new = sm.par_key;
        tim_new = sm.par_tim;
        // a sequential planner checks candidates in (cost, index) order until the first free one; rewire candidates:
        // near nodes that would get cheaper through the new node
        bool tried = false, cand = false;
        if (my_near) {
          const double c = my_ctc + my_cf;
          tried = c < ctc_new || (c == ctc_new && my_node <= par);  // all of these but the parent collide
          cand = my_node != par && (ctc_new + my_cb < my_ctc);
        }
        const int tries = __syncthreads_count(tried);
        if (cand) sm.u.d.cand[atomicAdd(&sm.ncand, 1)] = t;
        __syncthreads();
        const int ncand = sm.ncand;
        collide_calls = tries + ncand;
        // edges new -> candidate, started at the new node's arrival time
        if (P.m > 0 && ncand > 0) {
          for (int w = t; w < ncand * S; w += GROW_THREADS) {
            const int q = w / S, p = w - q * S;
            const int k = sm.u.d.cand[q];
            const Edge ed = make_edge(s, sm.u.d.st[k], sm.u.d.tb[k], tim_new);
            bool h = false;
            for (int j = 0; j < P.m; ++j) h |= edge_point_hits(ed, p, S, sm.obs[j], P.dynamic, P.R2);
            if (h) atomicOr(&sm.u.d.flag[k], 8);
          }
          __syncthreads();
        }
        if (cand && !(sm.u.d.flag[t] & 8)) {  // candidate and free: rewired.  Its children seed the write list
          my_q = atomicAdd(&sm.nrew, 1);
          sm.u.d.rewn[my_q] = my_node;
          atomicMax(&W.wmin[my_node], key);
          for (int c = __ldcg(&T.head[my_node]), guard = 0; c >= 0 && guard < T.cap; c = __ldcg(&T.next[c]), ++guard) {
            const int pos = atomicAdd(&sm.tail, 1);
            if (pos < WAVE_QCAP) sm.u.d.wl[pos] = c;
          }
        }
        __syncthreads();
        nrew = sm.nrew;
        // level-order walk of everything below the rewired nodes: write marks
        int level0 = 0;
        bool overflow = false;
        for (;;) {
          const int level1 = sm.tail;
          if (level1 > WAVE_QCAP || (level1 > level0 && nlvl >= GROW_LEVELS)) {
            overflow = true;
            break;
          }
          if (level0 == level1) break;
          if (t == 0) {
            sm.u.d.lvl[nlvl] = level0;
            sm.u.d.lvl[nlvl + 1] = level1;
          }
          ++nlvl;
          __syncthreads();  // every thread has read sm.tail before the next level pushes
          for (int q = level0 + t; q < level1; q += GROW_THREADS) {
            const int d = sm.u.d.wl[q];
            atomicMax(&W.wmin[d], key);
            atomicMax(&W.amin[d], key);
            atomicMax(&W.dsc[d], key);
            for (int c = __ldcg(&T.head[d]), guard = 0; c >= 0 && guard < T.cap; c = __ldcg(&T.next[c]), ++guard) {
              const int pos = atomicAdd(&sm.tail, 1);
              if (pos < WAVE_QCAP) sm.u.d.wl[pos] = c;
            }
          }
          level0 = level1;
          __syncthreads();
        }
        nwl = overflow ? 0 : level0;
        // a near node below one of this sample's own rewired nodes would see a changed cost mid-expansion
        int intra = 0;
        if (!overflow && nwl > 0) {
          __threadfence();
          __syncthreads();
          intra = __syncthreads_or(my_near && mark_sample(__ldcg(&W.dsc[my_node]), round) == cta);
        }
        outcome = (overflow || intra) ? OUT_COMPLEX : OUT_INSERT;
      } while (0);
      if (active && t == 0) {
        WaveRec r;
        r.outcome = outcome; r.n_b = n_b; r.par = par; r.park = park; r.K = K; r.nrew = nrew; r.nwl = nwl; r.pad = 0;
        r.ctc_new = ctc_new; r.tim_new = tim_new; r.collide_calls = collide_calls; r.d2_in_max = 0.0; r.d2_out_min = 0.0;
        W.rec[b] = r;
      }
      grid_barrier(cfg.bar, bar_target);
      phase_done(2);

      // -------------------------------------------------------------- validate
      int n_true = n_b;
      if (active) {
        bool fail = outcome == OUT_COMPLEX || outcome == OUT_CONFLICT;
        if (outcome == OUT_NOFREE || outcome == OUT_INSERT) {
          // the node index / tree size this sample really gets: earlier samples that found no free parent did not insert.
          // The decision was taken at the radius of n_b; it stands if no candidate changes sides between the two radii.
          const int ins = __syncthreads_count(t < cta && __ldcg(&W.rec[F0 + t].outcome) == OUT_INSERT);
          n_true = nsub + ins;
          bool conf = false;
          if (n_true != n_b && my_node >= 0) {
            const double rt = P.rho2[n_true];
            conf = (my_d <= rho2n) != (my_d <= rt);
          }
          if (my_near) conf |= mark_sample(__ldcg(&W.wmin[my_node]), round) < cta;
          if (outcome == OUT_INSERT) {
            if (t < nrew) conf |= mark_sample(__ldcg(&W.amin[sm.u.d.rewn[t]]), round) < cta;
            for (int q = t; q < nwl; q += GROW_THREADS) conf |= mark_sample(__ldcg(&W.amin[sm.u.d.wl[q]]), round) < cta;
          }
          fail = __syncthreads_or(conf) != 0;
        }
        if (fail && t == 0) atomicMin(&ctl->wave_Fs[slot], cta);
      }
      grid_barrier(cfg.bar, bar_target);
      phase_done(3);

      // -------------------------------------------------------------- apply
      const int F = min(nb, __ldcg(&ctl->wave_Fs[slot]));
      if (active && cta < F && outcome == OUT_INSERT) {
        const int nw = n_true;
        if (t == 0) {
          T.pos[nw] = make_double2(s.x, s.y);
          T.vel[nw] = make_double2(s.vx, s.vy);
          T.parent[nw] = par;
          T.ecost[nw] = __ldcg(&Bt.cf[base + park]);
          T.etime[nw] = sm.u.d.tf[park];
          T.ctc[nw] = ctc_new;
          T.tim[nw] = tim_new;
          T.head[nw] = nrew > 0 ? sm.u.d.rewn[0] : -1;  // the rewired nodes become the new node's children
          list_lock(W.lock, par);
          link_child_cg(T, par, nw);
          list_unlock(W.lock, par);
          Bt.map[b] = nw;
        }
        if (nrew > 0) {
          // detach each rewired node from its old parent (one thread per warp takes the lock: no spinning inside a warp)
          if ((t & 31) == 0) {
            for (int q = t >> 5; q < nrew; q += GROW_THREADS / 32) {
              const int j = sm.u.d.rewn[q];
              const int old = __ldcg(&T.parent[j]);
              list_lock(W.lock, old);
              unlink_child_cg(T, old, j);
              list_unlock(W.lock, old);
            }
          }
          __syncthreads();
          if (my_q >= 0) {
            const int j = my_node;
            const double tb = sm.u.d.tb[t];
            T.parent[j] = nw;
            T.next[j] = my_q + 1 < nrew ? sm.u.d.rewn[my_q + 1] : -1;
            T.prev[j] = my_q > 0 ? sm.u.d.rewn[my_q - 1] : -1;
            T.ecost[j] = my_cb;
            T.etime[j] = tb;
            T.ctc[j] = ctc_new + my_cb;
            T.tim[j] = tim_new + tb;
          }
          __syncthreads();
          // level order: a node is updated one level after its parent
          for (int l = 0; l < nlvl; ++l) {
            const int q0 = sm.u.d.lvl[l], q1 = sm.u.d.lvl[l + 1];
            for (int q = q0 + t; q < q1; q += GROW_THREADS) {
              const int d = sm.u.d.wl[q];
              const int pd = __ldcg(&T.parent[d]);
              __stcg(&T.ctc[d], __ldcg(&T.ctc[pd]) + __ldcg(&T.ecost[d]));
              __stcg(&T.tim[d], __ldcg(&T.tim[pd]) + __ldcg(&T.etime[d]));
            }
            __syncthreads();
          }
        }
      }
      // -------------------------------------------------------------- finish (last CTA, one warp)
      if (last && t < 32) {
        const int lane = t;
        if (F > 0) {  // fold the counters of the prefix the other CTAs are committing
          long long v[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
          for (int k = lane; k < F; k += 32) {
            const int o = __ldcg(&W.rec[F0 + k].outcome);
            v[0] += 1;
            if (o == OUT_INSERT) {
              const int rk = __ldcg(&W.rec[F0 + k].K);
              v[1] += 1; v[5] += __ldcg(&W.rec[F0 + k].nrew); v[6] += __ldcg(&W.rec[F0 + k].nwl); v[7] += rk;
              v[8] += 2 * rk - 1; v[9] += __ldcg(&W.rec[F0 + k].collide_calls);
            } else if (o == OUT_REJECT) {
              v[2] += 1;
            } else if (o == OUT_DUP) {
              v[3] += 1;
            } else {  // OUT_NOFREE
              const int rk = __ldcg(&W.rec[F0 + k].K);
              v[4] += 1; v[7] += rk; v[8] += rk; v[9] += __ldcg(&W.rec[F0 + k].collide_calls);
            }
          }
#pragma unroll
          for (int k = 0; k < 10; ++k)
            for (int off = 16; off > 0; off >>= 1) v[k] += __shfl_xor_sync(FULL, v[k], off);
          if (lane == 0) {
            ctl->n = nsub + (int)v[1];
            ctl->iter += F;
            for (int k = 0; k < 10; ++k) ctl->stats[k] += v[k];
            const State4 ls = Bt.samples[F0 + F - 1];
            ctl->last_sample[0] = ls.x; ctl->last_sample[1] = ls.y; ctl->last_sample[2] = ls.vx; ctl->last_sample[3] = ls.vy;
            ctl->ema16 = (3 * ctl->ema16 + 16 * F) / 4;
            ctl->rounds_wave += 1;
            ctl->wave_committed += F;
            int why = 0;
            if (F < nb) {
              const int o = __ldcg(&W.rec[F0 + F].outcome);
              why = o == OUT_CONFLICT ? 1 : o == OUT_COMPLEX ? 2 : (o == OUT_NOFREE ? 3 : 4);
            }
            ctl->why[why] += 1;
          }
        } else {  // the first sample needs the sequential path (empty near set, huge subtree, ...): commit it alone
          seq_commit(ctl, P, T, Bt, dyn_smem, F0, F0 + 1);
          if (lane == 0) {
            ctl->ema16 = (3 * ctl->ema16 + 16) / 4;
            ctl->rounds_seq += 1;
          }
        }
        __syncwarp();
        if (F0 + max(F, 1) >= Wn) clear_window(Bt);
        if (lane == 0) {
          ctl->wave_Fs[slot ^ 1] = 0x7fffffff;
          ctl->round = round + 1;
        }
      }
      grid_barrier(cfg.bar, bar_target);
      phase_done(4);
      F0 += max(F, 1);
    }
  }
}

}  // namespace krrt
