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
//   apply     same CTA, for the prefix: insert, rewire, store the propagated costs/times decide worked out (lists kept
//             in shared memory since decide); the last CTA folds the counters and, if the sample right behind the prefix
//             needs the in-order path, commits it in the same sub-round
//
// Why the prefix is exact.  RRT* is sequential: expansion k+1 sees the tree after k.  But an expansion only READS
// the cost/time of its near set and only WRITES the cost/time of the nodes it rewires and of their descendants.
// Each decision leaves marks: on the nodes it writes (with where the NEW cost is on record), on the node it chose as
// parent.  Sample b stays in the prefix iff, for every earlier sample of the sub-round,
//   - a near node of b that the earlier sample writes keeps b's decision unchanged at its new (lower) cost: it is not
//     b's parent, was not tried, is no rewire candidate, and is not cheaper than the chosen parent now;
//   - nothing b writes was written, or chosen as parent, by the earlier sample (a node b merely read and a later sample
//     writes is a write-after-read: harmless, b's decision is already taken);
//   - the earlier sample's new node, if it lands inside b's radius, would not be a duplicate, and where it would be
//     tried as parent or is a rewire candidate its edge collides (tested at validation, counted);
//   - b's near set is unchanged at the radius of the tree size it really sees (earlier samples that found no
//     collision-free parent did not insert).
// For such a prefix the sequential outcome IS the speculated one, counters included.  A rewire candidate that is a CHILD
// of a node the same sample rewires is settled inside decide (ascending-index order, second test at the lowered cost).
// A sample that cannot be decided this way (empty near set -> global nearest, more than GROW_QCAP descendants or
// GROW_LEVELS levels below its rewired nodes, a candidate deeper below a rewired node) is committed in order by one warp;
// while the tree is small (conflicts everywhere) whole windows are.  Trees and counters are bit-identical to the
// in-order commit whatever the prefix lengths are (DESIGN.md 5.1 lists each rule with its measured effect).
#pragma once
#include "krrt_kernels.cuh"

namespace krrt {

// -DKRRT_CHECK (lib/libkrrt_check.so, `make check`; tests/test_switch_matrix.py runs the GPU suite's hardest plans on it):
// every capacity the parallel commit relies on is asserted where it is used; a violation prints its code and traps, so the
// host sees KRRT_ERR_CUDA with a message instead of a silently corrupted tree or an undiagnosed `illegal instruction`.
#ifdef KRRT_CHECK
#define KRRT_DCHECK(cond, code)                                                                                      \
  do {                                                                                                               \
    if (!(cond)) {                                                                                                   \
      printf("krrt check %d failed: %s (cta %d thread %d, %s:%d)\n", (int)(code), #cond, (int)blockIdx.x, (int)threadIdx.x, \
             __FILE__, __LINE__);                                                                                    \
      __trap();                                                                                                      \
    }                                                                                                                \
  } while (0)
#else
#define KRRT_DCHECK(cond, code) do { } while (0)
#endif

constexpr int GROW_THREADS = 256;
#ifndef KRRT_GROW_MINB
#define KRRT_GROW_MINB 1   // CTAs of k_grow per SM the register allocation is held to.  2 (128 registers, 264 B of spills, 295 samples per
                           // sub-round) was measured: the spills cost what the longer prefixes gain (profiles/r2_grow_two_ctas_s2.txt)
#endif
constexpr int GROW_WMAX = 1024;   // window cap
constexpr int GROW_LEVELS = 62;   // BFS depth below the rewired nodes that the parallel path handles
constexpr int GROW_SEQ_WINDOW = 64;
constexpr int GROW_SMAX = 64;     // edge sample counts up to this use a shared table of h = p / (S - 1)

__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_u32(unsigned* p, unsigned v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// watchdog of the flag waits below (see grid_barrier)
__device__ __forceinline__ void flag_watchdog(unsigned& spins, unsigned long long& t_first, const char* what) {
  if ((++spins & 0x3fffu) == 0) {
    unsigned long long now;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
    if (t_first == 0) t_first = now;
    else if (now - t_first > 5000000000ull) {
      printf("krrt k_grow: %s wedged (cta %d, thread %d)\n", what, (int)blockIdx.x, (int)threadIdx.x);
      __trap();
    }
  }
}

// Grid barrier over n co-resident CTAs (cooperative launch, one CTA per SM): one counter, red.release to arrive
// (MEMBAR.ALL.GPU + REDG), ld.acquire to poll (LDG.STRONG.GPU + CCTL.IVALL, so later plain loads miss L1).  Measured on
// B200 (tests/native/barrier_bench.cu): 2200 cycles with 64 CTAs, against 3000 for fence + atomicAdd + spin + fence
// and 2400 for cooperative_groups grid.sync().
// `relaxed`: CTAs that only wait for a far-away event (the end of the window's sub-rounds) back off between polls.
struct GridBar {
  unsigned long long arrive, pad0[15];
};
struct BarState {
  unsigned long long arrived;
};
struct GrowCfg {
  int wave_enabled;
  int wave_min_n;
  const double* user_samples;  // device copy of the caller's explicit samples, or nullptr
  GridBar* bar;                // [0] all CTAs, [1] the CTAs that take part in a window's sub-rounds (zero at launch)
  GridIndex gi;                // spatial index of the neighbour scan
  int raw_forward;             // a near node rewired by an earlier sample of the sub-round: re-check with its new cost
  int pend_resolve;            // resolve neighbours decided in the same sub-round at validation (0: they end the prefix)
  int join_test;               // a newcomer that would be tried as parent / is a rewire candidate: test its edge at validation (0: ends the prefix)
  int flow;                    // sub-rounds as a dataflow: a sample waits for the samples BEFORE it only (0: two grid barriers)
  int prop_forward;            // a near node whose cost an earlier sample's PROPAGATION changes: re-check with the new cost
  int intra_fix;               // a rewire candidate that is a CHILD of a node the same sample rewires is settled in decide (0: in-order path)
  int tail_seq;                // a sample that needs the in-order path right behind the prefix is committed in the same sub-round
  IterLog log;                 // krrt_grow_logged: per-iteration records of the in-order commit (n_after == nullptr: off)
  int nb_mul, nb_add, nb_min;  // samples decided per sub-round = max(nb_min, nb_mul x mean prefix + nb_add)
  int pn_mul, pn_add, pn_min;  // CTAs that take part in a window's sub-rounds, likewise
};

// a non-negative cost / squared distance as an unsigned key that orders the same way (arg-min by REDUX on its two words)
#ifdef KRRT_F32_BUILD
__device__ __forceinline__ unsigned long long real_key(double x) { return (unsigned long long)__float_as_uint(x) << 32; }
__device__ __forceinline__ double key_real(unsigned long long k) { return __uint_as_float((unsigned)(k >> 32)); }
#else
__device__ __forceinline__ unsigned long long real_key(double x) { return (unsigned long long)__double_as_longlong(x); }
__device__ __forceinline__ double key_real(unsigned long long k) { return __longlong_as_double((long long)k); }
#endif

__device__ __forceinline__ void grid_barrier(GridBar* bar, BarState& st, int n, bool relaxed) {
  __syncthreads();
  st.arrived += (unsigned long long)n;
  if (threadIdx.x == 0) {
    asm volatile("red.release.gpu.global.add.u64 [%0], 1;" ::"l"(&bar->arrive) : "memory");
    // watchdog: no barrier of a healthy plan waits longer than a window (< 1 ms); a wait of 5 s means the grid is wedged,
    // and a trap (the host sees cudaErrorLaunchFailure -> KRRT_ERR_CUDA) is better than a kernel that never returns
    unsigned spins = 0;
    unsigned long long t_first = 0;
    while (ld_acquire_u64(&bar->arrive) < st.arrived) {
      if (relaxed) __nanosleep(400);
      if ((++spins & 0x3fffu) == 0) {
        unsigned long long now;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
        if (t_first == 0) t_first = now;
        else if (now - t_first > 5000000000ull) {
          printf("krrt k_grow: grid barrier wedged (cta %d, expected %llu, seen %llu)\n", (int)blockIdx.x, st.arrived,
                 ld_acquire_u64(&bar->arrive));
          __trap();
        }
      }
    }
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

// (both flags are loaded before either is looked at: a short-circuit would chain two L2 round trips)
__device__ __forceinline__ bool sample_insertable(const Batch& Bt, const Params& P, int e) {
  const int rej = __ldcg(&Bt.reject[e]), dup = __ldcg(&Bt.dup[e]);
  return !((!P.dynamic & (rej != 0)) | (dup != 0));
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
  const int h = __ldcg(&T.link[par].x);
  __stcg(&T.link[c].y, h);
  __stcg(&T.prev[c], -1);
  if (h >= 0) __stcg(&T.prev[h], c);
  __stcg(&T.link[par].x, c);
}
__device__ __forceinline__ void unlink_child_cg(const TreeView& T, int par, int c) {
  const int nx = __ldcg(&T.link[c].y), pv = __ldcg(&T.prev[c]);
  if (pv >= 0) __stcg(&T.link[pv].y, nx); else __stcg(&T.link[par].x, nx);
  if (nx >= 0) __stcg(&T.prev[nx], pv);
}

constexpr int GROW_XE = 8;  // newcomer edges one sample can settle at validation
#ifndef KRRT_QCAP
#define KRRT_QCAP 1024
#endif
constexpr int GROW_QCAP = KRRT_QCAP;   // strict descendants of one sample's rewired nodes that the parallel path handles
constexpr int WL_DEAD = (int)0x80000000;  // wl_par of a write-list entry that is rewired in its own right
constexpr int GROW_HASH = 512;    // near-set membership table (>= 2 x GROW_THREADS)

struct GrowShared {  // lives in dynamic shared memory, in front of the in-order commit's scratch
  union {
    State4 ss[GROW_WMAX];  // scan: the window's samples
    struct {
      State4 st[GROW_THREADS];   // state of candidate slot k's node
      double tf[GROW_THREADS];   // tau*(node -> sample)
      double tb[GROW_THREADS];   // tau*(sample -> node)
      double tim[GROW_THREADS];  // arrival time of the node
      double cb[GROW_THREADS];   // c*(sample -> node)
      int flag[GROW_THREADS];    // 1 in near set, 2 forward edge collides, 8 rewire edge collides; 16 child of a rewired node,
                                 // 32 rewired with a candidate among its children, 64 rewired twice over (see decide)
      int cand[GROW_THREADS];    // slots of the rewire candidates
      int rewn[GROW_THREADS];    // rewired nodes, in the order they become children of the new node
      int rslot[GROW_THREADS];   // their candidate slots
      int rpar[GROW_THREADS];    // their parents so far
      double rctc[GROW_THREADS]; // their new cost-to-come / arrival time (decide)
      double rtim[GROW_THREADS];
      int hash[GROW_HASH];       // open addressing over the rewire candidates' node indices ...
      int hslot[GROW_HASH];      // ... and their candidate slots
      int wl[GROW_QCAP];         // strict descendants of the rewired nodes, level by level ...
      int wl_head[GROW_QCAP];    // ... their first child (read together with the sibling link) ...
      int wl_par[GROW_QCAP];     // ... and their parent: index in wl, or -(q + 1) for rewired node q
      double wl_ctc[GROW_QCAP];  // new cost-to-come / arrival time (decide, on the way down)
      double wl_tim[GROW_QCAP];
      int lvl[GROW_LEVELS + 2];  // level boundaries in wl
    } d;
  } u;
  double4 obs[KRRT_MAX_OBS_DEV];
  double red_c[GROW_THREADS / 32];
  int red_i[GROW_THREADS / 32], red_k[GROW_THREADS / 32];
  double hS[GROW_SMAX];      // h = p / (S - 1) of the edge sample points
  double rho[GROW_THREADS];  // squared near radius at tree sizes nsub .. nsub + nb
  int woff[GROW_WMAX + 1];   // steer: exclusive scan of the window's candidate counts
  double par_cf;             // c*(parent -> sample)
  int ncand, nrew, tail, tries, K, flags, intra;
  // validate: edges to / from newcomers (earlier samples of the sub-round) that have to be collision-tested
  State4 xe_a[GROW_XE], xe_b[GROW_XE];
  double xe_tau[GROW_XE], xe_t0[GROW_XE];
  int xe_hit[GROW_XE], nxe, xcalls;
  int tmpF;                  // flow: first sample of the sub-round that has not passed validation (<< 1 | not known yet)
  int ndead, xprop, tries_adj, xflag;  // candidates among the children of rewired nodes (decide)
  long long fin[20];  // last CTA: counters of the call, flushed to the control block when the kernel ends
};
__host__ __device__ inline size_t grow_shared_bytes() { return (sizeof(GrowShared) + 15) / 16 * 16; }

__device__ __forceinline__ unsigned near_hash(int node) { return ((unsigned)node * 2654435761u) >> 23; }  // 9 bits

// children of `first`'s list go to the write list: one 8-byte load per sibling gives its own first child too
// (the child's edge cost / duration ride along: its new cost-to-come / arrival time = the parent's new ones + the edge's)
__device__ __forceinline__ void push_children(const TreeView& T, GrowShared& sm, int first, int parent_ref, double pc,
                                              double pt) {
  int c = first;
  for (int guard = 0; c >= 0 && guard < T.cap; ++guard) {
    const int2 lk = __ldcg(&T.link[c]);
    const double ec = __ldcg(&T.ecost[c]), et = __ldcg(&T.etime[c]);
    const int pos = atomicAdd(&sm.tail, 1);
    KRRT_DCHECK(c >= 0 && c < T.cap, 101);  // a child link that leaves the tree
    if (pos < GROW_QCAP) {
      sm.u.d.wl[pos] = c;
      sm.u.d.wl_head[pos] = lk.x;
      sm.u.d.wl_par[pos] = parent_ref;
      sm.u.d.wl_ctc[pos] = pc + ec;
      sm.u.d.wl_tim[pos] = pt + et;
    }
    c = lk.y;
  }
}

// ---------------------------------------------------------------------------------------------------
// In-order commit of window samples [b0, b1) by ONE warp (the whole window while the tree is small; a single sample
// when it needs the sequential path).  smem: obstacles + near list (commit_smem_bytes).
// nn_idx >= 0: the nearest node of sample b0 over the whole tree (found by a grid-wide scan), used if its near set is empty.
// kLog: the variant behind krrt_grow_logged, which also leaves a record per iteration (krrt_kernels.cuh: IterLog); the
// planner proper is the kLog = false instantiation and carries none of it.
template <bool kLog>
__device__ __noinline__ void seq_commit(Ctl* ctl, const Params& P, const TreeView& T, const Batch& Bt,
                                        unsigned char* smem, int b0, int b1, int nn_idx, double nn_d2, const IterLog& Lg) {
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
  // logged call: what iteration `iter` left behind (the tree's costs are copied only when the tree changed)
  const long long iter_begin = kLog ? __ldcg(&ctl->iter_begin) : 0;
  auto log_iter = [&](bool changed) {
    if (!kLog) return;
    const long long i = iter - iter_begin - 1;
    if (i < 0 || i >= Lg.rows - 1) return;
    __syncwarp();
    const int row = changed ? (int)i + 1 : (i > 0 ? Lg.snap_of[i - 1] : 0);
    if (changed) {
      for (int j = lane; j < n && j < Lg.stride; j += 32) {
        Lg.snap_ctc[(size_t)row * Lg.stride + j] = __ldcg(&T.ctc[j]);
        Lg.snap_tim[(size_t)row * Lg.stride + j] = __ldcg(&T.tim[j]);
      }
    }
    if (lane == 0) {
      Lg.n_after[i] = n;
      Lg.snap_of[i] = row;
      Lg.sample[4 * i + 0] = s.x; Lg.sample[4 * i + 1] = s.y; Lg.sample[4 * i + 2] = s.vx; Lg.sample[4 * i + 3] = s.vy;
    }
    __syncwarp();
  };
  for (int b = b0; b < b1; ++b) {
    s = Bt.samples[b];
    ++iter;
    st.iters++;
    if (!P.dynamic && __ldcg(&Bt.reject[b])) {
      st.fail_reject++;
      log_iter(false);
      continue;
    }
    if (n >= P.capacity) {
      full = 1;
      log_iter(false);
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
        dup |= valid && d == KR(0.0);
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
      log_iter(false);
      continue;
    }
    if (K == 0) {  // empty near set: exact nearest over the whole tree (rare once the tree is dense)
      int ni = nn_idx;
      double nd2 = nn_d2;
      if (ni < 0 || b != b0) warp_nearest(T, n, s, ni, nd2);
      if (nd2 == KR(0.0)) {
        st.fail_dup++;
        log_iter(false);
        continue;
      }
      if (lane == 0) {
        double t, c;
        steer(load_state(T, ni), s, P.r, t, c);
        N.idx[0] = ni;
        N.cf[0] = c;
        N.tf[0] = t;
        N.cb[0] = KR(0.0);
        N.tb[0] = KR(0.0);
      }
      K = 1;
    }
    __syncwarp();
    KRRT_DCHECK(K >= 1 && K <= cap && n >= 1 && n < P.capacity, 111);
    st.sum_near += K;
    int nw = warp_commit(T, N, K, n, s, E, st);
    if (nw >= 0) {
      if (lane == 0) Bt.map[b] = nw;
      n = nw + 1;
    }
    __syncwarp();
    log_iter(nw >= 0);
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
template <bool kLog>
__device__ __forceinline__ void grow_body(Ctl* ctl, const Params& P, const TreeView& T, const Batch& Bt, const Wave& W,
                                          const GrowCfg& cfg) {
  extern __shared__ __align__(16) unsigned char dyn_raw[];
  GrowShared& sm = *reinterpret_cast<GrowShared*>(dyn_raw);
  unsigned char* dyn_smem = dyn_raw + grow_shared_bytes();  // in-order commit scratch (last CTA only)
  const int t = threadIdx.x;
  const int G = gridDim.x;
  const int cta = blockIdx.x;
  const bool last = cta == G - 1;  // never owns a sample: finishes the sub-rounds, commits sequentially
  const int cap = Bt.near_cap;
  const int S = P.S;
  BarState bs_all = {0ull}, bs_sub = {0ull};
  if (t < 20) sm.fin[t] = 0;
  if (t < GROW_SMAX) sm.hS[t] = (S > 1) ? (double)t / (double)(S - 1) : KR(0.0);
  long long clk_mark = clock64();
  auto phase_done = [&](int which) {  // phase clocks of the last CTA (it waits at every barrier like the others)
    if (last && t == 0) {
      const long long now = clock64();
      ctl->phase_clk[which] += now - clk_mark;
      clk_mark = now;
    }
  };

#ifdef KRRT_DEBUG_CLK
  long long dbg_t0 = 0;
  long long dbg_acc[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  long long cta_dec = 0, cta_val = 0, cta_app = 0, cta_t = 0;
#define DBG_START() do { if (t == 0) dbg_t0 = clock64(); } while (0)
#define DBG_STAMP(k) do { if (cta == 0 && t == 0) { const long long now_ = clock64(); dbg_acc[k] += now_ - dbg_t0; dbg_t0 = now_; } } while (0)
#else
#define DBG_START()
#define DBG_STAMP(k)
#endif
  for (;;) {
    // ================================================================ window
    const int n0 = __ldcg(&ctl->n);
    const long long iter0 = __ldcg(&ctl->iter);
    const long long left = __ldcg(&ctl->iter_target) - iter0;
    const int full0 = __ldcg(&ctl->full), error0 = __ldcg(&ctl->error);
    const unsigned round0 = __ldcg(&ctl->round);
    if ((full0 | error0) != 0 || left <= 0 || round0 >= (1u << 23) - 2) break;
    const bool wave = cfg.wave_enabled && n0 >= cfg.wave_min_n;
    const int Wn = (int)min((long long)(wave ? Bt.B : min(Bt.B, GROW_SEQ_WINDOW)), left);
    const int user_off = (int)(iter0 - __ldcg(&ctl->iter_begin));

    // squared radius that covers every tree size this window can see
    double rmax = KR(0.0);
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
    // ---- spatial index: (re)built when the part of the tree it does not cover has outgrown 1/16 of the tree, or when its
    // cells are too small for this window's radius (krrt_set_params raised it).  Every CTA takes the same decision from the
    // same values (read after the previous window's end barrier; written below by one thread, a barrier before any use).
    int gn = 0, gw = 1, gh = 1;
    double g_lox = KR(0.0), g_loy = KR(0.0), g_invx = KR(0.0), g_invy = KR(0.0);
    if (cfg.gi.enabled && n0 >= cfg.gi.min_n) {
      gn = __ldcg(&ctl->grid_n);
      const double gcs = __ldcg(&ctl->grid_cs);
      const bool stale = gn == 0 || n0 - gn > max(2048, min(n0 >> 4, cfg.gi.tail_max)) || !(gcs * gcs >= KR(1.001) * rmax);
      if (stale) {
        // geometry: cells at least 1.001 x the near radius wide (the margin dwarfs the rounding of the cell arithmetic,
        // so two points closer than the radius on an axis are never more than one cell apart on it)
        const double rad = sqrt(rmax);
        const double cs = KR(1.001) * rad;
        g_lox = P.sp.lo[0];
        g_loy = P.sp.lo[1];
        const double wx = P.sp.hi[0] - P.sp.lo[0], wy = P.sp.hi[1] - P.sp.lo[1];
        gw = (cs > KR(0.0) && wx > KR(0.0)) ? (int)fmin((double)GRID_W_MAX, fmax(KR(1.0), floor(wx / cs))) : 1;
        gh = (cs > KR(0.0) && wy > KR(0.0)) ? (int)fmin((double)GRID_H_MAX, fmax(KR(1.0), floor(wy / cs))) : 1;
        g_invx = wx > KR(0.0) ? (double)gw / wx : KR(0.0);
        g_invy = wy > KR(0.0) ? (double)gh / wy : KR(0.0);
        const int ncell = gw * gh;
        for (int c = cta * GROW_THREADS + t; c <= ncell; c += G * GROW_THREADS) cfg.gi.cell[c] = 0;
        grid_barrier(cfg.bar, bs_all, G, false);
        for (int i = cta * GROW_THREADS + t; i < n0; i += G * GROW_THREADS) {
          const double2 p = __ldcg(&T.pos[i]);
          const int cx = min(gw - 1, max(0, (int)floor((p.x - g_lox) * g_invx)));
          const int cy = min(gh - 1, max(0, (int)floor((p.y - g_loy) * g_invy)));
          cfg.gi.off[i] = atomicAdd(&cfg.gi.cell[cy * gw + cx], 1);
        }
        grid_barrier(cfg.bar, bs_all, G, false);
        if (cta == 0) {  // exclusive prefix of the counts, in place: a run of cells per thread
          int* part = reinterpret_cast<int*>(dyn_smem);
          const int per = (ncell + GROW_THREADS - 1) / GROW_THREADS;
          const int c0 = min(ncell, t * per), c1 = min(ncell, c0 + per);
          int sum = 0;
          for (int c = c0; c < c1; ++c) sum += __ldcg(&cfg.gi.cell[c]);
          part[t] = sum;
          __syncthreads();
          if (t == 0) {
            int acc = 0;
            for (int k = 0; k < GROW_THREADS; ++k) {
              const int v = part[k];
              part[k] = acc;
              acc += v;
            }
            cfg.gi.cell[ncell] = acc;
          }
          __syncthreads();
          int acc = part[t];
          for (int c = c0; c < c1; ++c) {
            const int v = __ldcg(&cfg.gi.cell[c]);
            cfg.gi.cell[c] = acc;
            acc += v;
          }
        }
        grid_barrier(cfg.bar, bs_all, G, false);
        for (int i = cta * GROW_THREADS + t; i < n0; i += G * GROW_THREADS) {
          const double2 p = __ldcg(&T.pos[i]);
          const int cx = min(gw - 1, max(0, (int)floor((p.x - g_lox) * g_invx)));
          const int cy = min(gh - 1, max(0, (int)floor((p.y - g_loy) * g_invy)));
          const int slot = __ldcg(&cfg.gi.cell[cy * gw + cx]) + cfg.gi.off[i];
          KRRT_DCHECK(slot >= 0 && slot < n0 && gw * gh <= GRID_CELLS_MAX, 112);
          cfg.gi.pos[slot] = p;
          cfg.gi.vel[slot] = __ldcg(&T.vel[i]);
          cfg.gi.id[slot] = i;
        }
        if (last && t == 0) {
          ctl->grid_n = n0;
          ctl->grid_w = gw;
          ctl->grid_h = gh;
          ctl->grid_cs = cs;
          ctl->grid_lox = g_lox; ctl->grid_loy = g_loy; ctl->grid_invx = g_invx; ctl->grid_invy = g_invy;
          ctl->grid_rebuilds += 1;
        }
        gn = n0;
        grid_barrier(cfg.bar, bs_all, G, false);
      } else {
        gw = __ldcg(&ctl->grid_w);
        gh = __ldcg(&ctl->grid_h);
        g_lox = __ldcg(&ctl->grid_lox); g_loy = __ldcg(&ctl->grid_loy);
        g_invx = __ldcg(&ctl->grid_invx); g_invy = __ldcg(&ctl->grid_invy);
      }
    }
    // ---- scan, indexed part: item = (sample, cell row of its 3 x 3 block) = one contiguous range of the sorted copy
    if (gn > 0) {
      for (int item = cta; item < 3 * Wn; item += G) {
        const int b = item / 3, dy = item - 3 * b - 1;
        const State4 sb = sm.u.ss[b];
        const int cx = min(gw - 1, max(0, (int)floor((sb.x - g_lox) * g_invx)));
        const int cy = min(gh - 1, max(0, (int)floor((sb.y - g_loy) * g_invy)));
        const int row = cy + dy;
        if (row < 0 || row >= gh) continue;
        const int begin = __ldcg(&cfg.gi.cell[row * gw + max(cx - 1, 0)]);
        const int end = __ldcg(&cfg.gi.cell[row * gw + min(cx + 1, gw - 1) + 1]);
        // four slots per thread and trip, all loads issued before the first use: the range is read at L2 latency
        for (int slot0 = begin + t; slot0 < end; slot0 += 4 * GROW_THREADS) {
          double2 p[4], v[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int slot = slot0 + u * GROW_THREADS;
            if (slot < end) {
              p[u] = __ldcg(&cfg.gi.pos[slot]);
              v[u] = __ldcg(&cfg.gi.vel[slot]);
            }
          }
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int slot = slot0 + u * GROW_THREADS;
            if (slot < end) {
              const State4 a = {p[u].x, p[u].y, v[u].x, v[u].y};
              const double d = dist2(a, sb);
              if (d <= rmax) {
                if (d == KR(0.0)) Bt.dup[b] = 1;
                append_candidate(Bt, b, __ldcg(&cfg.gi.id[slot]), d);
              }
            }
          }
        }
        if (t == 0 && end > begin) atomicAdd(&ctl->scan_visited, (double)(end - begin));
      }
    }
    if (last && t == 0) atomicAdd(&ctl->scan_visited, (double)(n0 - gn) * (double)Wn);
    // ---- scan, direct part (the whole tree while there is no index): item = (tile of GROW_THREADS frozen nodes,
    // chunk of the window's samples)
    {
      const int tiles = (n0 - gn + GROW_THREADS - 1) / GROW_THREADS;
      const int chunks = max(1, min(8, (2 * G + tiles - 1) / max(tiles, 1)));
      const int per = (Wn + chunks - 1) / chunks;
      for (int item = cta; item < tiles * chunks; item += G) {
        const int tile = item / chunks, ch = item - tile * chunks;
        const int i = gn + tile * GROW_THREADS + t;
        const int c0 = ch * per, c1 = min(Wn, c0 + per);
        if (i < n0) {
          const State4 a = load_state_cg(T, i);
#pragma unroll 4
          for (int b = c0; b < c1; ++b) {
            const double d = dist2(a, sm.u.ss[b]);
            if (d <= rmax) {
              if (d == KR(0.0)) Bt.dup[b] = 1;
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
    grid_barrier(cfg.bar, bs_all, G, false);
    phase_done(0);
    // ---- steer: tau*, c* of candidate -> sample and sample -> candidate, one thread per (candidate, direction).  The
    // window's (sample, candidate, direction) items are spread evenly over the whole grid: every CTA scans the candidate
    // counts for itself (one L2 round trip) and finds its items' samples by bisection.
    {
      constexpr int CPT = GROW_WMAX / GROW_THREADS;  // consecutive counts per thread
      static_assert(CPT * GROW_THREADS == GROW_WMAX, "window cap is a multiple of the CTA size");
      int c[CPT], mine = 0;
#pragma unroll
      for (int k = 0; k < CPT; ++k) {
        const int b = CPT * t + k;
        c[k] = b < Wn ? min(__ldcg(&Bt.count[b]), cap) : 0;
        mine += c[k];
      }
      int incl = mine;
#pragma unroll
      for (int off = 1; off < 32; off <<= 1) {
        const int v = __shfl_up_sync(FULL, incl, off);
        if ((t & 31) >= off) incl += v;
      }
      if ((t & 31) == 31) sm.red_i[t >> 5] = incl;
      __syncthreads();
      int run = incl - mine;
#pragma unroll
      for (int w = 0; w < GROW_THREADS / 32; ++w)
        if (w < (t >> 5)) run += sm.red_i[w];
#pragma unroll
      for (int k = 0; k < CPT; ++k) {
        const int b = CPT * t + k;
        if (b <= Wn) sm.woff[b] = run;
        run += c[k];
      }
      if (t == GROW_THREADS - 1 && Wn == GROW_WMAX) sm.woff[GROW_WMAX] = run;
      __syncthreads();
      const int items = 2 * sm.woff[Wn];
      for (int g = cta * GROW_THREADS + t; g < items; g += G * GROW_THREADS) {
        const int pair = g >> 1;
        const bool fwd = (g & 1) == 0;
        int lo = 0, hi = Wn;  // the last sample whose offset is <= pair (samples without candidates share an offset)
        while (hi - lo > 1) {
          const int mid = (lo + hi) >> 1;
          if (sm.woff[mid] <= pair) lo = mid; else hi = mid;
        }
        const int b = lo, e = pair - sm.woff[lo];
        const size_t base = (size_t)b * cap;
        const State4 s = sm.u.ss[b];
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
    // What the sub-rounds of this window start from is read BEFORE the barrier: after it, a CTA that has nothing to decide
    // arrives at the sub-round barrier within one L2 round trip, and the last CTA updates ema16 a few microseconds later —
    // a CTA that read either value late would wait for one arrival too many (or disagree on who takes part) and hang the
    // grid.  Here nothing is in flight: every sub-round barrier of the previous window completed before its end barrier.
    const unsigned long long sub_base = __ldcg(&cfg.bar[1].arrive);
    const int ema_win = __ldcg(&ctl->ema16);
    grid_barrier(cfg.bar, bs_all, G, false);
    phase_done(1);

    if (!wave) {  // small tree: the whole window in order, one warp
      if (last && t < 32) {
        seq_commit<kLog>(ctl, P, T, Bt, dyn_smem, 0, Wn, -1, KR(0.0), cfg.log);
        __syncwarp();
        clear_window(Bt);
        if (t == 0) ctl->rounds_seq += 1;
      }
      grid_barrier(cfg.bar, bs_all, G, !last);
      phase_done(5);
      continue;
    }

    // ================================================================ sub-rounds
    for (int j = t; j < P.m; j += GROW_THREADS) sm.obs[j] = P.obs[j];
    // the CTAs that take part in this window's sub-rounds: enough for twice the prefixes seen lately + the last CTA;
    // the others go straight to the window's end and doze there
    // (at most GROW_THREADS - 1 of them: validation looks at the earlier samples of the sub-round one thread each)
    const int Pn = min(min(G - 1, GROW_THREADS - 1), max(cfg.pn_min, (cfg.pn_mul * ema_win) / 16 + cfg.pn_add));
    const bool takes_part = cta < Pn || last;
    // the set changes from window to window: pick the barrier's counters up where they stood before the window's barrier
    bs_sub.arrived = sub_base;
    int F0 = 0;
    while (takes_part && F0 < Wn) {
      const int nsub = __ldcg(&ctl->n);
      const long long iter_sub = __ldcg(&ctl->iter);
      const int ema_sub = __ldcg(&ctl->ema16);
      const unsigned round = __ldcg(&ctl->round);
      const int full_sub = __ldcg(&ctl->full), error_sub = __ldcg(&ctl->error);  // (no short-circuit: one round trip)
      if ((full_sub | error_sub) != 0 || round >= (1u << 23) - 2) break;
      const int slot = (int)(round & 1u);
      const int nb = min(min(Pn, Wn - F0), max(cfg.nb_min, (cfg.nb_mul * ema_sub) / 16 + cfg.nb_add));
      const bool active = cta < nb;
      const int b = F0 + cta;
      const size_t base = (size_t)b * cap;
      const unsigned key = mark_key(round, cta);
      // what this CTA carries from decide to validate and apply
      int outcome = OUT_REJECT, n_b = 0, K = 0, par = -1, park = -1, nrew = 0, nwl = 0, nlvl = 0, nprop = 0;
      long long collide_calls = 0;
      double ctc_new = KR(0.0), tim_new = KR(0.0), rho2n = KR(0.0);
      int my_node = -1, my_q = -1;
      bool my_tried = false, my_cand = false;  // this thread's near node was tried as parent / is a rewire candidate
      int pend_e = -1;  // this thread's candidate is an earlier sample of the SAME sub-round (window index): see validate
      double pend_tf = KR(0.0), pend_tb = KR(0.0);  // tau* of that candidate -> sample / sample -> candidate
      double my_d = KR(0.0), my_cf = KR(0.0), my_cb = KR(0.0), my_ctc = KR(0.0);
      bool my_near = false;
      State4 s = {0, 0, 0, 0};

      // -------------------------------------------------------------- decide
      DBG_START();
#ifdef KRRT_DEBUG_CLK
      const long long dec_begin = clock64();
      cta_t = dec_begin;
#endif
      if (active) do {
        // first round trip: everything that is addressed by the sample alone.  Slots past the candidate count hold stale
        // but allocated data; they are ignored below.
        const bool spec = t < min(cap, 64);  // these slots are warm in L2 (every window uses them); the rest wait for the count
        int ref = 0;
        double my_tf = KR(0.0), my_tb = KR(0.0);
        if (spec) {
          ref = __ldcg(&Bt.ref[base + t]);
          my_d = __ldcg(&Bt.d2[base + t]);
          my_cf = __ldcg(&Bt.cf[base + t]);
          my_cb = __ldcg(&Bt.cb[base + t]);
          my_tf = __ldcg(&Bt.tf[base + t]);
          my_tb = __ldcg(&Bt.tb[base + t]);
        }
        const int cnt_raw = __ldcg(&Bt.count[b]);
        const int rej_b = __ldcg(&Bt.reject[b]);
        bool earlier_ins = false;
        if (t < cta) earlier_ins = sample_insertable(Bt, P, F0 + t);
        const bool rejected = !P.dynamic && rej_b;
        KRRT_DCHECK(nb >= 1 && nb < GROW_THREADS && cta < 512, 106);  // one thread per earlier sample; 9-bit sample id of the marks
        if (t <= nb) sm.rho[t] = P.rho2[min(nsub + t, P.capacity)];
        s = Bt.samples[b];
        if (t == 0) {
          sm.ncand = 0;
          sm.nrew = 0;
          sm.tail = 0;
          sm.tries = 0;
          sm.K = 0;
          sm.flags = 0;
          sm.intra = 0;
          sm.ndead = 0;
          sm.xprop = 0;
          sm.tries_adj = 0;
          sm.xflag = 0;
          sm.nxe = 0;
          sm.xcalls = 0;
        }
        if (rejected) break;  // OUT_REJECT
        DBG_STAMP(1);
        outcome = OUT_COMPLEX;
        // second round trip: the candidate's node (a sample committed in an earlier sub-round goes through the map)
        bool pend = false;
        const bool listed = t < cnt_raw && t < cap;
        if (listed && !spec) {
          ref = __ldcg(&Bt.ref[base + t]);
          my_d = __ldcg(&Bt.d2[base + t]);
          my_cf = __ldcg(&Bt.cf[base + t]);
          my_cb = __ldcg(&Bt.cb[base + t]);
          my_tf = __ldcg(&Bt.tf[base + t]);
          my_tb = __ldcg(&Bt.tb[base + t]);
        }
        if (listed) {
          if (ref & IN_BATCH) {
            const int e = ref & ~IN_BATCH;
            const int mapped = __ldcg(&Bt.map[e]);            // (all three loads leave together)
            const bool e_ins = sample_insertable(Bt, P, e);
            if (e < F0) my_node = mapped;  // committed in an earlier sub-round (-1: it did not insert)
            else if (cfg.pend_resolve) {  // decided in this sub-round: settled at validation
              pend_e = e_ins ? e : -1;
              pend_tf = my_tf;
              pend_tb = my_tb;
            }
            else pend = e_ins;             // may land inside the radius: wait for the next sub-round
          } else {
            my_node = ref;
          }
        }
        State4 x = {0, 0, 0, 0};
        double my_tim = KR(0.0);
        int my_head = -1, my_par0 = -1;
        KRRT_DCHECK(my_node < T.cap, 107);  // a candidate reference that leaves the tree
        if (my_node >= 0) {
          x = load_state_cg(T, my_node);
          my_ctc = __ldcg(&T.ctc[my_node]);
          my_tim = __ldcg(&T.tim[my_node]);
          my_head = __ldcg(&T.link[my_node].x);  // only needed if the node gets rewired; it rides along for free
          my_par0 = __ldcg(&T.parent[my_node]);  // (likewise: apply detaches the node from this parent's list)
        }
        sm.u.d.hash[t] = -1;
        sm.u.d.hash[t + GROW_THREADS] = -1;
        DBG_STAMP(13);
        // tree size this sample sees if every earlier sample of the sub-round that can insert does insert
        const int ins = __syncthreads_count(earlier_ins);
        n_b = nsub + ins;
        if (n_b >= P.capacity || cnt_raw > min(cap, GROW_THREADS)) break;  // tree full / overflow / huge near set
        rho2n = sm.rho[ins];
        DBG_STAMP(0);
        my_near = my_node >= 0 && my_d <= rho2n;
        {
          const unsigned nearb = __ballot_sync(FULL, my_near);
          const unsigned dupb = __ballot_sync(FULL, my_near && my_d == KR(0.0));
          const unsigned pendb = __ballot_sync(FULL, pend);
          if ((t & 31) == 0) {
            if (nearb) atomicAdd(&sm.K, __popc(nearb));
            if (dupb | pendb) atomicOr(&sm.flags, (dupb ? 1 : 0) | (pendb ? 2 : 0));
          }
        }
        if (my_near) {
          sm.u.d.st[t] = x;
          sm.u.d.tim[t] = my_tim;
          sm.u.d.tf[t] = my_tf;
          sm.u.d.tb[t] = my_tb;
          sm.u.d.cb[t] = my_cb;
        }
        sm.u.d.flag[t] = my_near ? 1 : 0;
        __syncthreads();
        K = sm.K;
        if (sm.flags & 1) {
          outcome = OUT_DUP;
          break;
        }
        if (sm.flags & 2) {
          outcome = OUT_CONFLICT;
          break;
        }
        if (K == 0) {  // nearest-node fallback: a grid-wide scan, then the in-order path
          outcome = OUT_NEAREST;
          break;
        }
        DBG_STAMP(2);
        // forward edges of ALL near nodes at once: one thread per (edge, sample point)
        if (P.m > 0) {
          const int tot = min(cnt_raw, GROW_THREADS) * S;
          for (int w = t; w < tot; w += GROW_THREADS) {
            const int k = w / S, p = w - k * S;
            if (!(sm.u.d.flag[k] & 1)) continue;
            const Edge ed = make_edge(sm.u.d.st[k], s, sm.u.d.tf[k], sm.u.d.tim[k]);
            const double h = S <= GROW_SMAX ? sm.hS[p] : ((S > 1) ? (double)p / (double)(S - 1) : KR(0.0));
            if (edge_point_hits_any(ed, h, sm.obs, P.m, P.dynamic, P.R2)) atomicOr(&sm.u.d.flag[k], 2);
          }
          __syncthreads();
        }
        DBG_STAMP(3);
        // parent = arg-min (cost, node index) over collision-free near nodes.  Costs are >= 0, so their bit patterns
        // order like unsigned integers: three REDUX per warp, then every thread folds the per-warp winners
        {
          const bool ok = my_near && !(sm.u.d.flag[t] & 2);
          const unsigned long long kb = ok ? real_key(my_ctc + my_cf) : ~0ull;
          const unsigned hi = (unsigned)(kb >> 32), lo = (unsigned)kb;
          const unsigned mh = __reduce_min_sync(FULL, hi);
          const unsigned ml = __reduce_min_sync(FULL, hi == mh ? lo : 0xffffffffu);
          const bool top = ok && hi == mh && lo == ml;
          const unsigned mn = __reduce_min_sync(FULL, top ? (unsigned)my_node : 0xffffffffu);
          if (top && (unsigned)my_node == mn) {
            sm.red_c[t >> 5] = my_ctc + my_cf;
            sm.red_i[t >> 5] = my_node;
            sm.red_k[t >> 5] = t;
          } else if ((t & 31) == 0 && mn == 0xffffffffu) {
            sm.red_i[t >> 5] = -1;
          }
          __syncthreads();
          double bc = INFINITY;
          int bi = -1, bk = -1;
#pragma unroll
          for (int w = 0; w < GROW_THREADS / 32; ++w) {
            const int wi = sm.red_i[w];
            if (wi < 0) continue;
            const double wc = sm.red_c[w];
            if (bi < 0 || wc < bc || (wc == bc && wi < bi)) {
              bc = wc;
              bi = wi;
              bk = sm.red_k[w];
            }
          }
          park = bk;
          par = bi;
          ctc_new = bc;
        }
        DBG_STAMP(4);
        if (park < 0) {
          outcome = OUT_NOFREE;
          collide_calls = K;  // a sequential planner tries every candidate before giving up
          break;
        }
        tim_new = sm.u.d.tim[park] + sm.u.d.tf[park];
        if (t == park) {
          sm.par_cf = my_cf;  // (read by thread 0 in apply, barriers in between)
          // parent mark: this sample's cost is DERIVED from the parent's.  A later sample of the sub-round that changes the
          // parent's cost (rewires it, or an ancestor of it) would have to propagate into this sample's new node, which
          // its write list cannot contain.  Plain reads of the other near nodes leave no such trace: a later write to
          // them is a write-after-read, harmless because this decision is already taken (sequential order).
          atomicMax(&W.amin[par], key);
        }
        // a sequential planner checks candidates in (cost, index) order until the first free one; rewire candidates:
        // near nodes that would get cheaper through the new node
        bool tried = false, cand = false;
        if (my_near) {
          const double c = my_ctc + my_cf;
          tried = c < ctc_new || (c == ctc_new && my_node <= par);  // all of these but the parent collide
          cand = my_node != par && (ctc_new + my_cb < my_ctc);
        }
        {
          const unsigned triedb = __ballot_sync(FULL, tried);
          if ((t & 31) == 0 && triedb) atomicAdd(&sm.tries, __popc(triedb));
        }
        my_tried = tried;
        my_cand = cand;
        if (cand) {
          KRRT_DCHECK(sm.ncand < GROW_THREADS && 2 * GROW_THREADS <= GROW_HASH, 102);  // candidate list / hash load <= 1/2
          sm.u.d.cand[atomicAdd(&sm.ncand, 1)] = t;
          // membership table of the REWIRE CANDIDATES (cleared before the barrier above): a candidate below another rewired
          // node would be tested against a cost that the first rewire has just lowered (rewires go in ascending index,
          // each sees the costs the previous one left) — that sample takes the in-order path.  A near node that is no
          // candidate at its old cost is none at a lower one, and the parent is chosen before any rewire: no need to
          // look for those below the rewired nodes.
          unsigned hsl = near_hash(my_node);
          while (atomicCAS(&sm.u.d.hash[hsl], -1, my_node) != -1) hsl = (hsl + 1) & (GROW_HASH - 1);
          sm.u.d.hslot[hsl] = t;
        }
        __syncthreads();
        const int ncand = sm.ncand;
        DBG_STAMP(5);
        collide_calls = sm.tries + ncand;
        // edges new -> candidate, started at the new node's arrival time
        if (P.m > 0 && ncand > 0) {
          for (int w = t; w < ncand * S; w += GROW_THREADS) {
            const int q = w / S, p = w - q * S;
            const int k = sm.u.d.cand[q];
            const Edge ed = make_edge(s, sm.u.d.st[k], sm.u.d.tb[k], tim_new);
            const double h = S <= GROW_SMAX ? sm.hS[p] : ((S > 1) ? (double)p / (double)(S - 1) : KR(0.0));
            if (edge_point_hits_any(ed, h, sm.obs, P.m, P.dynamic, P.R2)) atomicOr(&sm.u.d.flag[k], 8);
          }
          __syncthreads();
        }
        DBG_STAMP(6);
        if (cand && !(sm.u.d.flag[t] & 8)) {  // candidate and free: rewired.  Its children seed the write list
          my_q = atomicAdd(&sm.nrew, 1);
          KRRT_DCHECK(my_q >= 0 && my_q < GROW_THREADS && my_node >= 0 && my_node < T.cap, 103);
          sm.u.d.rewn[my_q] = my_node;
          sm.u.d.rslot[my_q] = t;
          sm.u.d.rpar[my_q] = my_par0;
          sm.u.d.rctc[my_q] = ctc_new + my_cb;
          sm.u.d.rtim[my_q] = tim_new + sm.u.d.tb[t];
          atomicMax(&W.wmin[my_node], key);
          atomicMax(&W.rmin[my_node], ((unsigned long long)key << 32) | (unsigned)t);
          push_children(T, sm, my_head, -(my_q + 1), ctc_new + my_cb, tim_new + sm.u.d.tb[t]);
        }
        __syncthreads();
        nrew = sm.nrew;
        DBG_STAMP(7);
        // level-order walk of everything below the rewired nodes: write marks.  A near node below one of this sample's
        // own rewired nodes would see a changed cost mid-expansion (intra): that sample takes the in-order path
        int level0 = 0;
        bool overflow = false;
        for (;;) {
          const int level1 = sm.tail;
          if (level1 > GROW_QCAP || (level1 > level0 && nlvl >= GROW_LEVELS)) {
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
          KRRT_DCHECK(level1 <= GROW_QCAP && nlvl <= GROW_LEVELS + 1 && cta < 1024, 104);  // write list, level table, wlc row
          for (int q = level0 + t; q < level1; q += GROW_THREADS) {
            const int d = sm.u.d.wl[q];
            KRRT_DCHECK(d >= 0 && d < T.cap, 105);
            atomicMax(&W.wmin[d], key);
            bool dead = false;
            for (unsigned hsl = near_hash(d);; hsl = (hsl + 1) & (GROW_HASH - 1)) {
              const int hv = sm.u.d.hash[hsl];
              if (hv == d) {
                // d is a rewire candidate below a node this sample rewires.  A CHILD of rewired node r is settled here
                // (rewires go in ascending index, each sees the costs the previous one left):
                //  d < r: d is tested first, at its old cost — as decided above; if it is rewired it leaves r's subtree
                //  d > r: r goes first and d's cost drops to r's new cost + d's edge; d is tested again at that cost
                // If d is rewired it is a child of the new node in its own right: its entry below r is dead.
                // Anything deeper, chains (r itself a candidate child of a rewired node) and a rewired d that fails
                // the second test take the in-order path.
                const int pr = sm.u.d.wl_par[q];
                if (!cfg.intra_fix || pr >= 0) {
                  sm.intra = 1;
                } else {
                  const int ks = sm.u.d.hslot[hsl];
                  const int rq = -pr - 1;
                  const bool kfree = !(sm.u.d.flag[ks] & 8);
                  atomicOr(&sm.u.d.flag[sm.u.d.rslot[rq]], 32);
                  atomicOr(&sm.u.d.flag[ks], 16);
                  if (d > sm.u.d.rewn[rq]) {
                    const double ck = sm.u.d.wl_ctc[q];  // == r's new cost + d's edge
                    if (ctc_new + sm.u.d.cb[ks] < ck) {
                      if (kfree) {  // r's propagation reaches d and its subtree, then d's own does: counted twice
                        atomicOr(&sm.u.d.flag[ks], 64);
                        sm.xflag = 1;
                      }
                    } else if (kfree) {
                      sm.intra = 1;
                    } else {
                      atomicAdd(&sm.tries_adj, 1);  // no longer a candidate: its (colliding) edge is never tested
                    }
                  }
                  if (kfree) {
                    dead = true;
                    sm.u.d.wl_par[q] = WL_DEAD;
                    atomicAdd(&sm.ndead, 1);
                  }
                }
              }
              if (hv == d || hv < 0) break;
            }
            if (!dead) {
              // the new cost goes on record for later samples of the sub-round that have d in their near set
              atomicMax(&W.rmin[d], ((unsigned long long)key << 32) | (unsigned)(256 + q));
              __stcg(&W.wlc[(size_t)cta * GROW_QCAP + q], sm.u.d.wl_ctc[q]);
              push_children(T, sm, sm.u.d.wl_head[q], q, sm.u.d.wl_ctc[q], sm.u.d.wl_tim[q]);
            }
          }
          level0 = level1;
          __syncthreads();
        }
        nwl = overflow ? 0 : level0;
        nprop = nwl;  // what a sequential planner counts as propagated
        if (!overflow && sm.ndead > 0) {
          if (my_q >= 0 && (sm.u.d.flag[t] & 48) == 48) sm.intra = 1;  // chain: the in-order path
          if (sm.xflag) {
            int cnt = (my_q >= 0 && (sm.u.d.flag[t] & 64)) ? 1 : 0;
            for (int q = t; q < nwl; q += GROW_THREADS) {
              int p = sm.u.d.wl_par[q];
              if (p == WL_DEAD) continue;
              while (p >= 0) p = sm.u.d.wl_par[p];  // (a dead entry has no children)
              if (sm.u.d.flag[sm.u.d.rslot[-p - 1]] & 64) ++cnt;
            }
            if (cnt) atomicAdd(&sm.xprop, cnt);
          }
          __syncthreads();
          nprop = nwl - sm.ndead + sm.xprop;
          collide_calls -= sm.tries_adj;
        } else if (!overflow && sm.tries_adj > 0) {
          collide_calls -= sm.tries_adj;
        }
        DBG_STAMP(8);
        const int intra = sm.intra;
        outcome = (overflow || intra) ? OUT_COMPLEX : OUT_INSERT;
        DBG_STAMP(9);
      } while (0);
      if (active && t == 0) {
        WaveRec r;
#ifdef KRRT_DEBUG_CLK
        r.pad = (int)(clock64() - dec_begin);
        if (W.dbg_log) {
          const int e_ = atomicAdd(&W.dbg_log[0], 1);
          if (e_ < (1 << 20)) {
            int* L_ = W.dbg_log + 8 + 8 * (size_t)e_;
            L_[0] = cta; L_[1] = r.pad; L_[2] = outcome; L_[3] = K; L_[4] = __ldcg(&Bt.count[b]); L_[5] = nrew; L_[6] = nwl; L_[7] = nlvl;
          }
        }
#else
        r.pad = 0;
#endif
        r.outcome = outcome; r.n_b = n_b; r.par = par; r.park = park; r.K = K; r.nrew = nrew; r.nwl = nprop;
        r.ctc_new = ctc_new; r.tim_new = tim_new; r.collide_calls = collide_calls; r.d2_in_max = KR(0.0); r.d2_out_min = KR(0.0);
        W.rec[b] = r;
      }
#ifdef KRRT_DEBUG_CLK
      if (active || last) cta_dec += clock64() - cta_t;
#endif
      DBG_START();
      const unsigned tag = round + 1u;
      if (cfg.flow) {
        // Validation of sample c looks at what the samples BEFORE it decided (records, marks of lower samples): it waits
        // for those only, not for the whole sub-round — a straggler (a sample with a deep subtree to walk) holds up the
        // samples behind it, and those are the ones least likely to make it into the prefix anyway
        if (active) {
          __syncthreads();  // every mark of this CTA is issued, the record is written
          if (t == 0) {
            __threadfence();
            st_release_u32(&W.done[cta], tag);
          }
          if (t < cta) {
            unsigned spins = 0;
            unsigned long long t_first = 0;
            while (ld_acquire_u32(&W.done[t]) != tag) flag_watchdog(spins, t_first, "wait for earlier decisions");
          }
          __syncthreads();
        }
      } else {
        grid_barrier(cfg.bar + 1, bs_sub, Pn + 1, false);
      }
      DBG_STAMP(12);
      phase_done(2);
#ifdef KRRT_DEBUG_CLK
      cta_t = clock64();
#endif

      // -------------------------------------------------------------- validate
#ifdef KRRT_WHY  // diagnostic build: what made the sample that ended the prefix fail (counted after F is known)
      int why_bits = 0;
#define WHY(cond, bit) do { if (cond) why_bits |= (bit); } while (0)
#else
#define WHY(cond, bit)
#endif
      int n_true = n_b;
      if (active) {
        bool fail = outcome == OUT_COMPLEX || outcome == OUT_CONFLICT || outcome == OUT_NEAREST;
        if (outcome == OUT_NOFREE || outcome == OUT_INSERT) {
          // the node index / tree size this sample really gets: earlier samples that found no free parent did not insert.
          // The decision was taken at the radius of n_b; it stands if no candidate changes sides between the two radii.
          // one round trip for everything that is addressed by what decide left behind: the earlier samples' outcomes, the
          // marks on this thread's near node, the record of its same-sub-round neighbour, the marks on the nodes it writes
          const int o_t = t < cta ? __ldcg(&W.rec[F0 + t].outcome) : -1;
          unsigned mw = 0, mk_w = 0xffffffffu, mk_a = 0xffffffffu, ml_w = 0xffffffffu, ml_a = 0xffffffffu;
          unsigned long long mr = 0;
          if (my_near) {
            mw = __ldcg(&W.wmin[my_node]);
            mr = __ldcg(&W.rmin[my_node]);
          }
          int oe = -1;
          double ce_p = KR(0.0), te_p = KR(0.0);
          State4 xe_p = {0, 0, 0, 0};
          if (pend_e >= 0) {
            oe = __ldcg(&W.rec[pend_e].outcome);
            ce_p = __ldcg(&W.rec[pend_e].ctc_new);
            te_p = __ldcg(&W.rec[pend_e].tim_new);
            xe_p = Bt.samples[pend_e];
          }
          const bool chk_r = outcome == OUT_INSERT && t < nrew, chk_l = outcome == OUT_INSERT && t < nwl;
          if (chk_r) {
            const int j = sm.u.d.rewn[t];
            mk_w = __ldcg(&W.wmin[j]);
            mk_a = __ldcg(&W.amin[j]);
          }
          if (chk_l) {
            const int j = sm.u.d.wl[t];
            ml_w = __ldcg(&W.wmin[j]);
            ml_a = __ldcg(&W.amin[j]);
          }
          const int ins = __syncthreads_count(o_t == OUT_INSERT);
          n_true = nsub + ins;
          const double rt = sm.rho[ins];  // == P.rho2[n_true]: n_true < capacity, or n_b would have been too
          bool conf = false;
          if (n_true != n_b && my_node >= 0) conf = (my_d <= rho2n) != (my_d <= rt);
          WHY(conf, 1);
          // A near node whose cost an earlier sample of the sub-round changes.  If that sample REWIRED it, the new cost is
          // that sample's cost + c*(sample -> node), both on record (the rewire mark carries the node's slot in the writer's
          // candidate list); it only went down, so this decision stands unless the node is the parent, was tried as parent, is a rewire candidate, or
          // would now be tried before the chosen parent.  A cost changed by propagation ends the prefix as before.
          if (my_near) {
            const int w = mark_sample(mw, round);
            if (w < cta) {
              bool c = true;
              if (cfg.raw_forward && outcome == OUT_INSERT && mark_sample((unsigned)(mr >> 32), round) == w &&
                  t != park && !my_tried && !my_cand) {
                const unsigned where = (unsigned)mr;
                if (where < 256u) {
                  // rewired: new cost of the node = the rewiring sample's cost + its c*(sample -> node), both on record
                  const double ce = __ldcg(&W.rec[F0 + w].ctc_new);
                  const double c_e = __ldcg(&Bt.cb[(size_t)(F0 + w) * cap + where]);
                  const double cn = (ce + c_e) + my_cf;
                  c = cn < ctc_new || (cn == ctc_new && my_node < par);
                } else if (cfg.prop_forward) {
                  // below a rewired node: the writer left the propagated cost in its row
                  const double cn = __ldcg(&W.wlc[(size_t)w * GROW_QCAP + (where - 256u)]) + my_cf;
                  c = cn < ctc_new || (cn == ctc_new && my_node < par);
                }
              }
              conf |= c;
              WHY(c, 2);
              WHY(c && mark_sample((unsigned)(mr >> 32), round) != w, 256);  // ... by propagation, not by a rewire
            }
          }
          // A candidate that is an earlier sample of this very sub-round: its decision is on record now.  If it inserts
          // inside this sample's radius it joins the near set (counters), and the decision taken without it stands as
          // long as the newcomer would be neither the parent (it is tried before the chosen one only if strictly cheaper:
          // it has the highest index) nor a rewire candidate, and is not a duplicate.  Nobody changes its cost in between:
          // a later sample that rewires it fails here, one that rewires an ancestor fails on the read marks of its parent.
          bool joins = false, need_a = false, need_b = false;
          if (pend_e >= 0) {
            const double ce = ce_p;
            if (oe == OUT_INSERT && my_d <= rt) {
              joins = true;
              conf |= my_d == KR(0.0);
              // the newcomer would be tried as parent (all candidates are if none is free) / is a rewire candidate
              need_a = outcome == OUT_NOFREE || ce + my_cf < ctc_new;
              need_b = outcome == OUT_INSERT && ctc_new + my_cb < ce;
              if (!cfg.join_test) {
                conf |= need_a || need_b;
                need_a = need_b = false;
              }
              WHY(outcome == OUT_NOFREE, 4); WHY(my_d == KR(0.0), 8); WHY(ce + my_cf < ctc_new, 16); WHY(ctc_new + my_cb < ce, 32);
            }
          }
          const int n_join = __syncthreads_count(joins);
          // If that edge COLLIDES the decision stands — a sequential planner tests it (one more collision test) and moves on;
          // if it is free the sample has another parent, or rewires the newcomer: redone in the next sub-round.
          if (__syncthreads_or(need_a || need_b)) {
            int my_x = -1;
            if (need_a || need_b) {
              my_x = atomicAdd(&sm.nxe, 1);
              KRRT_DCHECK(my_x >= 0 && pend_e >= F0 && pend_e < F0 + cta, 110);  // a newcomer is an EARLIER sample of this sub-round
              if (my_x < GROW_XE) {
                sm.xe_a[my_x] = need_a ? xe_p : s;
                sm.xe_b[my_x] = need_a ? s : xe_p;
                sm.xe_tau[my_x] = need_a ? pend_tf : pend_tb;
                sm.xe_t0[my_x] = need_a ? te_p : tim_new;
                sm.xe_hit[my_x] = 0;
              } else {
                conf = true;
              }
            }
            __syncthreads();
            const int ne = min(sm.nxe, GROW_XE);
            for (int w = t; w < ne * S; w += GROW_THREADS) {
              const int q = w / S, p = w - q * S;
              const Edge ed = make_edge(sm.xe_a[q], sm.xe_b[q], sm.xe_tau[q], sm.xe_t0[q]);
              const double h = S <= GROW_SMAX ? sm.hS[p] : ((S > 1) ? (double)p / (double)(S - 1) : KR(0.0));
              if (edge_point_hits_any(ed, h, sm.obs, P.m, P.dynamic, P.R2)) sm.xe_hit[q] = 1;
            }
            __syncthreads();
            if (my_x >= 0 && my_x < GROW_XE) {
              if (sm.xe_hit[my_x]) atomicAdd(&sm.xcalls, 1);
              else conf = true;
            }
            __syncthreads();
          }
          if (t == 0 && (n_join > 0 || sm.xcalls > 0)) {
            W.rec[b].K = K + n_join;
            W.rec[b].collide_calls = collide_calls + sm.xcalls;
          }
          if (outcome == OUT_INSERT) {
            // nodes this sample writes: none may have been written, or chosen as parent, by an earlier sample
            if (chk_r) {
              const bool c = min(mark_sample(mk_w, round), mark_sample(mk_a, round)) < cta;
              conf |= c;
              WHY(c, 64);
            }
            if (chk_l) {
              const bool c = min(mark_sample(ml_w, round), mark_sample(ml_a, round)) < cta;
              conf |= c;
              WHY(c, 128);
            }
            for (int q = t + GROW_THREADS; q < nwl; q += GROW_THREADS) {
              const int j = sm.u.d.wl[q];
              const bool c = min(mark_sample(__ldcg(&W.wmin[j]), round), mark_sample(__ldcg(&W.amin[j]), round)) < cta;
              conf |= c;
              WHY(c, 128);
            }
          }
          fail = __syncthreads_or(conf) != 0;
        }
        if (t == 0) {
          if (cfg.flow) {
            __threadfence();
            st_release_u32(&W.ok[cta], (tag << 1) | (fail ? 1u : 0u));
          } else if (fail) {
            atomicMin(&ctl->wave_Fs[slot], cta);
          }
        }
      }
#ifdef KRRT_DEBUG_CLK
      if (active || last) cta_val += clock64() - cta_t;
#endif
      int F_flow = nb;
      if (cfg.flow) {
        // the prefix: the first sample that failed validation, known as soon as every sample before it has passed —
        // every CTA of the sub-round works it out for itself from the same flags
        unsigned spins = 0, okv = 0;
        unsigned long long t_first = 0;
        bool have = false;
        for (;;) {
          if (t == 0) sm.tmpF = nb << 1;
          __syncthreads();
          if (t < nb) {
            if (!have) {  // (a flag that has been seen is not fetched again: 148 CTAs poll the same five lines)
              okv = ld_acquire_u32(&W.ok[t]);
              have = (okv >> 1) == tag;
            }
            if (!have) atomicMin(&sm.tmpF, (t << 1) | 1);       // not validated yet
            else if (okv & 1u) atomicMin(&sm.tmpF, t << 1);     // failed
          }
          __syncthreads();
          const int key = sm.tmpF;
          __syncthreads();
          if (!(key & 1)) {
            F_flow = key >> 1;
            break;
          }
          flag_watchdog(spins, t_first, "wait for the prefix");
        }
      } else {
        grid_barrier(cfg.bar + 1, bs_sub, Pn + 1, false);
      }
      DBG_STAMP(12);
      phase_done(3);
#ifdef KRRT_DEBUG_CLK
      cta_t = clock64();
#endif

      // -------------------------------------------------------------- apply
      const int F = cfg.flow ? F_flow : min(nb, __ldcg(&ctl->wave_Fs[slot]));
#ifdef KRRT_WHY
      if (active && cta == F) {
        if (why_bits) atomicOr(&sm.flags, why_bits << 8);  // (sm.flags is re-initialised by the next decide)
        __syncthreads();
        const int all = sm.flags >> 8;
        if (t < 8 && ((all >> t) & 1)) atomicAdd((unsigned long long*)&ctl->dbg_clk[t], 1ull);
        if (t == 9 && ((all >> 8) & 1)) atomicAdd((unsigned long long*)&ctl->dbg_clk[15], 1ull);
        if (t == 8) atomicAdd((unsigned long long*)&ctl->dbg_clk[8 + min(outcome, 6)], 1ull);  // (9..14: by outcome)
      }
#endif
      // the first sample has an empty near set: every CTA of the sub-round scans its share of the tree for the nearest
      // node (lowest index among equal distances), the last CTA folds the partials and commits the sample in order
      const bool need_nn = F == 0 && __ldcg(&W.rec[F0].outcome) == OUT_NEAREST;
      // the sample that ended the prefix needs the in-order path anyway (it rewires a node below another node it
      // rewires): the last CTA commits it right behind the prefix, one barrier later, instead of spending a whole
      // sub-round on it
      // — or has an empty near set: the grid-wide nearest-node scan runs behind the prefix as well
      const int o_tail = (cfg.tail_seq && F > 0 && F < nb) ? __ldcg(&W.rec[F0 + F].outcome) : OUT_INSERT;
      const bool tail_seq = o_tail == OUT_COMPLEX || o_tail == OUT_NEAREST;
      const bool tail_nn = o_tail == OUT_NEAREST;
      // this CTA's share of a nearest-node scan over nodes [0, n_scan) for window sample b_scan
      auto nn_scan = [&](int b_scan, int n_scan) {
        const State4 s0 = Bt.samples[b_scan];
        const int pidx = last ? Pn : cta;
        double bd = INFINITY;
        int bi = 0x7fffffff;
        for (int i = pidx * GROW_THREADS + t; i < n_scan; i += (Pn + 1) * GROW_THREADS) {
          const double d = dist2(load_state_cg(T, i), s0);
          if (d < bd) {
            bd = d;
            bi = i;
          }
        }
        // (distance, index) arg-min: distances are >= 0, their bit patterns order like unsigned integers
        const unsigned long long kb = real_key(bd);
        const unsigned hi = (unsigned)(kb >> 32), lo = (unsigned)kb;
        const unsigned mh = __reduce_min_sync(FULL, hi);
        const unsigned ml = __reduce_min_sync(FULL, hi == mh ? lo : 0xffffffffu);
        const unsigned mn = __reduce_min_sync(FULL, (hi == mh && lo == ml) ? (unsigned)bi : 0xffffffffu);
        if ((t & 31) == 0) {
          sm.red_c[t >> 5] = key_real(((unsigned long long)mh << 32) | ml);
          sm.red_i[t >> 5] = (int)mn;
        }
        __syncthreads();
        if (t == 0) {
          double c = sm.red_c[0];
          int ci = sm.red_i[0];
          for (int w = 1; w < GROW_THREADS / 32; ++w)
            if (sm.red_c[w] < c || (sm.red_c[w] == c && sm.red_i[w] < ci)) {
              c = sm.red_c[w];
              ci = sm.red_i[w];
            }
          W.nn_d2[pidx] = c;
          W.nn_idx[pidx] = ci;
        }
      };
      // the last CTA's first warp folds the partials
      auto nn_fold = [&](int& ni, double& nd2) {
        const int lane = t;
        nd2 = INFINITY;
        ni = 0x7fffffff;
        for (int k = lane; k <= Pn; k += 32) {
          const double c = __ldcg(&W.nn_d2[k]);
          const int ci = __ldcg(&W.nn_idx[k]);
          if (c < nd2 || (c == nd2 && ci < ni)) {
            nd2 = c;
            ni = ci;
          }
        }
        for (int off = 16; off > 0; off >>= 1) {
          const double c = __shfl_xor_sync(FULL, nd2, off);
          const int ci = __shfl_xor_sync(FULL, ni, off);
          if (c < nd2 || (c == nd2 && ci < ni)) {
            nd2 = c;
            ni = ci;
          }
        }
      };
      if (need_nn) {
        nn_scan(F0, nsub);
        grid_barrier(cfg.bar + 1, bs_sub, Pn + 1, false);
      }
      DBG_START();
      if (active && cta < F && outcome == OUT_INSERT) {
        const int nw = n_true;
        KRRT_DCHECK(nw > 0 && nw < P.capacity && par >= 0 && par < nw && park >= 0 && park < GROW_THREADS, 108);
        // the descendants' new cost-to-come / arrival time were worked out on the way down (decide)
        for (int q = t; q < nwl; q += GROW_THREADS) {
          if (sm.u.d.wl_par[q] == WL_DEAD) continue;
          const int d = sm.u.d.wl[q];
          __stcg(&T.ctc[d], sm.u.d.wl_ctc[q]);
          __stcg(&T.tim[d], sm.u.d.wl_tim[q]);
        }
        if (t == 0) {
          T.pos[nw] = make_double2(s.x, s.y);
          T.vel[nw] = make_double2(s.vx, s.vy);
          T.parent[nw] = par;
          T.ecost[nw] = sm.par_cf;
          T.etime[nw] = sm.u.d.tf[park];
          T.ctc[nw] = ctc_new;
          T.tim[nw] = tim_new;
          T.link[nw].x = nrew > 0 ? sm.u.d.rewn[0] : -1;  // the rewired nodes become the new node's children
          list_lock(W.lock, par);
          link_child_cg(T, par, nw);
          list_unlock(W.lock, par);
          Bt.map[b] = nw;
        }
        DBG_STAMP(10);
        if (nrew > 0) {
          // detach each rewired node from its old parent (one thread per warp takes the lock: no spinning inside a warp;
          // warps 1.. : warp 0's first lane is busy inserting the new node at the same time)
          if ((t & 31) == 0 && t > 0) {
            for (int q = (t >> 5) - 1; q < nrew; q += GROW_THREADS / 32 - 1) {
              const int j = sm.u.d.rewn[q];
              const int old = sm.u.d.rpar[q];  // (nobody else of the prefix has rewired j: that would be a conflict)
              KRRT_DCHECK(old >= 0 && old < T.cap && j > 0 && j < T.cap && __ldcg(&T.parent[j]) == old, 109);
              list_lock(W.lock, old);
              unlink_child_cg(T, old, j);
              list_unlock(W.lock, old);
            }
          }
          __syncthreads();
          if (my_q >= 0) {
            const int j = my_node;
            T.parent[j] = nw;
            T.link[j].y = my_q + 1 < nrew ? sm.u.d.rewn[my_q + 1] : -1;
            T.prev[j] = my_q > 0 ? sm.u.d.rewn[my_q - 1] : -1;
            T.ecost[j] = my_cb;
            T.etime[j] = sm.u.d.tb[t];
            T.ctc[j] = sm.u.d.rctc[my_q];
            T.tim[j] = sm.u.d.rtim[my_q];
          }
          DBG_STAMP(11);
        }
      }
      // -------------------------------------------------------------- finish (last CTA, one warp)
      if (last && t < 32) {
        const int lane = t;
#ifdef KRRT_DEBUG_CLK
        {
          int mx = 0;
          for (int k = lane; k < nb; k += 32) mx = max(mx, __ldcg(&W.rec[F0 + k].pad));
          for (int off = 16; off > 0; off >>= 1) mx = max(mx, __shfl_xor_sync(FULL, mx, off));
          if (lane == 0) {
            ctl->dbg_clk[14] += mx;
            ctl->dbg_clk[15] += nb;
          }
        }
#endif
        if (F > 0) {  // fold the counters of the prefix the other CTAs are committing
          long long v[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
          for (int k = lane; k < F; k += 32) {
            const WaveRec* rp = &W.rec[F0 + k];
            const int o = __ldcg(&rp->outcome), rk = __ldcg(&rp->K), rr = __ldcg(&rp->nrew), rw = __ldcg(&rp->nwl);
            const long long rc = __ldcg(&rp->collide_calls);
            v[0] += 1;
            if (o == OUT_INSERT) {
              v[1] += 1; v[5] += rr; v[6] += rw; v[7] += rk; v[8] += 2 * rk - 1; v[9] += rc;
            } else if (o == OUT_REJECT) {
              v[2] += 1;
            } else if (o == OUT_DUP) {
              v[3] += 1;
            } else {  // OUT_NOFREE
              v[4] += 1; v[7] += rk; v[8] += rk; v[9] += rc;
            }
          }
#pragma unroll
          for (int k = 0; k < 10; ++k)
            for (int off = 16; off > 0; off >>= 1) v[k] += __shfl_xor_sync(FULL, v[k], off);
          int o_next = OUT_INSERT;
          if (F < nb) o_next = __ldcg(&W.rec[F0 + F].outcome);
          if (lane == 0) {  // only stores on this path: the other CTAs wait for this warp at the barrier
            ctl->n = nsub + (int)v[1];
            ctl->iter = iter_sub + F;
            for (int k = 0; k < 10; ++k) sm.fin[k] += v[k];
            const State4 ls = Bt.samples[F0 + F - 1];
            ctl->last_sample[0] = ls.x; ctl->last_sample[1] = ls.y; ctl->last_sample[2] = ls.vx; ctl->last_sample[3] = ls.vy;
            ctl->ema16 = (3 * ema_sub + 16 * F) / 4;
            sm.fin[10] += 1;  // rounds_wave
            sm.fin[12] += F;  // wave_committed
            int why = 0;
            if (F < nb) why = o_next == OUT_CONFLICT ? 1 : (o_next == OUT_COMPLEX || o_next == OUT_NEAREST) ? 2 : (o_next == OUT_NOFREE ? 3 : 4);
            sm.fin[13 + why] += 1;
          }
        } else {  // the first sample needs the sequential path (empty near set, huge subtree, ...): commit it alone
          int ni = -1;
          double nd2 = KR(0.0);
          if (need_nn) nn_fold(ni, nd2);
          seq_commit<kLog>(ctl, P, T, Bt, dyn_smem, F0, F0 + 1, ni, nd2, cfg.log);
          if (lane == 0) {
            ctl->ema16 = (3 * ema_sub + 16) / 4;
            sm.fin[11] += 1;  // rounds_seq
          }
        }
      }
      if (tail_seq) grid_barrier(cfg.bar + 1, bs_sub, Pn + 1, false);  // the prefix is in the tree, ctl is up to date
      if (tail_nn) {
        nn_scan(F0 + F, __ldcg(&ctl->n));
        grid_barrier(cfg.bar + 1, bs_sub, Pn + 1, false);
      }
      if (last && t < 32) {
        if (tail_seq) {
          int ni = -1;
          double nd2 = KR(0.0);
          if (tail_nn) nn_fold(ni, nd2);
          seq_commit<kLog>(ctl, P, T, Bt, dyn_smem, F0 + F, F0 + F + 1, ni, nd2, cfg.log);
          if (t == 0) sm.fin[18] += 1;  // tail commits
        }
        __syncwarp();
        if (F0 + max(F, 1) + (tail_seq ? 1 : 0) >= Wn) clear_window(Bt);
        if (t == 0) {
          ctl->wave_Fs[slot ^ 1] = 0x7fffffff;
          ctl->round = round + 1;
        }
      }
#ifdef KRRT_DEBUG_CLK
      if (active || last) cta_app += clock64() - cta_t;
#endif
      DBG_START();
      grid_barrier(cfg.bar + 1, bs_sub, Pn + 1, false);
      DBG_STAMP(12);
      phase_done(4);
      F0 += max(F, 1) + (tail_seq ? 1 : 0);
    }
    grid_barrier(cfg.bar, bs_all, G, !takes_part);  // window end
    phase_done(4);
  }
  if (last && t == 0) {
    for (int k = 0; k < 10; ++k) ctl->stats[k] += sm.fin[k];
    ctl->rounds_wave += sm.fin[10];
    ctl->rounds_seq += sm.fin[11] + sm.fin[18];
    ctl->tail_commits += sm.fin[18];
    ctl->wave_committed += sm.fin[12];
    for (int k = 0; k < 5; ++k) ctl->why[k] += sm.fin[13 + k];
  }
#ifdef KRRT_DEBUG_CLK
  if (cta == 0 && t == 0)
    for (int k = 0; k < 14; ++k) ctl->dbg_clk[k] += dbg_acc[k];
  if (t == 0 && cta < 160) {
    ctl->dbg_cta[0][cta] += cta_dec;
    ctl->dbg_cta[1][cta] += cta_val;
    ctl->dbg_cta[2][cta] += cta_app;
  }
#endif
}

// a fresh tree holding only the root (rrt.setStart, robosoccer_motion_planning_node.cpp:135)
__global__ void k_set_start(Ctl* ctl, TreeView T, State4 x0) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  T.pos[0] = make_double2(x0.x, x0.y);
  T.vel[0] = make_double2(x0.vx, x0.vy);
  T.parent[0] = -1;
  T.ctc[0] = KR(0.0); T.tim[0] = KR(0.0); T.ecost[0] = KR(0.0); T.etime[0] = KR(0.0);
  T.link[0] = make_int2(-1, -1); T.prev[0] = -1;
  Ctl z;
  memset(&z, 0, sizeof(z));
  z.n = 1;
  z.goal_index = -1;
  z.round = ctl->round;  // the epoch of the conflict marks survives a restart (marks are never cleared)
  *ctl = z;
}

// before every k_grow launch: barrier counter, empty candidate lists, armed prefix slots; `first`: a new krrt_grow call
__global__ void k_begin_grow(Ctl* ctl, int first, int has_goal, State4 g, int n_iter, Batch Bt, GridBar* bar,
                             int reset_round, TreeView T, IterLog Lg) {
  for (int b = threadIdx.x; b < Bt.B; b += blockDim.x) {
    Bt.count[b] = 0;
    Bt.dup[b] = 0;
  }
  if (first && Lg.n_after) {  // logged call: row 0 of the snapshots is the tree before the call
    const int n = ctl->n;
    for (int j = threadIdx.x; j < n && j < Lg.stride; j += blockDim.x) {
      Lg.snap_ctc[j] = T.ctc[j];
      Lg.snap_tim[j] = T.tim[j];
    }
  }
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  bar[0].arrive = 0ull;
  bar[1].arrive = 0ull;
  ctl->wave_Fs[0] = 0x7fffffff;
  ctl->wave_Fs[1] = 0x7fffffff;
  if (reset_round) ctl->round = 0;
  if (!first) return;
  ctl->iter_begin = ctl->iter;
  ctl->iter_target = ctl->iter + n_iter;
  if (ctl->ema16 <= 0) ctl->ema16 = 16 * 16;
  ctl->full = 0;
  ctl->scan_pairs = KR(0.0);
  ctl->scan_nodes = KR(0.0);
  ctl->scan_visited = KR(0.0);
  ctl->windows = 0;
  for (int k = 0; k < 8; ++k) ctl->phase_clk[k] = 0;
  for (int k = 0; k < 16; ++k) ctl->dbg_clk[k] = 0;
  for (int k = 0; k < 3 * 160; ++k) (&ctl->dbg_cta[0][0])[k] = 0;
  ctl->has_goal = has_goal;
  ctl->goal[0] = g.x; ctl->goal[1] = g.y; ctl->goal[2] = g.vx; ctl->goal[3] = g.vy;
}

__global__ void __launch_bounds__(GROW_THREADS, KRRT_GROW_MINB) k_grow(Ctl* ctl, Params P, TreeView T, Batch Bt, Wave W,
                                                            GrowCfg cfg) {
  grow_body<false>(ctl, P, T, Bt, W, cfg);
}
// krrt_grow_logged: the same kernel with the in-order commit keeping per-iteration records (cfg.wave_enabled == 0)
__global__ void __launch_bounds__(GROW_THREADS, 1) k_grow_logged(Ctl* ctl, Params P, TreeView T, Batch Bt, Wave W, GrowCfg cfg) {
  grow_body<true>(ctl, P, T, Bt, W, cfg);
}

}  // namespace krrt
