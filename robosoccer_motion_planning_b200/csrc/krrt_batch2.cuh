// krrt_batch2.cuh — k_plan_batch2: the batched small-tree planner laid out for occupancy.
//
// Same contract as k_plan_batch (krrt_batch.cuh: one warp per query, the whole tree in shared memory, results and
// counters bit-identical with the sequential planner), replacing solve_rrt() (robosoccer_motion_planning_node.cpp:167-196)
// run once per robot / seed / start-goal pair.  What changed, and why (profiles/r1_plan_batch_s21.md: 9 warps per SM,
// every warp a chain of dependent fp64 instructions, issue slots 29 % busy, 2100 warp instructions per expansion):
//   * the near list (node, c/tau forward, c/tau backward) lives in REGISTERS, two candidates per lane, instead of five
//     shared-memory arrays of `capacity` entries; a near set of more than 64 nodes hands the query to k_plan_batch
//     (redo list), so the result never depends on the limit;
//   * child / sibling / parent / previous-sibling links and the propagation queue are 16-bit;
//     -> 74 B per node instead of 121 B: 15 warps per SM at 200 iterations instead of 9;
//   * the window of 32 future samples is drawn by the 32 lanes at once (Philox is counter-based) and handed out by
//     shuffles, instead of every lane drawing every sample;
//   * an edge is tested with lane = (sample point, obstacle group): the point is evaluated once per lane, h = k / (S - 1)
//     comes from a table, and there is no integer division in the loop;
//   * the winner of the parent / rewire selections is found with three REDUX instead of a five-step shuffle tree.
// Every floating-point operation, and the order of the operations that decide something, is that of krrt_warp.cuh.
#pragma once
#include "krrt_batch.cuh"

namespace krrt {

constexpr unsigned B2_NIL = 0xffffu;
constexpr int B2_SLOTS = 2;          // near candidates per lane (registers)
constexpr int B2_KMAX = 32 * B2_SLOTS;
constexpr int B2_HMAX = 64;          // edge sample counts up to this use the table of h = k / (S - 1)
#ifdef KRRT_F32_BUILD
#ifndef KRRT_B2_MAX_WARPS_F32
#define KRRT_B2_MAX_WARPS_F32 20   // 42 B per node: 26 warps of 200-node trees would fit; measured 16 / 20 / 24 / 32: 2.44 / 2.49 / 2.34 / 2.25 M
                                   // queries/s on C4 (profiles/r2_c4_warps_s4.txt) - a warp is a latency chain, more of them share the issue slots
#endif
#define KRRT_B2_MAX_WARPS KRRT_B2_MAX_WARPS_F32
#else
#define KRRT_B2_MAX_WARPS 16
#endif
constexpr int B2_MAX_WARPS = KRRT_B2_MAX_WARPS;

struct Batch2Job {
  BatchJob b;        // inputs, outputs, parameters (next_query / stats as in k_plan_batch)
  int* redo_list;    // queries whose near set outgrew the registers: planned by k_plan_batch afterwards
  int* redo_count;
};

struct B2Tree {
  double2* pos;
  double2* vel;
  double* ctc;
  double* tim;
  double* ecost;
  double* etime;
  unsigned short* child;   // first child
  unsigned short* sib;     // next sibling
  unsigned short* parent;
  unsigned short* prev;    // previous sibling
  unsigned short* scr;     // near-set node list while a sample is steered, then the propagation queue
  int cap;
};

__host__ __device__ inline size_t batch2_warp_smem_bytes(int cap, int m) {
  cap += 1;  // the goal node's slot
  size_t b = 0;
  b += (size_t)cap * sizeof(double2) * 2;               // pos, vel
  b += (size_t)cap * sizeof(double) * 4;                // ctc, tim, ecost, etime
  b += (size_t)(m > 0 ? m : 1) * sizeof(double4);       // obstacles
  b += (((size_t)cap + 7) / 8 * 8) * 2 * 5;             // child, sib, parent, prev, scr (each 16-byte aligned)
  return (b + 15) / 16 * 16;
}

__device__ __forceinline__ State4 b2_state(const B2Tree& T, int i) {
  const double2 p = T.pos[i], v = T.vel[i];
  State4 s = {p.x, p.y, v.x, v.y};
  return s;
}

__device__ __forceinline__ State4 goal_load_state(const B2Tree& T, int i) { return b2_state(T, i); }

#ifdef KRRT_F32_BUILD
// binary32 build: the key is one word — two REDUX + one ballot
__device__ __forceinline__ unsigned b2_key(double x) {
  const unsigned u = __float_as_uint(x);
  return (u >> 31) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ unsigned b2_argmin(bool have, double key, unsigned idx, int& owner) {
  const unsigned kb = have ? b2_key(key) : 0xffffffffu;
  const unsigned mk = __reduce_min_sync(FULL, kb);
  const bool top = have && kb == mk;
  const unsigned mn = __reduce_min_sync(FULL, top ? idx : 0xffffffffu);
  const unsigned own = __ballot_sync(FULL, top && idx == mn);
  owner = own ? (__ffs(own) - 1) : 0;
  return mn;
}
#else
// order-preserving map of a double onto an unsigned 64-bit key (total order of the finite values and infinities)
__device__ __forceinline__ unsigned long long b2_key(double x) {
  const unsigned long long u = (unsigned long long)__double_as_longlong(x);
  return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}

// lexicographic warp arg-min of (key, idx): returns the winning idx (0xffffffff when no lane has a candidate) and
// the lane that owns it.  Three REDUX + one ballot.
__device__ __forceinline__ unsigned b2_argmin(bool have, double key, unsigned idx, int& owner) {
  const unsigned long long kb = have ? b2_key(key) : ~0ull;
  const unsigned hi = (unsigned)(kb >> 32), lo = (unsigned)kb;
  const unsigned mh = __reduce_min_sync(FULL, hi);
  const unsigned ml = __reduce_min_sync(FULL, hi == mh ? lo : 0xffffffffu);
  const bool top = have && hi == mh && lo == ml;
  const unsigned mn = __reduce_min_sync(FULL, top ? idx : 0xffffffffu);
  const unsigned own = __ballot_sync(FULL, top && idx == mn);
  owner = own ? (__ffs(own) - 1) : 0;
  return mn;
}
#endif

// lane -> (sample point, obstacle group) map of the edge test
struct B2Lanes {
  int k0, g, G, kstep;   // first sample point of this lane, its obstacle group, groups, point stride (k0 < 0: lane idle)
};
__device__ __forceinline__ B2Lanes b2_lanes(int S, int m) {
  B2Lanes L;
  int G = S > 0 ? 32 / S : 1;
  if (G > m) G = m;
  if (G < 1) G = 1;
  L.G = G;
  L.kstep = 32 / G;
  const int lane = threadIdx.x & 31;
  L.k0 = lane < G * L.kstep ? lane / G : -1;
  L.g = lane - (lane / G) * G;
  return L;
}

// S x m disc tests of the edge a -> b (arrival time tau, departure at t0); uniform result.  The same operations per
// (point, obstacle) pair as edge_point_hits (krrt_math.cuh), OR-ed in another order.
__device__ __forceinline__ bool b2_edge_collide(const State4& a, const State4& b, double tau, double t0,
                                                const EnvView& E, const B2Lanes& L, const double* s_h) {
  const Edge e = make_edge(a, b, tau, t0);
  bool hit = false;
  if (L.k0 >= 0) {
    for (int k = L.k0; k < E.S; k += L.kstep) {
      const double h = E.S <= B2_HMAX ? s_h[k] : ((E.S > 1) ? (double)k / (double)(E.S - 1) : KR(0.0));
      const double px = axis_pos(e.kx, h), py = axis_pos(e.ky, h);
      const double t = e.t0 + e.tau * h;
      for (int j = L.g; j < E.m; j += L.G) {
        const double4 o = E.obs[j];
        double ox = o.x, oy = o.y;
        if (E.dynamic) {
          ox = fma(o.z, t, ox);
          oy = fma(o.w, t, oy);
        }
        const double dx = px - ox, dy = py - oy;
        hit |= fma(dy, dy, dx * dx) < E.R2;
      }
    }
  }
  return __any_sync(FULL, hit);
}

// child-list maintenance (one lane)
__device__ __forceinline__ void b2_link_child(const B2Tree& T, int par, int c) {
  const unsigned h = T.child[par];
  T.sib[c] = (unsigned short)h;
  T.prev[c] = (unsigned short)B2_NIL;
  if (h != B2_NIL) T.prev[h] = (unsigned short)c;
  T.child[par] = (unsigned short)c;
}
__device__ __forceinline__ void b2_unlink_child(const B2Tree& T, int c) {
  const unsigned par = T.parent[c], nx = T.sib[c], pv = T.prev[c];
  if (pv != B2_NIL) T.sib[pv] = (unsigned short)nx; else T.child[par] = (unsigned short)nx;
  if (nx != B2_NIL) T.prev[nx] = (unsigned short)pv;
}

// cost-to-come and arrival time of every descendant of j from its parent: level order, 32 nodes per step
// (warp_propagate of krrt_warp.cuh with 16-bit links; returns the number of descendants)
__device__ __forceinline__ int b2_propagate(const B2Tree& T, int j) {
  const int lane = threadIdx.x & 31;
  int qhead = 0, qtail = 0;
  if (lane == 0) {
    unsigned c = T.child[j];
    while (c != B2_NIL && qtail < T.cap) {
      T.scr[qtail++] = (unsigned short)c;
      c = T.sib[c];
    }
  }
  qtail = __shfl_sync(FULL, qtail, 0);
  __syncwarp();
  while (qhead < qtail) {
    const int cnt = min(32, qtail - qhead);
    unsigned c = B2_NIL;
    if (lane < cnt) {
      const int d = T.scr[qhead + lane];
      const int par = T.parent[d];
      T.ctc[d] = T.ctc[par] + T.ecost[d];
      T.tim[d] = T.tim[par] + T.etime[d];
      c = T.child[d];
    }
    qhead += cnt;
    for (;;) {
      const unsigned mk = __ballot_sync(FULL, c != B2_NIL);
      if (mk == 0 || qtail + __popc(mk) > T.cap) break;  // (exact bound: see warp_propagate)
      if (c != B2_NIL) {
        T.scr[qtail + __popc(mk & ((1u << lane) - 1))] = (unsigned short)c;
        c = T.sib[c];
      }
      qtail += __popc(mk);
    }
    __syncwarp();
  }
  return qtail;
}


__global__ void __launch_bounds__(B2_MAX_WARPS * 32, 1) k_plan_batch2(Batch2Job JJ) {
  const BatchJob& J = JJ.b;
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ double s_h[B2_HMAX];
  for (int k = threadIdx.x; k < B2_HMAX; k += blockDim.x) s_h[k] = (J.S > 1) ? (double)k / (double)(J.S - 1) : KR(0.0);
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int cap = J.cap + 1;  // array stride: the tree's J.cap nodes + the goal node's slot
  const size_t cap8 = ((size_t)cap + 7) / 8 * 8;
  unsigned char* base = smem + (size_t)warp * batch2_warp_smem_bytes(J.cap, J.m);
  B2Tree T;
  T.pos = reinterpret_cast<double2*>(base);
  T.vel = T.pos + cap;
  T.ctc = reinterpret_cast<double*>(T.vel + cap);
  T.tim = T.ctc + cap;
  T.ecost = T.tim + cap;
  T.etime = T.ecost + cap;
  double4* s_obs = reinterpret_cast<double4*>(T.etime + cap);
  T.child = reinterpret_cast<unsigned short*>(s_obs + (J.m > 0 ? J.m : 1));
  T.sib = T.child + cap8;
  T.parent = T.sib + cap8;
  T.prev = T.parent + cap8;
  T.scr = T.prev + cap8;
  T.cap = J.cap;
  const EnvView E = {s_obs, J.m, J.dynamic, J.R2, J.S};
  const B2Lanes L = b2_lanes(J.S, J.m);
  const unsigned lt_mask = (1u << lane) - 1;

  for (;;) {
    int qi = 0;
    if (lane == 0) qi = atomicAdd(J.next_query, 1);
    qi = __shfl_sync(FULL, qi, 0);
    if (qi >= J.q) break;
    const uint32_t query_id = J.query_base + (uint32_t)qi;
    for (int j = lane; j < J.m; j += 32) {
      const krrt_io_real* o = J.obstacles + ((size_t)qi * J.m + j) * 4;
      s_obs[j] = make_double4((double)o[0], (double)o[1], (double)o[2], (double)o[3]);
    }
    if (lane == 0) {
      const krrt_io_real* xs = J.starts + 4 * (size_t)qi;
      T.pos[0] = make_double2((double)xs[0], (double)xs[1]);
      T.vel[0] = make_double2((double)xs[2], (double)xs[3]);
      T.parent[0] = (unsigned short)B2_NIL;
      T.ctc[0] = KR(0.0); T.tim[0] = KR(0.0); T.ecost[0] = KR(0.0); T.etime[0] = KR(0.0);
      T.child[0] = (unsigned short)B2_NIL; T.sib[0] = (unsigned short)B2_NIL; T.prev[0] = (unsigned short)B2_NIL;
    }
    __syncwarp();
    // the query's counters (warp-uniform; added to the job's totals when the query completes here)
    int c_iters = 0, c_ins = 0, c_rej = 0, c_dup = 0, c_nofree = 0, c_rew = 0, c_prop = 0, c_near = 0, c_steer = 0,
        c_coll = 0;
    int n = 1;
    bool overflow = false, full = false;
    for (int it0 = 0; it0 < J.n_iter && !full && !overflow; it0 += 32) {
      // the next 32 samples, one per lane
      State4 mine;
      {
        const krrt_io_real* gs = J.goals + 4 * (size_t)qi;
        const State4 goal = {(double)gs[0], (double)gs[1], (double)gs[2], (double)gs[3]};
        mine = draw_sample(J.sp, query_id, (uint64_t)(it0 + lane), true, goal);
      }
      const int lim = min(32, J.n_iter - it0);
      for (int l = 0; l < lim; ++l) {
        State4 s;
        s.x = __shfl_sync(FULL, mine.x, l);
        s.y = __shfl_sync(FULL, mine.y, l);
        s.vx = __shfl_sync(FULL, mine.vx, l);
        s.vy = __shfl_sync(FULL, mine.vy, l);
        c_iters++;
        if (!J.dynamic && warp_point_in_obstacle(s, E)) {
          c_rej++;
          continue;
        }
        if (n >= J.cap) {
          full = true;
          break;
        }
        const double rho2n = (double)J.rho2[n];
        // near set (ascending by construction) + nearest + duplicate detection
        int K = 0;
        bool dup = false;
        double bd = INFINITY;
        int bi = -1;
        for (int b0 = 0; b0 < n; b0 += 32) {
          const int i = b0 + lane;
          bool valid = false;
          if (i < n) {
            const double d = dist2(b2_state(T, i), s);
            if (bi < 0 || d < bd) {
              bd = d;
              bi = i;
            }
            valid = d <= rho2n;
            dup |= d == KR(0.0);
          }
          const unsigned mk = __ballot_sync(FULL, valid);
          if (valid) T.scr[K + __popc(mk & lt_mask)] = (unsigned short)i;
          K += __popc(mk);
        }
        if (__any_sync(FULL, dup)) {
          c_dup++;
          continue;
        }
        if (K == 0) {  // empty near set: the nearest node is the only candidate (lowest index among equal distances)
          int owner;
          const unsigned w = b2_argmin(bi >= 0, bd, (unsigned)bi, owner);
          if (lane == 0) T.scr[0] = (unsigned short)w;
          K = 1;
        }
        __syncwarp();
        if (K > B2_KMAX) {  // does not fit the registers: the whole query goes to k_plan_batch
          overflow = true;
          break;
        }
        // this lane's candidates: k = lane and lane + 32 of the near list
        const int ci0 = lane < K ? (int)T.scr[lane] : -1;
        const int ci1 = lane + 32 < K ? (int)T.scr[lane + 32] : -1;
        double cf0 = KR(0.0), tf0 = KR(0.0), cb0 = KR(0.0), tb0 = KR(0.0), cf1 = KR(0.0), tf1 = KR(0.0), cb1 = KR(0.0), tb1 = KR(0.0);
        // both steering directions of every near node: 2K independent tasks over the lanes (task t < K: node -> sample of
        // candidate t; task t >= K: sample -> node of candidate t - K), ONE call site so the lanes iterate Newton together;
        // a backward result is handed to the lane that owns the candidate by a shuffle
        const int npass = (2 * K + 31) >> 5;
        for (int p = 0; p < npass; ++p) {
          const int t = lane + 32 * p;
          const bool valid = t < 2 * K;
          const bool fwd = t < K;
          double tau = KR(0.0), c = KR(0.0);
          if (valid) {
            const State4 x = b2_state(T, T.scr[fwd ? t : t - K]);
            steer(fwd ? x : s, fwd ? s : x, J.r, tau, c);
          }
          if (fwd) {  // t = lane + 32 p is this lane's own candidate of slot p (p <= 1 because K <= 64)
            if (p == 0) { tf0 = tau; cf0 = c; } else { tf1 = tau; cf1 = c; }
          }
          // backward task of candidate k runs as task K + k: pass (K + k) >> 5 on lane (K + k) & 31
          {
            const int t0b = K + lane, t1b = K + lane + 32;
            const double a_tau = __shfl_sync(FULL, tau, t0b & 31), a_c = __shfl_sync(FULL, c, t0b & 31);
            if ((t0b >> 5) == p && ci0 >= 0) { tb0 = a_tau; cb0 = a_c; }
            if ((t1b >> 5) == p && ci1 >= 0) { tb1 = a_tau; cb1 = a_c; }   // same source lane: (K + lane + 32) & 31
          }
        }
        __syncwarp();  // the node list in T.scr is dead from here on (the propagation queue reuses it)
        c_near += K;

        // ---------------------------------------------------------------- commit (warp_commit of krrt_warp.cuh)
        c_steer += K;
        // parent = arg-min over collision-free candidates of ctc[i] + c*(i -> new); lowest index on ties
        int par = -1;
        double cfp = KR(0.0), tfp = KR(0.0);
        unsigned tried = 0;
        for (;;) {
          double bc = INFINITY;
          int bn = -1, bslot = 0;
          if (ci0 >= 0 && !(tried & 1u)) {
            bc = T.ctc[ci0] + cf0;
            bn = ci0;
          }
          if (ci1 >= 0 && !(tried & 2u)) {
            const double cand = T.ctc[ci1] + cf1;
            if (bn < 0 || cand < bc) {  // ci1 > ci0: an equal cost keeps the lower index
              bc = cand;
              bn = ci1;
              bslot = 1;
            }
          }
          int owner;
          const unsigned w = b2_argmin(bn >= 0, bc, (unsigned)bn, owner);
          if (w == 0xffffffffu) break;
          if (lane == owner) tried |= 1u << bslot;
          const double w_tf = __shfl_sync(FULL, bslot ? tf1 : tf0, owner);
          const double w_cf = __shfl_sync(FULL, bslot ? cf1 : cf0, owner);
          c_coll++;
          if (!b2_edge_collide(b2_state(T, (int)w), s, w_tf, T.tim[w], E, L, s_h)) {
            par = (int)w;
            cfp = w_cf;
            tfp = w_tf;
            break;
          }
        }
        if (par < 0) {
          c_nofree++;
          continue;
        }
        const int nw = n;
        const double ctc_new = T.ctc[par] + cfp;
        const double tim_new = T.tim[par] + tfp;
        if (lane == 0) {
          T.pos[nw] = make_double2(s.x, s.y);
          T.vel[nw] = make_double2(s.vx, s.vy);
          T.parent[nw] = (unsigned short)par;
          T.ecost[nw] = cfp;
          T.etime[nw] = tfp;
          T.ctc[nw] = ctc_new;
          T.tim[nw] = tim_new;
          T.child[nw] = (unsigned short)B2_NIL;
          b2_link_child(T, par, nw);
        }
        __syncwarp();
        c_ins++;
        c_steer += K - 1;
        // rewire in ascending node index; each decision sees the costs left by the previous one
        int last = -1;
        for (;;) {
          unsigned bj = 0xffffffffu;
          int bslot = 0;
          if (ci0 > last && ci0 != par && ctc_new + cb0 < T.ctc[ci0]) bj = (unsigned)ci0;
          if (bj == 0xffffffffu && ci1 > last && ci1 != par && ctc_new + cb1 < T.ctc[ci1]) {
            bj = (unsigned)ci1;
            bslot = 1;
          }
          const unsigned w = __reduce_min_sync(FULL, bj);
          if (w == 0xffffffffu) break;
          last = (int)w;
          const unsigned own = __ballot_sync(FULL, bj == w);
          const int owner = __ffs(own) - 1;
          const double w_tb = __shfl_sync(FULL, bslot ? tb1 : tb0, owner);
          const double w_cb = __shfl_sync(FULL, bslot ? cb1 : cb0, owner);
          c_coll++;
          if (!b2_edge_collide(s, b2_state(T, (int)w), w_tb, tim_new, E, L, s_h)) {
            if (lane == 0) {
              b2_unlink_child(T, (int)w);
              T.parent[w] = (unsigned short)nw;
              b2_link_child(T, nw, (int)w);
              T.ecost[w] = w_cb;
              T.etime[w] = w_tb;
              T.ctc[w] = ctc_new + w_cb;
              T.tim[w] = tim_new + w_tb;
            }
            __syncwarp();
            c_rew++;
            c_prop += b2_propagate(T, (int)w);
          }
        }
        n = nw + 1;
        __syncwarp();
      }
    }
    if (overflow) {  // nothing of this query is reported from here: k_plan_batch plans it from scratch
      if (lane == 0) JJ.redo_list[atomicAdd(JJ.redo_count, 1)] = qi;
      __syncwarp();
      continue;
    }
    // goal node: connected to the cheapest collision-free offer of the tree (default), or the cheapest node within
    // goal_tol of the goal (warp_find_goal); lowest index on ties
    int gi;
    {
      const krrt_io_real* gs = J.goals + 4 * (size_t)qi;
      const State4 goal = {(double)gs[0], (double)gs[1], (double)gs[2], (double)gs[3]};
      if (J.goal_mode == GOAL_MODE_TOLERANCE) {
        double bc = INFINITY;
        int bn = -1;
        for (int i = lane; i < n; i += 32) {
          if (dist2(b2_state(T, i), goal) <= J.goal_tol2) {
            const double c = T.ctc[i];
            if (c < bc) {
              bc = c;
              bn = i;
            }
          }
        }
        int owner;
        const unsigned w = b2_argmin(bn >= 0, bc, (unsigned)bn, owner);
        gi = w == 0xffffffffu ? -1 : (int)w;
      } else {
        int gpar = -1;
        gi = warp_goal_connect(T, n, goal, J.r, E, &gpar);
        if (gi == n && lane == 0) T.parent[n] = (unsigned short)gpar;
        __syncwarp();
      }
    }
    ResultRec* R = J.out + qi;
    krrt_io_real* wp = &R->waypoints[0][0];
    for (int k = lane; k < 32 * 5; k += 32) wp[k] = KR(0.0);
    __syncwarp();
    if (lane == 0) {
      R->solved = gi >= 0;
      R->goal_index = gi;
      R->n_nodes = n;
      int len = 0;
      for (int i = gi; i >= 0 && len <= n; i = (T.parent[i] == B2_NIL ? -1 : (int)T.parent[i])) ++len;
      R->path_len = len;
      R->cost = gi >= 0 ? T.ctc[gi] : KR(0.0);
      R->time = gi >= 0 ? T.tim[gi] : KR(0.0);
      int p = len - 1;
      for (int i = gi; i >= 0 && p >= 0; i = (T.parent[i] == B2_NIL ? -1 : (int)T.parent[i]), --p) {
        if (p < 32) {
          const double2 a = T.pos[i], v = T.vel[i];
          R->waypoints[p][0] = a.x; R->waypoints[p][1] = a.y;
          R->waypoints[p][2] = v.x; R->waypoints[p][3] = v.y;
          R->waypoints[p][4] = T.tim[i];
        }
      }
      if (J.stats) {
        const int cv[10] = {c_iters, c_ins, c_rej, c_dup, c_nofree, c_rew, c_prop, c_near, c_steer, c_coll};
#pragma unroll
        for (int k = 0; k < 10; ++k)
          if (cv[k]) atomicAdd(reinterpret_cast<unsigned long long*>(J.stats + k), (unsigned long long)cv[k]);
      }
    }
    __syncwarp();
    batch_query_done(J, qi);
  }
}

}  // namespace krrt
