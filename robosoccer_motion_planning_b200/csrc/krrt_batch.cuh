// krrt_batch.cuh — k_plan_batch: many independent small planning queries, one warp per query, the whole tree
// (states, costs, child lists, near list) in shared memory.  Replaces solve_rrt()
// (robosoccer_motion_planning_node.cpp:167-196: setStart, setObstacles, for(i<iteration) grow(xg)) run once per
// robot / seed / start-goal pair; the live system plans ~75-200-node trees at a high rate (config.yaml:3).
// Warps pull queries from a global counter, so the grid is sized to the machine, not to q.
#pragma once
#include "krrt_warp.cuh"

namespace krrt {

struct ResultRec {  // layout == krrt_result (include/krrt.h): binary64 in both precisions (krrt_real.cuh)
  int solved, goal_index, n_nodes, path_len;
  krrt_io_real cost, time;
  krrt_io_real waypoints[32][5];
};

struct BatchJob {
  const krrt_io_real* starts;     // q x 4   (inputs and the radius table are binary64 in both precisions: rounded to the
  const krrt_io_real* goals;      // q x 4    arithmetic type when loaded)
  const krrt_io_real* obstacles;  // q x m x 4
  ResultRec* out;           // q
  int* next_query;          // work counter
  const int* list;          // when set: plan only the queries list[0 .. *list_count) (k_plan_batch2's redo list)
  const int* list_count;
  long long* stats;         // [10] accumulated over queries
  const krrt_io_real* rho2;       // [cap + 1]
  int q, m, dynamic, n_iter, cap;  // cap: most nodes a tree can hold; the arrays have cap + 1 slots (goal node)
  uint32_t query_base;
  double r, R2, goal_tol2;
  int S;
  int goal_mode;  // KRRT_GOAL_CONNECT / KRRT_GOAL_TOLERANCE (include/krrt.h)
  // completion counters for a consumer on another stream (krrt_plan_batch_device_chunks): chunk_done[i / chunk_queries]
  // is incremented once query i's record is visible device-wide (nullptr: none)
  unsigned* chunk_done;
  int chunk_queries;
  SamplerParams sp;
};

constexpr int GOAL_MODE_TOLERANCE = 1;  // == KRRT_GOAL_TOLERANCE

// Goal connection at the end of a query (k_goal_connect of krrt_kernels.cuh for one warp; T in shared memory): every
// inserted node offers node -> goal, the cheapest collision-free offer wins (lowest index among equal costs), the goal
// node is written to slot n (all but its parent link, whose type is the caller's: goal_parent) unless a node that
// coincides with the goal wins.  Returns the goal index or -1.
// (not inlined, and handed COPIES of the views: the planner's own tree view stays in registers — taking its address
// would move it to local memory for the whole kernel)
__device__ __forceinline__ State4 goal_load_state(const TreeView& T, int i) { return load_state(T, i); }

// The goal's parent is the arg-min of (cost through the node, node index) over the COLLISION-FREE offers — the first free
// offer a sequential planner meets when it walks the offers sorted by (cost, index).  No counter depends on how many
// offers were tested, so every lane settles its own nodes (i = 1 + lane, + 32, ...): steer, and test the edge — all S x m
// disc tests by this lane alone — only if the offer beats the cheapest free offer any lane has found so far (the bound is
// refreshed once per pass; an offer that merely ties it is tested: it may have the lower index).  The colliding offers
// that are cheaper than the winner have to be tested by any scheme; here 32 of them are tested at a time instead of one
// after the other with 3 disc tests per lane (C4: 2.35 -> 2.1x ms per 4096 plans).  Kept SMALL (one steer site, not
// inlined): the other warps of the CTA are in the planner's main loop and share the instruction cache with this code.
template <typename Tree>
__device__ __noinline__ int warp_goal_connect(const Tree T, int n, const State4 goal, double r, const EnvView E,
                                              int* goal_parent) {
  const int lane = threadIdx.x & 31;
  double best_c = INFINITY, best_tg = KR(0.0), best_cg = KR(0.0);  // this lane's cheapest free offer
  int best_i = -1;
  bool best_same = false;
  double bound = INFINITY;  // cheapest free offer of the whole warp as of the last pass
#pragma unroll 1
  for (int first = 1; first < n; first += 32) {
    const int i = first + lane;
    if (i < n) {
      const State4 x = goal_load_state(T, i);
      const bool same = dist2(x, goal) == KR(0.0);  // a node that coincides with the goal offers a zero-length edge
      double tg = KR(0.0), cg = KR(0.0);
      if (!same) steer(x, goal, r, tg, cg);
      const double c = T.ctc[i] + cg;
      if (c < best_c && !(c > bound)) {  // (strict on the lane's own best: its earlier offers have lower indices)
        bool hit = false;
        if (!same && E.m > 0) {
          const Edge e = make_edge(x, goal, tg, T.tim[i]);
#pragma unroll 1
          for (int k = 0; k < E.S && !hit; ++k) {
            const double h = (E.S > 1) ? (double)k / (double)(E.S - 1) : KR(0.0);
            hit = edge_point_hits_any(e, h, E.obs, E.m, E.dynamic, E.R2);
          }
        }
        if (!hit) {
          best_c = c;
          best_i = i;
          best_tg = tg;
          best_cg = cg;
          best_same = same;
        }
      }
    }
    double wb = best_c;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) wb = fmin(wb, __shfl_xor_sync(FULL, wb, off));
    bound = wb;
  }
  double kc = best_c;
  int ki = best_i, kl = lane;
  warp_argmin(kc, ki, kl);
  *goal_parent = ki;
  if (ki < 0) return -1;
  if (__shfl_sync(FULL, best_same ? 1 : 0, kl)) return ki;  // ... and IS the goal node when it wins
  const double w_tg = __shfl_sync(FULL, best_tg, kl), w_cg = __shfl_sync(FULL, best_cg, kl);
  if (lane == 0) {
    T.pos[n] = make_double2(goal.x, goal.y);
    T.vel[n] = make_double2(goal.vx, goal.vy);
    T.ecost[n] = w_cg;
    T.etime[n] = w_tg;
    T.ctc[n] = kc;
    T.tim[n] = T.tim[ki] + w_tg;
  }
  __syncwarp();
  return n;
}

// every lane's stores of the record are ordered before the counter (fence by each writing lane, then one atomic).
// System scope: the consumer that waits on the counter ships the records with an NCCL gather, whose readers can be
// copy engines and peer GPUs (symmetric windows), not only this device's SMs.  Once per query.
__device__ __forceinline__ void batch_query_done(const BatchJob& J, int qi) {
  if (!J.chunk_done) return;
  __threadfence_system();
  __syncwarp();
  if ((threadIdx.x & 31) == 0) atomicAdd(&J.chunk_done[qi / J.chunk_queries], 1u);
}

__host__ __device__ inline size_t batch_warp_smem_bytes(int cap, int m) {
  cap += 1;  // the goal node's slot
  size_t b = 0;
  b += (size_t)cap * sizeof(double2) * 2;           // pos, vel
  b += (size_t)(m > 0 ? m : 1) * sizeof(double4);   // obstacles
  b += (size_t)cap * sizeof(double) * 8;            // ctc, tim, ecost, etime, cf, tf, cb, tb
  b += (size_t)cap * 4 * 6;                         // link (first child, next sibling), parent, prev, queue, near idx
  b += ((size_t)cap + 15) / 16 * 16;                // tried
  return (b + 15) / 16 * 16;
}

__global__ void __launch_bounds__(256) k_plan_batch(BatchJob J) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int lane = lane_id();
  const int warp = threadIdx.x >> 5;
  const int cap = J.cap + 1;  // array stride: the tree's cap nodes + the goal node's slot
  unsigned char* base = smem + (size_t)warp * batch_warp_smem_bytes(J.cap, J.m);
  TreeView T;
  NearView N;
  T.pos = reinterpret_cast<double2*>(base);
  T.vel = T.pos + cap;
  double4* s_obs = reinterpret_cast<double4*>(T.vel + cap);
  T.ctc = reinterpret_cast<double*>(s_obs + (J.m > 0 ? J.m : 1));
  T.tim = T.ctc + cap;
  T.ecost = T.tim + cap;
  T.etime = T.ecost + cap;
  N.cf = T.etime + cap;
  N.tf = N.cf + cap;
  N.cb = N.tf + cap;
  N.tb = N.cb + cap;
  T.link = reinterpret_cast<int2*>(N.tb + cap);
  T.parent = reinterpret_cast<int*>(T.link + cap);
  T.prev = T.parent + cap;
  T.queue = T.prev + cap;
  N.idx = T.queue + cap;
  N.tried = reinterpret_cast<unsigned char*>(N.idx + cap);
  T.cap = J.cap;
  EnvView E = {s_obs, J.m, J.dynamic, J.R2, J.S};
  Stats st = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};

  const int n_work = J.list ? *J.list_count : J.q;
  for (;;) {
    int qi = 0;
    if (lane == 0) qi = atomicAdd(J.next_query, 1);
    qi = __shfl_sync(FULL, qi, 0);
    if (qi >= n_work) break;
    if (J.list) qi = J.list[qi];
    const uint32_t query_id = J.query_base + (uint32_t)qi;
    const krrt_io_real* gs = J.goals + 4 * (size_t)qi;
    const krrt_io_real* xs = J.starts + 4 * (size_t)qi;
    const State4 goal = {(double)gs[0], (double)gs[1], (double)gs[2], (double)gs[3]};
    for (int j = lane; j < J.m; j += 32) {
      const krrt_io_real* o = J.obstacles + ((size_t)qi * J.m + j) * 4;
      s_obs[j] = make_double4((double)o[0], (double)o[1], (double)o[2], (double)o[3]);
    }
    if (lane == 0) {
      T.pos[0] = make_double2((double)xs[0], (double)xs[1]);
      T.vel[0] = make_double2((double)xs[2], (double)xs[3]);
      T.parent[0] = -1;
      T.ctc[0] = KR(0.0); T.tim[0] = KR(0.0); T.ecost[0] = KR(0.0); T.etime[0] = KR(0.0);
      T.link[0] = make_int2(-1, -1); T.prev[0] = -1;
    }
    __syncwarp();
    int n = 1;
    for (int it = 0; it < J.n_iter; ++it) {
      const State4 s = draw_sample(J.sp, query_id, (uint64_t)it, true, goal);
      st.iters++;
      if (!J.dynamic && warp_point_in_obstacle(s, E)) {
        st.fail_reject++;
        continue;
      }
      if (n >= J.cap) break;
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
          double d = dist2(load_state(T, i), s);
          if (bi < 0 || d < bd) {
            bd = d;
            bi = i;
          }
          valid = d <= rho2n;
          dup |= d == KR(0.0);
        }
        unsigned mk = __ballot_sync(FULL, valid);
        if (valid) N.idx[K + __popc(mk & ((1u << lane) - 1))] = i;
        K += __popc(mk);
      }
      if (__any_sync(FULL, dup)) {
        st.fail_dup++;
        continue;
      }
      __syncwarp();
      if (K == 0) {  // empty near set: the nearest node is the only candidate
        int aux = 0;
        warp_argmin(bd, bi, aux);
        if (lane == 0) N.idx[0] = bi;
        K = 1;
        __syncwarp();
      }
      // both steering directions of every near node: 2K independent tasks over the lanes, ONE call site so the
      // lanes of a warp iterate the Newton loop together
      for (int t = lane; t < 2 * K; t += 32) {
        const bool fwd = t < K;
        const int k = fwd ? t : t - K;
        const State4 x = load_state(T, N.idx[k]);
        double tau, c;
        steer(fwd ? x : s, fwd ? s : x, J.r, tau, c);
        if (fwd) {
          N.tf[k] = tau; N.cf[k] = c;
        } else {
          N.tb[k] = tau; N.cb[k] = c;
        }
      }
      __syncwarp();
      st.sum_near += K;
      int nw = warp_commit(T, N, K, n, s, E, st);
      if (nw >= 0) n = nw + 1;
      __syncwarp();
    }
    // result record
    int gi;
    if (J.goal_mode == GOAL_MODE_TOLERANCE) {
      gi = warp_find_goal(T, n, goal, J.goal_tol2);
    } else {
      int gpar = -1;
      gi = warp_goal_connect(T, n, goal, J.r, E, &gpar);
      if (gi == n && lane == 0) T.parent[n] = gpar;
      __syncwarp();
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
      for (int i = gi; i >= 0 && len <= n; i = T.parent[i]) ++len;
      R->path_len = len;
      R->cost = gi >= 0 ? T.ctc[gi] : KR(0.0);
      R->time = gi >= 0 ? T.tim[gi] : KR(0.0);
      int p = len - 1;
      for (int i = gi; i >= 0 && p >= 0; i = T.parent[i], --p) {
        if (p < 32) {
          double2 a = T.pos[i], v = T.vel[i];
          R->waypoints[p][0] = a.x; R->waypoints[p][1] = a.y;
          R->waypoints[p][2] = v.x; R->waypoints[p][3] = v.y;
          R->waypoints[p][4] = T.tim[i];
        }
      }
    }
    __syncwarp();
    batch_query_done(J, qi);
  }
  if (lane == 0 && J.stats) {
    const long long* sv = &st.iters;
    for (int k = 0; k < 10; ++k) atomicAdd(reinterpret_cast<unsigned long long*>(J.stats + k), (unsigned long long)sv[k]);
  }
}

}  // namespace krrt
