// krrt_warp.cuh — warp-cooperative RRT* commit: parent choice, insert, rewire, cost/time propagation,
// time-space edge collision.  One warp owns one tree while it commits; the same code serves the
// single big tree (arrays in HBM/L2, near list pre-evaluated by the scan/steer kernels) and the
// batched small trees (arrays in shared memory).  Pointers are generic so either placement works.
//
// Stands in for the inside of RRTStar::grow (robosoccer_motion_planning_node.cpp:177): the Tree slot's
// parent/insert, the rewire step, and the Environment slot's collision predicate
// (robosoccer_visual.hpp:16-18,35).  Sequential semantics are preserved exactly: a warp handles one
// expansion at a time, candidates are ordered by node index, ties go to the lowest index.
#pragma once
#include "krrt_math.cuh"

namespace krrt {

constexpr unsigned FULL = 0xffffffffu;
constexpr int KRRT_MAX_OBS_DEV = 64;  // == KRRT_MAX_OBS in include/krrt.h

struct TreeView {
  double2* pos;  // (x, y)
  double2* vel;  // (vx, vy)
  double* ctc;   // cost-to-come
  double* tim;   // arrival time since root
  double* ecost; // cost of the edge parent -> node
  double* etime; // duration of that edge
  int* parent;
  int2* link;    // child lists: .x = first child, .y = next sibling (one 8-byte load walks a sibling AND finds its children)
  int* prev;
  int* queue;    // scratch for subtree propagation (capacity ints)
  int cap;       // node capacity (bounds every list walk: a corrupted list can never hang the kernel)
};

struct NearView {
  int* idx;
  double* cf;  // c*(near -> new)
  double* tf;  // tau*(near -> new)
  double* cb;  // c*(new -> near)
  double* tb;  // tau*(new -> near)
  unsigned char* tried;
};

struct EnvView {
  const double4* obs;  // (x, y, vx, vy)
  int m;
  int dynamic;
  double R2;
  int S;
};

struct Stats {
  long long iters, inserted, fail_reject, fail_dup, fail_nofree;
  long long rewired, propagated, sum_near, steer_calls, collide_calls;
};

__device__ __forceinline__ State4 load_state(const TreeView& T, int i) {
  double2 p = T.pos[i], v = T.vel[i];
  State4 s = {p.x, p.y, v.x, v.y};
  return s;
}

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

// S x m disc tests of one edge spread over the 32 lanes; uniform result
__device__ __forceinline__ bool warp_edge_collide(const State4& a, const State4& b, double tau, double t0,
                                                  const EnvView& E) {
  Edge e = make_edge(a, b, tau, t0);
  bool hit = false;
  int total = E.S * E.m;
  for (int p = lane_id(); p < total; p += 32) {
    int k = p / E.m, j = p - k * E.m;
    hit |= edge_point_hits(e, k, E.S, E.obs[j], E.dynamic, E.R2);
  }
  return __any_sync(FULL, hit);
}

__device__ __forceinline__ bool warp_point_in_obstacle(const State4& s, const EnvView& E) {
  bool hit = false;
  for (int j = lane_id(); j < E.m; j += 32) {
    double dx = s.x - E.obs[j].x, dy = s.y - E.obs[j].y;
    hit |= fma(dy, dy, dx * dx) < E.R2;
  }
  return __any_sync(FULL, hit);
}

// (key, idx) lexicographic arg-min over the warp; idx < 0 means "no candidate on this lane"
__device__ __forceinline__ void warp_argmin(double& key, int& idx, int& aux) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    double ok = __shfl_xor_sync(FULL, key, off);
    int oi = __shfl_xor_sync(FULL, idx, off);
    int oa = __shfl_xor_sync(FULL, aux, off);
    if (oi >= 0 && (idx < 0 || ok < key || (ok == key && oi < idx))) {
      key = ok;
      idx = oi;
      aux = oa;
    }
  }
}
__device__ __forceinline__ void warp_min_int(int& key, int& aux) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    int ok = __shfl_xor_sync(FULL, key, off);
    int oa = __shfl_xor_sync(FULL, aux, off);
    if (ok < key) {
      key = ok;
      aux = oa;
    }
  }
}

// child-list maintenance (one lane)
__device__ __forceinline__ void link_child(const TreeView& T, int par, int c) {
  int h = T.link[par].x;
  T.link[c].y = h;
  T.prev[c] = -1;
  if (h >= 0) T.prev[h] = c;
  T.link[par].x = c;
}
__device__ __forceinline__ void unlink_child(const TreeView& T, int c) {
  int par = T.parent[c], nx = T.link[c].y, pv = T.prev[c];
  if (pv >= 0) T.link[pv].y = nx; else T.link[par].x = nx;
  if (nx >= 0) T.prev[nx] = pv;
}

// recompute cost-to-come and arrival time of every descendant of j from its parent: level-order walk,
// 32 nodes per step.  A node is always queued after its parent has been updated.
__device__ __forceinline__ long long warp_propagate(const TreeView& T, int j) {
  const int lane = lane_id();
  int qhead = 0, qtail = 0;
  {
    int c = (lane == 0) ? T.link[j].x : -1;
    if (lane == 0) {
      while (c >= 0 && qtail < T.cap) {
        T.queue[qtail++] = c;
        c = T.link[c].y;
      }
    }
    qtail = __shfl_sync(FULL, qtail, 0);
  }
  __syncwarp();
  while (qhead < qtail) {
    int cnt = min(32, qtail - qhead);
    int c = -1;
    if (lane < cnt) {
      int d = T.queue[qhead + lane];
      int par = T.parent[d];
      T.ctc[d] = T.ctc[par] + T.ecost[d];
      T.tim[d] = T.tim[par] + T.etime[d];
      c = T.link[d].x;
    }
    qhead += cnt;
    for (;;) {
      unsigned m = __ballot_sync(FULL, c >= 0);
      // (a node is queued once, so a sound tree never fills the queue; the bound only stops a corrupted list.  It must be
      //  the exact one: "qtail + 32" would drop the grandchildren of every rewired node in a tree of fewer than 64 slots)
      if (m == 0 || qtail + __popc(m) > T.cap) break;
      if (c >= 0) {
        T.queue[qtail + __popc(m & ((1u << lane) - 1))] = c;
        c = T.link[c].y;
      }
      qtail += __popc(m);
    }
    __syncwarp();
  }
  return qtail;
}

// One expansion with a ready near list N[0..K) (K >= 1, every entry within the radius, cf/tf filled for all,
// cb/tb filled for all but a lone nearest-fallback entry).  Returns the new node's index (== n) or -1.
__device__ __forceinline__ int warp_commit(const TreeView& T, const NearView& N, int K, int n, const State4& s,
                                           const EnvView& E, Stats& st) {
  const int lane = lane_id();
  for (int k = lane; k < K; k += 32) N.tried[k] = 0;
  __syncwarp();
  st.steer_calls += K;

  // parent = arg-min over collision-free candidates of ctc[i] + c*(i -> new); lowest index on ties
  int par = -1, park = -1;
  for (;;) {
    double bc = INFINITY;
    int bi = -1, bk = -1;
    for (int k = lane; k < K; k += 32) {
      if (N.tried[k]) continue;
      int i = N.idx[k];
      double cand = T.ctc[i] + N.cf[k];
      if (cand < bc || (bi >= 0 && cand == bc && i < bi)) {
        bc = cand;
        bi = i;
        bk = k;
      }
    }
    warp_argmin(bc, bi, bk);
    if (bi < 0) break;
    if (lane == 0) N.tried[bk] = 1;
    __syncwarp();
    st.collide_calls++;
    State4 a = load_state(T, bi);
    if (!warp_edge_collide(a, s, N.tf[bk], T.tim[bi], E)) {
      par = bi;
      park = bk;
      break;
    }
  }
  if (par < 0) {
    st.fail_nofree++;
    return -1;
  }

  const int nw = n;
  const double cfp = N.cf[park], tfp = N.tf[park];
  const double ctc_new = T.ctc[par] + cfp;
  const double tim_new = T.tim[par] + tfp;
  if (lane == 0) {
    T.pos[nw] = make_double2(s.x, s.y);
    T.vel[nw] = make_double2(s.vx, s.vy);
    T.parent[nw] = par;
    T.ecost[nw] = cfp;
    T.etime[nw] = tfp;
    T.ctc[nw] = ctc_new;
    T.tim[nw] = tim_new;
    T.link[nw].x = -1;
    link_child(T, par, nw);
  }
  __syncwarp();
  st.inserted++;
  st.steer_calls += K - 1;

  // rewire in ascending node index; each decision sees the costs left by the previous one
  int last = -1;
  for (;;) {
    int bj = 0x7fffffff, bk = -1;
    for (int k = lane; k < K; k += 32) {
      int j = N.idx[k];
      if (j <= last || j == par || j >= bj) continue;
      double cand = ctc_new + N.cb[k];
      if (cand < T.ctc[j]) {
        bj = j;
        bk = k;
      }
    }
    warp_min_int(bj, bk);
    if (bk < 0) break;
    last = bj;
    st.collide_calls++;
    State4 xj = load_state(T, bj);
    double tb = N.tb[bk], cb = N.cb[bk];
    if (!warp_edge_collide(s, xj, tb, tim_new, E)) {
      if (lane == 0) {
        unlink_child(T, bj);
        T.parent[bj] = nw;
        link_child(T, nw, bj);
        T.ecost[bj] = cb;
        T.etime[bj] = tb;
        T.ctc[bj] = ctc_new + cb;
        T.tim[bj] = tim_new + tb;
      }
      __syncwarp();
      st.rewired++;
      st.propagated += warp_propagate(T, bj);
    }
  }
  return nw;
}

// exact nearest over nodes [0,n): lowest index among equal distances
__device__ __forceinline__ void warp_nearest(const TreeView& T, int n, const State4& s, int& best, double& best_d2) {
  double bd = INFINITY;
  int bi = -1, aux = 0;
  for (int i = lane_id(); i < n; i += 32) {
    double d = dist2(load_state(T, i), s);
    if (bi < 0 || d < bd) {
      bd = d;
      bi = i;
    }
  }
  warp_argmin(bd, bi, aux);
  best = bi;
  best_d2 = bd;
}

// goal node = cheapest node within goal_tol of the goal, lowest index on ties
__device__ __forceinline__ int warp_find_goal(const TreeView& T, int n, const State4& goal, double tol2) {
  double bc = INFINITY;
  int bi = -1, aux = 0;
  for (int i = lane_id(); i < n; i += 32) {
    if (dist2(load_state(T, i), goal) <= tol2) {
      double c = T.ctc[i];
      if (c < bc) {
        bc = c;
        bi = i;
      }
    }
  }
  warp_argmin(bc, bi, aux);
  return bi;
}

}  // namespace krrt
