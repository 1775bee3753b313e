// krrt_kernels.cuh — sm_100a kernels of the kinodynamic RRT* hot path.
//
// Single big tree (configs C1-C3): samples are counter-based (Philox), so the tree-independent work of a whole
// window of future expansions is evaluated in parallel against the tree as it stands ("frozen"); validated prefixes of
// the window are then committed in parallel or, while the tree is small, the window is committed strictly in order by
// one warp (krrt_grow.cuh: one persistent kernel per krrt_grow call).  Every arithmetic operation a sequential planner
// would do is done with the same operands; only the point in time differs, so trees are bit-identical with a
// sequential run.
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
  // round control (all decided on the device: k_grow runs until iter == iter_target)
  long long iter_begin;   // iteration counter when the current krrt_grow call began (index base of user samples)
  long long iter_target;  // iteration counter at which the current krrt_grow call is complete
  unsigned round;         // sub-round id: epoch of the conflict marks (krrt_grow.cuh)
  int wave;               // this round commits a validated prefix in parallel (tree large enough)
  int wave_F;             // length of that prefix
  int ema16;              // 16 x running mean of the prefix length: sizes the next round's batch
  long long rounds_wave, rounds_seq, wave_committed;  // diagnostics since set_start
  long long tail_commits;  // of rounds_seq: single samples committed in order right behind a parallel prefix (no sub-round of their own)
  long long why[5];  // what ended the prefix: batch end, in-batch neighbour, needs sequential path, no free parent, read/write conflict
  int wave_Fs[2];         // prefix length of the current sub-round (slot = round & 1; the other slot is being re-armed)
  long long windows;      // windows scanned since the current krrt_grow call began
  long long phase_clk[8]; // SM clocks the last CTA spent per phase since the call began: 0 scan, 1 steer, 2 decide,
                          // 3 validate, 4 apply + finish, 5 sequential windows
  // spatial index of the neighbour scan (krrt_grow.cuh: uniform grid over the position plane, rebuilt as the tree grows)
  int grid_n;             // nodes [0, grid_n) are in the sorted copy; the rest of the tree is scanned directly (0: no index)
  int grid_w, grid_h;     // cells per axis
  int grid_pad;
  double grid_cs;         // the near radius the cell size was chosen for (cells are at least 1.001 x that wide)
  double grid_lox, grid_loy, grid_invx, grid_invy;  // cell = clamp(floor((p - lo) * inv), 0, cells - 1)
  double scan_visited;    // (node, sample) pairs the scan really evaluated since the current krrt_grow call began
  long long grid_rebuilds;
  long long dbg_clk[16];
  long long dbg_cta[3][160];  // per CTA index: decide / validate / apply(+finish) cycles, sub-rounds it was active in  // KRRT_DEBUG_CLK builds: step clocks inside decide / apply of CTA 0, max over CTAs
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
  int goal_mode;       // KRRT_GOAL_CONNECT / KRRT_GOAL_TOLERANCE (include/krrt.h)
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

// Per-iteration record of a logged call (krrt_grow_logged: the caller's own `for(i<iteration) rrt.grow(xg)` loop,
// robosoccer_motion_planning_node.cpp:176-185, served from ONE device call).  The in-order commit leaves, per iteration, the
// tree size, the sample, and — whenever the tree changed — a snapshot of every node's cost-to-come and arrival time
// (row 0: the tree before the call); k_goal_log then works out what grow(xg) would have returned after each iteration.
struct IterLog {
  int* n_after;      // [rows - 1] tree size after iteration i of the call (nullptr: logging off)
  int* snap_of;      // [rows - 1] snapshot row that holds the tree's costs after iteration i
  double* sample;    // [rows - 1][4]
  double* snap_ctc;  // [rows][stride]
  double* snap_tim;
  int stride;        // nodes a snapshot row holds
  int rows;          // 1 + iterations the buffers hold
};
struct IterRec {     // layout == krrt_iter_record (include/krrt.h)
  int n_nodes, solved, goal_index, inserted;
  double sample[4];
  double goal_cost;
};

// ---------------------------------------------------------------------------------------------------
// Per-sample decision of a sub-round (krrt_grow.cuh) and the conflict marks it is validated with.
enum { OUT_REJECT = 0, OUT_DUP = 1, OUT_NOFREE = 2, OUT_INSERT = 3, OUT_COMPLEX = 4, OUT_CONFLICT = 5,
       OUT_NEAREST = 6 };  // empty near set: the in-order path with a grid-wide nearest-node scan in front

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
  unsigned* wmin;   // [capacity] mark: lowest sample of this sub-round that WRITES the node's cost/time
  unsigned* amin;   // [capacity] mark: lowest sample of this sub-round that chose the node as PARENT (derived its cost from it)
  unsigned long long* rmin;  // [capacity] (mark << 32 | where): lowest sample of this sub-round that changes the node's cost, and
                             // where the new cost is on record: where < 256: the node is REWIRED, its slot in that sample's
                             // candidate list (c*(sample -> node)); else 256 + its place in that sample's row of wlc
  unsigned* done;   // [1024] sub-round tag of the last decision CTA c finished (flow)
  unsigned* ok;     // [1024] (tag << 1 | failed) of the last validation CTA c finished (flow)
  double* wlc;      // [gridDim][GROW_QCAP] new cost-to-come of the nodes below the nodes CTA c's sample rewires
  int* lock;        // [capacity] child-list lock of a parent
  double* nn_d2;    // [gridDim] per-CTA partial of a grid-wide nearest-node scan ...
  int* nn_idx;      // ... (squared distance, node index)
  int* dbg_log;     // KRRT_DEBUG_CLK builds: [0] = entries, then 8 ints per decided sample (else nullptr)
};

// Uniform grid over the position plane for the neighbour scan of a big tree: a copy of the node states sorted by cell
// (cells of at least the near radius, so a sample's neighbours lie in its 3 x 3 block = three contiguous cell ranges).
constexpr int GRID_W_MAX = 96, GRID_H_MAX = 64;
constexpr int GRID_CELLS_MAX = GRID_W_MAX * GRID_H_MAX;
struct GridIndex {
  double2* pos;   // [capacity] sorted copy of T.pos
  double2* vel;   // [capacity] ... of T.vel
  int* id;        // [capacity] node index of each sorted slot
  int* off;       // [capacity] rebuild scratch: rank of node i inside its cell
  int* cell;      // [GRID_CELLS_MAX + 1] rebuild: counts; afterwards exclusive starts, cell[cells] = grid_n
  int enabled;    // KRRT_GRID_INDEX=0 in the environment: every window scans the whole tree
  int min_n;      // smallest tree that gets an index
  int tail_max;   // most nodes left outside the index (scanned directly) before it is rebuilt: min(n / 16, tail_max)
};

// marks carry the round id, so they never need clearing: (round << 9) | (511 - sample), combined with atomicMax
__device__ __forceinline__ unsigned mark_key(unsigned round, int b) { return (round << 9) | (unsigned)(511 - b); }
__device__ __forceinline__ int mark_sample(unsigned m, unsigned round) {
  return (m >> 9) == round ? 511 - (int)(m & 511u) : 0x7fffffff;
}

// shared-memory scratch of the in-order commit warp (krrt_grow.cuh: seq_commit): obstacles + near list
__host__ __device__ inline size_t commit_smem_bytes(int near_cap) {
  return sizeof(double4) * KRRT_MAX_OBS_DEV + (size_t)near_cap * (4 * 8 + 4 + 1) + 16;
}

// goal lookup after a grow call, goal_mode KRRT_GOAL_TOLERANCE: cheapest node within goal_tol of the goal
// (rrt.goalIndex(), robosoccer_visual.hpp:19)
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

// Goal connection after a grow call, goal_mode KRRT_GOAL_CONNECT (default; the `solved` of rrt.grow(xg),
// robosoccer_motion_planning_node.cpp:177, and rrt.goalIndex(), robosoccer_visual.hpp:19,40-45).  Every node the planner
// inserted (i >= 1) offers the edge i -> goal: optimal steering, cost ctc[i] + c*(i -> goal), collision-tested as any
// other edge, departing at the node's arrival time.  The goal's parent is the cheapest collision-free offer on the tree
// as it stands (lowest index among equal costs); a node that coincides with the goal offers a zero-length edge and, if it
// wins, IS the goal node.  Otherwise the goal node is written right behind the last tree node (slot n: the arrays hold
// capacity + 1 entries): path / edge / trajectory extraction take its index like any other, the next grow call does not
// see it (no scan reaches slot n, it is in no child list) and overwrites it.
// One thread per node; a thread skips the collision test of an offer that cannot beat the best free one it has seen.
constexpr int GOAL_THREADS = 256;
constexpr int GOAL_MAX_CTAS = 1024;
struct GoalScratch {
  double* part_c;    // [GOAL_MAX_CTAS] per-CTA best offer ...
  int* part_i;       // ... and the node that makes it (-1: none)
  unsigned* ticket;  // CTAs done (the last one folds the partials; re-armed for the next launch)
};

__global__ void __launch_bounds__(GOAL_THREADS) k_goal_connect(Ctl* ctl, Params P, TreeView T, GoalScratch Gs) {
  __shared__ double4 s_obs[KRRT_MAX_OBS_DEV];
  __shared__ double s_c[GOAL_THREADS / 32];
  __shared__ int s_i[GOAL_THREADS / 32];
  __shared__ int s_last;
  const int t = threadIdx.x;
  if (!ctl->has_goal) {
    if (t == 0 && blockIdx.x == 0) ctl->goal_index = -1;
    return;
  }
  const int n = ctl->n;
  const State4 g = {ctl->goal[0], ctl->goal[1], ctl->goal[2], ctl->goal[3]};
  for (int j = t; j < P.m; j += GOAL_THREADS) s_obs[j] = P.obs[j];
  __syncthreads();
  double bc = INFINITY;
  int bi = -1, aux = 0;
  for (int i = 1 + blockIdx.x * GOAL_THREADS + t; i < n; i += gridDim.x * GOAL_THREADS) {
    const State4 x = load_state(T, i);
    const bool same = dist2(x, g) == KR(0.0);
    double tg = KR(0.0), cg = KR(0.0);
    if (!same) steer(x, g, P.r, tg, cg);
    const double cand = T.ctc[i] + cg;
    if (bi >= 0 && cand >= bc) continue;  // (i only grows: an equal offer loses on the index)
    bool hit = false;
    if (!same && P.m > 0) {
      const Edge e = make_edge(x, g, tg, T.tim[i]);
      for (int k = 0; k < P.S && !hit; ++k) {
        const double h = (P.S > 1) ? (double)k / (double)(P.S - 1) : KR(0.0);
        hit = edge_point_hits_any(e, h, s_obs, P.m, P.dynamic, P.R2);
      }
    }
    if (!hit) {
      bc = cand;
      bi = i;
    }
  }
  warp_argmin(bc, bi, aux);
  if ((t & 31) == 0) {
    s_c[t >> 5] = bc;
    s_i[t >> 5] = bi;
  }
  __syncthreads();
  if (t < 32) {
    bc = t < GOAL_THREADS / 32 ? s_c[t] : INFINITY;
    bi = t < GOAL_THREADS / 32 ? s_i[t] : -1;
    warp_argmin(bc, bi, aux);
    if (t == 0) {
      Gs.part_c[blockIdx.x] = bc;
      Gs.part_i[blockIdx.x] = bi;
      __threadfence();
      s_last = atomicAdd(Gs.ticket, 1u) == gridDim.x - 1;
    }
  }
  __syncthreads();
  if (!s_last || t >= 32) return;
  __threadfence();
  bc = INFINITY;
  bi = -1;
  for (int k = t; k < (int)gridDim.x; k += 32) {
    const double c = __ldcg(&Gs.part_c[k]);
    const int ci = __ldcg(&Gs.part_i[k]);
    if (ci >= 0 && (bi < 0 || c < bc || (c == bc && ci < bi))) {
      bc = c;
      bi = ci;
    }
  }
  warp_argmin(bc, bi, aux);
  if (t == 0) {
    *Gs.ticket = 0u;
    int gi = -1;
    if (bi >= 0) {
      const State4 x = load_state(T, bi);
      if (dist2(x, g) == KR(0.0)) {
        gi = bi;
      } else {
        double tg, cg;
        steer(x, g, P.r, tg, cg);
        gi = n;
        T.pos[n] = make_double2(g.x, g.y);
        T.vel[n] = make_double2(g.vx, g.vy);
        T.parent[n] = bi;
        T.ecost[n] = cg;
        T.etime[n] = tg;
        T.ctc[n] = T.ctc[bi] + cg;
        T.tim[n] = T.tim[bi] + tg;
      }
    }
    ctl->goal_index = gi;
  }
}

// grow(xg)'s answer after EVERY iteration of a logged call: block i redoes the goal connection (k_goal_connect's rule) on
// the tree as it stood after iteration i — node states never change, costs and arrival times come from the snapshot row.
constexpr int GLOG_THREADS = 128;
__global__ void __launch_bounds__(GLOG_THREADS) k_goal_log(const Ctl* ctl, Params P, TreeView T, IterLog L, int n_iter,
                                                           int n_before, IterRec* out) {
  __shared__ double4 s_obs[KRRT_MAX_OBS_DEV];
  __shared__ double s_c[GLOG_THREADS / 32];
  __shared__ int s_i[GLOG_THREADS / 32];
  const int t = threadIdx.x, it = blockIdx.x;
  if (it >= n_iter) return;
  // the call ends at the iteration that finds the tree full, and the log ends with it: an iteration behind that one
  // repeats its record (include/krrt.h) instead of reading log entries nobody wrote
  const long long ran = ctl->iter - ctl->iter_begin;
  const int done = (int)(ran < 0 ? 0 : (ran > n_iter ? n_iter : ran));
  if (done == 0) {  // (a call that stopped with an error before its first iteration: the tree as it was)
    if (t == 0) {
      IterRec r;
      r.n_nodes = n_before; r.solved = 0; r.goal_index = -1; r.inserted = 0; r.goal_cost = KR(0.0);
      for (int k = 0; k < 4; ++k) r.sample[k] = KR(0.0);
      out[it] = r;
    }
    return;
  }
  const int src = it < done ? it : done - 1;
  const int n = L.n_after[src];
  const int row = L.snap_of[src];
  const int n_prev = it >= done ? n : (it > 0 ? L.n_after[it - 1] : n_before);
  const double* ctc = L.snap_ctc + (size_t)row * L.stride;
  const double* tim = L.snap_tim + (size_t)row * L.stride;
  for (int j = t; j < P.m; j += GLOG_THREADS) s_obs[j] = P.obs[j];
  __syncthreads();
  double bc = INFINITY;
  int bi = -1, aux = 0;
  const bool has_goal = ctl->has_goal != 0;
  const State4 g = {ctl->goal[0], ctl->goal[1], ctl->goal[2], ctl->goal[3]};
  if (has_goal && P.goal_mode == GOAL_MODE_TOLERANCE) {
    for (int i = t; i < n; i += GLOG_THREADS)
      if (dist2(load_state(T, i), g) <= P.goal_tol2) {
        const double c = ctc[i];
        if (c < bc) {
          bc = c;
          bi = i;
        }
      }
  } else if (has_goal) {
    for (int i = 1 + t; i < n; i += GLOG_THREADS) {
      const State4 x = load_state(T, i);
      const bool same = dist2(x, g) == KR(0.0);
      double tg = KR(0.0), cg = KR(0.0);
      if (!same) steer(x, g, P.r, tg, cg);
      const double cand = ctc[i] + cg;
      if (bi >= 0 && cand >= bc) continue;
      bool hit = false;
      if (!same && P.m > 0) {
        const Edge e = make_edge(x, g, tg, tim[i]);
        for (int k = 0; k < P.S && !hit; ++k) {
          const double h = (P.S > 1) ? (double)k / (double)(P.S - 1) : KR(0.0);
          hit = edge_point_hits_any(e, h, s_obs, P.m, P.dynamic, P.R2);
        }
      }
      if (!hit) {
        bc = cand;
        bi = i;
      }
    }
  }
  warp_argmin(bc, bi, aux);
  if ((t & 31) == 0) {
    s_c[t >> 5] = bc;
    s_i[t >> 5] = bi;
  }
  __syncthreads();
  if (t < 32) {
    bc = t < GLOG_THREADS / 32 ? s_c[t] : INFINITY;
    bi = t < GLOG_THREADS / 32 ? s_i[t] : -1;
    warp_argmin(bc, bi, aux);
    if (t == 0) {
      IterRec r;
      r.n_nodes = n;
      r.inserted = n > n_prev ? 1 : 0;
      r.solved = bi >= 0 ? 1 : 0;
      r.goal_index = -1;
      r.goal_cost = KR(0.0);
      if (bi >= 0) {
        const bool node_is_goal = P.goal_mode == GOAL_MODE_TOLERANCE || dist2(load_state(T, bi), g) == KR(0.0);
        r.goal_index = node_is_goal ? bi : n;
        r.goal_cost = bc;
      }
      for (int k = 0; k < 4; ++k) r.sample[k] = L.sample[4 * (size_t)src + k];
      out[it] = r;
    }
  }
}

// radius + nearest query of one state against the tree; ascending indices (single CTA, ordered compaction)
__global__ void __launch_bounds__(1024) k_near_query(TreeView T, int n, State4 x, double radius2, int* out_idx,
                                                      int cap, int* out_count, int* out_nearest,
                                                      double* out_nearest_d2) {
  __shared__ int s_cnt[32];
  __shared__ double s_d[32];
  __shared__ int s_i[32];
  __shared__ int s_total;
  const int lane = lane_id(), warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_total = 0;
  __syncthreads();
  double bd = INFINITY;
  int bi = -1, aux = 0;
  for (int base = 0; base < n; base += 1024) {
    int i = base + threadIdx.x;
    bool valid = false;
    if (i < n) {
      double d = dist2(load_state(T, i), x);
      if (bi < 0 || d < bd) {
        bd = d;
        bi = i;
      }
      valid = d <= radius2;
    }
    unsigned mk = __ballot_sync(FULL, valid);
    if (lane == 0) s_cnt[warp] = __popc(mk);
    __syncthreads();
    int off = s_total;
    for (int w = 0; w < warp; ++w) off += s_cnt[w];
    if (valid) {
      int k = off + __popc(mk & ((1u << lane) - 1));
      if (k < cap) out_idx[k] = i;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      int tot = 0;
      for (int w = 0; w < 32; ++w) tot += s_cnt[w];
      s_total += tot;
    }
    __syncthreads();
  }
  warp_argmin(bd, bi, aux);
  if (lane == 0) {
    s_d[warp] = bd;
    s_i[warp] = bi;
  }
  __syncthreads();
  if (threadIdx.x < 32) {
    bd = s_d[threadIdx.x];
    bi = s_i[threadIdx.x];
    warp_argmin(bd, bi, aux);
    if (threadIdx.x == 0) {
      *out_count = s_total;
      *out_nearest = bi;
      *out_nearest_d2 = bd;
    }
  }
}

}  // namespace krrt
