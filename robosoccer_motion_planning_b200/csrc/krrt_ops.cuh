// krrt_ops.cuh — stateless batched operators behind krrt_steer_cost / krrt_edge_collide / krrt_near /
// krrt_solve / krrt_trajectory (the Cost, Environment, Tree-query and Connector slots called on their own:
// robosoccer_motion_planning_node.cpp:242-245, robosoccer_visual.hpp:35).
#pragma once
#include "krrt_warp.cuh"

namespace krrt {

__global__ void __launch_bounds__(128) k_steer_batch(const double* __restrict__ xi, const double* __restrict__ xf,
                                                       int n, double r, double* __restrict__ t,
                                                       double* __restrict__ c) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double4 a = reinterpret_cast<const double4*>(xi)[i];
  const double4 b = reinterpret_cast<const double4*>(xf)[i];
  State4 x0 = {a.x, a.y, a.z, a.w}, x1 = {b.x, b.y, b.z, b.w};
  double tau, cost;
  steer(x0, x1, r, tau, cost);
  t[i] = tau;
  c[i] = cost;
}

// one warp per edge: S x m disc tests over the lanes
__global__ void __launch_bounds__(128) k_edge_collide_batch(const double* __restrict__ xi,
                                                              const double* __restrict__ xf,
                                                              const double* __restrict__ tau,
                                                              const double* __restrict__ t0, int n,
                                                              const double4* __restrict__ obs, int m, int dynamic,
                                                              double R2, int S, unsigned char* __restrict__ hit) {
  __shared__ double4 s_obs[KRRT_MAX_OBS_DEV];
  for (int j = threadIdx.x; j < m; j += blockDim.x) s_obs[j] = obs[j];
  __syncthreads();
  int e = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (e >= n) return;
  const double4 a = reinterpret_cast<const double4*>(xi)[e];
  const double4 b = reinterpret_cast<const double4*>(xf)[e];
  State4 x0 = {a.x, a.y, a.z, a.w}, x1 = {b.x, b.y, b.z, b.w};
  EnvView E = {s_obs, m, dynamic, R2, S};
  bool h = warp_edge_collide(x0, x1, tau[e], t0[e], E);
  if (lane_id() == 0) hit[e] = h ? 1 : 0;
}

// (k_near_query, the radius + nearest query behind krrt_near, lives in krrt_kernels.cuh: both precisions use it)

// root -> idx chain (tree.states(idx), …node.cpp:237): states (len x 4) and arrival times, root first
__global__ void k_path_extract(TreeView T, int n, int idx, double* states, double* times, int cap, int* out_len) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  if (idx < 0 || idx >= n) {
    *out_len = 0;
    return;
  }
  int len = 0;
  for (int i = idx; i >= 0 && len < n; i = T.parent[i]) ++len;  // (n counts the goal node's slot when idx is the goal)
  int p = len - 1;
  for (int i = idx; i >= 0 && p >= 0; i = T.parent[i], --p) {
    if (p < cap) {
      double2 a = T.pos[i], v = T.vel[i];
      states[4 * p + 0] = a.x; states[4 * p + 1] = a.y; states[4 * p + 2] = v.x; states[4 * p + 3] = v.y;
      times[p] = T.tim[i];
    }
  }
  *out_len = len;
}

// rows of solver.solve(xi, xf, dt): t = 0, dt, 2dt, ... < tau, then tau.  last = index of the closing row.
__device__ __forceinline__ int solve_last_row(double tau, double dt) {
  int k = (int)floor(tau / dt);
  if (k < 0) k = 0;
  while ((double)k * dt < tau) ++k;
  while (k > 0 && !((double)(k - 1) * dt < tau)) --k;
  return k;
}

// per path edge e (states[e] -> states[e+1]): tau* and row count
__global__ void k_traj_count(const double* __restrict__ states, int n_edges, double r, double dt, double* tau,
                             int* rows) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n_edges) return;
  const double4 a = reinterpret_cast<const double4*>(states)[e];
  const double4 b = reinterpret_cast<const double4*>(states)[e + 1];
  State4 x0 = {a.x, a.y, a.z, a.w}, x1 = {b.x, b.y, b.z, b.w};
  double t, c;
  steer(x0, x1, r, t, c);
  tau[e] = t;
  rows[e] = solve_last_row(t, dt) + 1;
}

// exclusive prefix of rows (few edges: one thread), total in offs[n_edges]
__global__ void k_traj_offsets(const int* rows, int n_edges, int* offs) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  int acc = 0;
  for (int e = 0; e < n_edges; ++e) {
    offs[e] = acc;
    acc += rows[e];
  }
  offs[n_edges] = acc;
}

// one CTA per edge: rows (x, y, vx, vy, t0 + t)
__global__ void __launch_bounds__(128) k_traj_fill(const double* __restrict__ states, const double* __restrict__ t0s,
                                                     const double* __restrict__ tau, const int* __restrict__ offs,
                                                     double dt, double* out, int cap_rows,
                                                     double* __restrict__ inputs) {
  const int e = blockIdx.x;
  const double4 a = reinterpret_cast<const double4*>(states)[e];
  const double4 b = reinterpret_cast<const double4*>(states)[e + 1];
  State4 x0 = {a.x, a.y, a.z, a.w}, x1 = {b.x, b.y, b.z, b.w};
  const double T = tau[e], t0 = t0s[e];
  const int o = offs[e], cnt = offs[e + 1] - o;
  Edge ed = make_edge(x0, x1, T, t0);
  for (int k = threadIdx.x; k < cnt; k += blockDim.x) {
    if (o + k >= cap_rows) break;
    double t = (k == cnt - 1) ? T : (double)k * dt;
    double h = (T == 0.0) ? 0.0 : t / T;
    double* row = out + 5 * (size_t)(o + k);
    row[0] = axis_pos(ed.kx, h);
    row[1] = axis_pos(ed.ky, h);
    row[2] = axis_vel(ed.kx, h, x0.vx, T);
    row[3] = axis_vel(ed.ky, h, x0.vy, T);
    row[4] = t0 + t;
    if (inputs) {  // (ax, ay) at the same row
      inputs[2 * (size_t)(o + k) + 0] = axis_acc(ed.kx, h, T);
      inputs[2 * (size_t)(o + k) + 1] = axis_acc(ed.ky, h, T);
    }
  }
}

// Connector slot: S samples (x, y, vx, vy, t0 + tau h) at h = k / (S - 1) of the optimal edge xi -> xf, one thread
// per sample.  tau == nullptr: solve the steering here.  The same h grid as the collision predicate.
__device__ __forceinline__ void write_edge_sample(const State4& x0, const State4& x1, double tau, double t0, int k,
                                                  int S, double* row) {
  const double h = (S > 1) ? (double)k / (double)(S - 1) : 0.0;
  Edge e = make_edge(x0, x1, tau, t0);
  row[0] = axis_pos(e.kx, h);
  row[1] = axis_pos(e.ky, h);
  row[2] = axis_vel(e.kx, h, x0.vx, tau);
  row[3] = axis_vel(e.ky, h, x0.vy, tau);
  row[4] = t0 + tau * h;
}

__global__ void __launch_bounds__(128) k_edge_samples(const double* __restrict__ xi, const double* __restrict__ xf,
                                                        const double* __restrict__ t0, int n, int S, double r,
                                                        double* __restrict__ rows) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= (long long)n * S) return;
  const int e = (int)(g / S), k = (int)(g - (long long)e * S);
  const double4 a = reinterpret_cast<const double4*>(xi)[e];
  const double4 b = reinterpret_cast<const double4*>(xf)[e];
  State4 x0 = {a.x, a.y, a.z, a.w}, x1 = {b.x, b.y, b.z, b.w};
  double tau, c;
  steer(x0, x1, r, tau, c);
  write_edge_sample(x0, x1, tau, t0[e], k, S, rows + 5 * g);
}

// tree.trajectories: for every node i the stored edge parent[i] -> i (node 0: S copies of the root at t = 0)
__global__ void __launch_bounds__(128) k_tree_edges(TreeView T, int n, int S, double* __restrict__ rows) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= (long long)n * S) return;
  const int i = (int)(g / S), k = (int)(g - (long long)i * S);
  const int p = T.parent[i];
  const State4 x1 = load_state(T, i);
  if (p < 0) {
    double* row = rows + 5 * g;
    row[0] = x1.x; row[1] = x1.y; row[2] = x1.vx; row[3] = x1.vy; row[4] = 0.0;
    return;
  }
  write_edge_sample(load_state(T, p), x1, T.etime[i], T.tim[p], k, S, rows + 5 * g);
}

// Obstacle ingest: f frames of OminiVisionInfo robotinfo[] (float32, cm and cm/s:
// nubot_common/msgs/RobotInfo.msg:2,9-12) -> the planner's own state + obstacle rows in metres, as
// RobotSubscriber::callback (robotsubscriber.cpp:44-77: own team, x / 100.), ::rival_callback
// (robotsubscriber.cpp:29-42: rivals flipped, -x * 1e-2 — a different rounding than / 100., kept) and
// ::getObstacles (robotsubscriber.h:50-55: friends without self, then rivals) build them.
// One thread per (frame, output slot); slot n_team + n_rival of each frame writes the frame's scalars.
// HBM-bound: 24 n_team + 16 n_rival + 4 B in, 32 (n_team + n_rival) + 48 B out per frame.
struct IngestJob {
  const float* team;        // f x n_team x 5
  const int32_t* agent_id;  // f x n_team
  const float* rival;       // f x n_rival x 4
  const int32_t* self_id;   // f
  double* self_state;       // f x 4
  double* heading;          // f
  int32_t* self_index;      // f
  double* obstacles;        // f x (n_team + n_rival) x 4
  int32_t* n_obs;           // f
  int f, n_team, n_rival;
};

__global__ void __launch_bounds__(256) k_ingest_vision(IngestJob J) {
  const int slots = J.n_team + J.n_rival;
  const long long total = (long long)J.f * (slots + 1);
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < total;
       g += (long long)gridDim.x * blockDim.x) {
    const int q = (int)(g / (slots + 1)), s = (int)(g - (long long)q * (slots + 1));
    const int32_t* ids = J.agent_id + (size_t)q * J.n_team;
    const int32_t me = J.self_id[q];
    int n_self = 0, last = -1;
    for (int i = 0; i < J.n_team; ++i) {
      if (ids[i] == me) {
        ++n_self;
        last = i;
      }
    }
    const int n_friend = J.n_team - n_self;
    if (s == slots) {  // the frame's scalars
      double4 st = make_double4(0.0, 0.0, 0.0, 0.0);
      double hd = 0.0;
      if (last >= 0) {
        const float* r = J.team + ((size_t)q * J.n_team + last) * 5;
        st = make_double4((double)r[0] / 100., (double)r[1] / 100., (double)r[2] / 100., (double)r[3] / 100.);
        hd = (double)r[4];
      }
      reinterpret_cast<double4*>(J.self_state)[q] = st;
      J.heading[q] = hd;
      J.self_index[q] = last;
      J.n_obs[q] = n_friend + J.n_rival;
      continue;
    }
    double4 o = make_double4(0.0, 0.0, 0.0, 0.0);
    if (s < n_friend) {  // the s-th team entry that is not the robot itself
      int seen = 0, src = -1;
      for (int i = 0; i < J.n_team; ++i) {
        if (ids[i] != me) {
          if (seen == s) src = i;
          ++seen;
        }
      }
      const float* r = J.team + ((size_t)q * J.n_team + src) * 5;
      o = make_double4((double)r[0] / 100., (double)r[1] / 100., (double)r[2] / 100., (double)r[3] / 100.);
    } else if (s < n_friend + J.n_rival) {
      const float4 r = reinterpret_cast<const float4*>(J.rival)[(size_t)q * J.n_rival + (s - n_friend)];
      o = make_double4(-(double)r.x * 1e-2, -(double)r.y * 1e-2, -(double)r.z * 1e-2, -(double)r.w * 1e-2);
    }
    reinterpret_cast<double4*>(J.obstacles)[(size_t)q * slots + s] = o;
  }
}

}  // namespace krrt
