// krrt_yaw.cuh — k_yaw_profiles: batched yaw (heading) profiles, SURVEY.md §8(f) row 2.
// Stands in for Trajectory1D::AngleController::angleControl + getAngleTrajectory
// (src/trajectory1d.cpp:23-51,203-314,326-394) as called once per planned path by
// robosoccer_motion_planning_node.cpp:306-314 and by the compute_trajectory service
// (trajectory_generator_node.cpp:6-53), for q independent (heading, rate, target yaw) queries at once.
//
// One warp per query: lane 0 walks the bang-bang case machine (a handful of segments) into shared memory, then the
// 32 lanes evaluate the n sample points and write the rows coalesced.  Output-bandwidth bound: 32 B per sample row
// (w, dw, t, u).  Arithmetic: binary64, the reference's operations in the reference's order, no contraction
// (-fmad=false), so rows compare bit-for-bit with the reference library (tests/test_yaw.py).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

namespace krrt {

constexpr int YAW_MAX_SEG = 102;      // the reference gives up ("control sequence error!") past 100 entries
constexpr int YAW_WARPS = 4;
constexpr double YAW_PI = 3.14159265358979323846;  // M_PI

struct YawSeg {
  double effort, time, w, dw;  // Trajectory1D::Control: effort, time, term = (w, dw)
};

struct YawJob {
  const double* w;
  const double* dw;
  const double* yaw;
  int q, n;
  double v_max, a_max;
  double* states;      // q x max(n,1) x 2
  double* times;       // q x max(n,1)
  double* inputs;      // q x max(n,1)
  double* final_time;  // q
  int32_t* status;     // q: 0 ok, 1 "case not match!", 2 "control sequence error!"
  int32_t* n_seg;      // q: entries of the control sequence incl. the initial state
};

__device__ __forceinline__ double yaw_sign(double x) { return x < 0 ? -1.0 : 1.0; }  // trajectory1d.cpp:6

__device__ __forceinline__ double yaw_wrap(double a) {  // trajectory1d.cpp:338-342
  if (a > YAW_PI) a = -YAW_PI + (a - YAW_PI);
  else if (a < -YAW_PI) a = YAW_PI + (a + YAW_PI);
  return a;
}

// Controller::applyControl + case1 .. case3 (trajectory1d.cpp:203-314); returns 0 / 1 ("case not match!")
__device__ __forceinline__ int yaw_apply(YawSeg* seg, int n, double w, double dw, double fin, double v_max,
                                         double a_max) {
  const double wf = fin - w;
  const double sg = yaw_sign(wf);
  const double final_n = wf * sg;
  const double dw0 = dw * sg;
  const double tmp = dw0 * dw0 / (2.0 * a_max);
  const double wf_diff = final_n - tmp;
  YawSeg c;
  if (dw0 < 0) {  // case1
    c.effort = a_max;
    c.time = -dw0 / a_max;
    c.w = -dw0 * dw0 / (2.0 * a_max);
    c.dw = 0;
  } else if (dw0 < v_max && wf_diff > 0.00001) {  // case21
    const double dw1 = sqrt((final_n * a_max) + (dw0 * dw0 / 2.0));
    const double t1 = (v_max - dw0) / a_max;
    const double t2 = (dw1 - dw0) / a_max;
    c.effort = a_max;
    if (t1 < t2) {
      c.time = t1;
      c.dw = v_max;
      c.w = (v_max * v_max - dw0 * dw0) / (2.0 * a_max);
    } else {
      c.time = t2;
      c.dw = dw1;
      c.w = (final_n / 2.0) + (dw0 * dw0 / (2.0 * a_max));
    }
  } else if (dw0 == v_max && wf_diff > 0.00001) {  // case22
    c.effort = 0.0;
    c.dw = v_max;
    c.w = final_n - (v_max * v_max) / (2 * a_max);
    c.time = (final_n / v_max) - (v_max / (2 * a_max));
  } else if (dw0 <= v_max && wf_diff <= 0.00001) {  // case23
    c.effort = -a_max;
    c.dw = 0.0;
    c.w = (dw0 * dw0) / (2 * a_max);
    c.time = dw0 / a_max;
  } else if (dw0 > v_max) {  // case3
    c.effort = -a_max;
    c.dw = v_max;
    c.w = (dw0 * dw0 - v_max * v_max) / (2 * a_max);
    c.time = (dw0 - v_max) / a_max;
  } else {
    return 1;
  }
  c.effort *= sg;
  c.dw *= sg;
  c.w *= sg;
  c.w += w;
  seg[n] = c;
  return 0;
}

// Controller::getState (trajectory1d.cpp:53-81)
__device__ __forceinline__ void yaw_state(const YawSeg* seg, int n, double final_time, double distance, double time,
                                          double& w, double& dw, double& effort) {
  w = 0.0;
  dw = 0.0;
  effort = 0.0;
  if (time < final_time) {
    double t = 0.0, dt = time;
    for (int i = 1; i < n; ++i) {
      const YawSeg c = seg[i];
      effort = c.effort;
      t += c.time;
      if (time < t) {
        const YawSeg s0 = seg[i - 1];
        dw = s0.dw + c.effort * dt;
        w = s0.w + s0.dw * dt + c.effort * dt * dt / 2.0;
        break;
      }
      dt -= c.time;
    }
  } else {
    w = distance;
  }
}

__global__ void __launch_bounds__(YAW_WARPS * 32) k_yaw_profiles(YawJob J) {
  __shared__ YawSeg s_seg[YAW_WARPS][YAW_MAX_SEG];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  YawSeg* seg = s_seg[warp];
  const int rows = J.n > 0 ? J.n : 1;
  for (int k = blockIdx.x * YAW_WARPS + warp; k < J.q; k += gridDim.x * YAW_WARPS) {
    int n = 1, rc = 0;
    double final_time = 0.0, distance = 0.0, f_offset = 0.0;
    if (lane == 0) {
      double w = J.w[k], dw = J.dw[k];
      double to = J.yaw[k];
      // AngleController::shortest_path (trajectory1d.cpp:362-394) + angleControl (:326-332)
      {
        const double f = w, t = to;
        const double from_n = yaw_wrap(w), to_n = yaw_wrap(to);
        double d = to_n - from_n;
        double t_offset = 0.0;
        if (fabs(d) > YAW_PI) {
          const int s = (d > 0 ? 1 : -1);
          d = -s * (2 * YAW_PI - (fabs(from_n) + fabs(to_n)));
          f_offset = -f;
          t_offset = d - t;
        }
        w = w + f_offset;
        to = to + t_offset;
      }
      // Controller::optimalControl (trajectory1d.cpp:23-51)
      seg[0].effort = 0.0;
      seg[0].time = 0.0;
      seg[0].w = w;
      seg[0].dw = dw;
      for (;;) {
        const double w_diff = to - w;
        if ((fabs(w_diff) <= 0.00001) && (fabs(dw) <= 0.001)) break;
        rc = yaw_apply(seg, n, w, dw, to, J.v_max, J.a_max);
        if (rc) break;
        ++n;
        distance = seg[n - 1].w;
        if (n > 100) {
          rc = 2;
          break;
        }
        final_time += fabs(seg[n - 1].time);
        w = seg[n - 1].w;
        dw = seg[n - 1].dw;
      }
      J.status[k] = rc;
      J.n_seg[k] = n;
      J.final_time[k] = final_time;
    }
    __syncwarp();
    n = __shfl_sync(0xffffffffu, n, 0);
    rc = __shfl_sync(0xffffffffu, rc, 0);
    final_time = __shfl_sync(0xffffffffu, final_time, 0);
    distance = __shfl_sync(0xffffffffu, distance, 0);
    f_offset = __shfl_sync(0xffffffffu, f_offset, 0);
    double2* S = reinterpret_cast<double2*>(J.states) + (size_t)rows * k;
    double* T = J.times + (size_t)rows * k;
    double* U = J.inputs + (size_t)rows * k;
    if (rc) {
      for (int i = lane; i < rows; i += 32) {
        S[i] = make_double2(0.0, 0.0);
        T[i] = 0.0;
        U[i] = 0.0;
      }
    } else if (J.n <= 0) {  // Controller::getTrajectory's empty case (trajectory1d.cpp:97-102)
      if (lane == 0) {
        S[0] = make_double2(yaw_wrap(seg[0].w - f_offset), seg[0].dw);
        T[0] = 0.0;
        U[0] = 0.0;
      }
    } else {
      // getTrajectory(ctrl, 0, final_time, n, inputs) (trajectory1d.cpp:83-96) + the wrap of getAngleTrajectory (:334-346)
      const double dt = (final_time - 0.0) / J.n;
      for (int i = lane + 1; i <= J.n; i += 32) {
        const double now = (double)i * dt;
        double w, dw, effort;
        yaw_state(seg, n, final_time, distance, now, w, dw, effort);
        S[i - 1] = make_double2(yaw_wrap(w - f_offset), dw);
        T[i - 1] = 0.0 + now;
        U[i - 1] = effort;
      }
    }
    __syncwarp();
  }
}

}  // namespace krrt
