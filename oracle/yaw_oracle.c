/* yaw_oracle.c — CPU restatement of the reference's 1-D bang-bang yaw profile generator
 * (Trajectory1D::Controller / AngleController).  TEST INFRASTRUCTURE ONLY (see yaw_oracle.h).
 *
 * PARITY PINNED: unlike the RRT* path, the reference's own sources for this path are in the mounted tree and
 * compile on their own (oracle/Makefile target `ref` builds oracle/_ref/libyaw_ref.so from
 * /root/reference/src/robosoccer_motion_planning/src/trajectory1d.cpp where it lies).  tests/test_yaw.py checks
 * this restatement bit-for-bit against that library, against fixtures generated from it
 * (tests/golden/yaw_ref.npz) and against the known answers of the reference's unit test
 * (test/unit_test.cpp:7-86).
 *
 * All citations: /root/reference/src/robosoccer_motion_planning/src/trajectory1d.cpp.
 * Arithmetic: binary64, only + - * / sqrt fabs, written operation by operation in the reference's order
 * (compile with -ffp-contract=off; the reference library is built without FMA).
 */
#include "yaw_oracle.h"

#include <math.h>
#include <string.h>

#define YO_SIGN(X) ((X) < 0 ? -1.0 : 1.0) /* trajectory1d.cpp:6 */

/* Controller::case1 .. case3 (trajectory1d.cpp:236-314) on the normalised state (w0 = 0, dw0, final > 0 side) */
static void yo_case(int which, double dw0, double fin, double v_max, double a_max, yo_segment* c) {
  switch (which) {
    case YO_ACCELERATION1: /* :236-248 */
      c->control_case = YO_ACCELERATION1;
      c->effort = a_max;
      c->time = -dw0 / a_max;
      c->w = -dw0 * dw0 / (2.0 * a_max);
      c->dw = 0;
      break;
    case YO_ACCELERATION2_1: { /* case21 :250-274 */
      double dw1 = sqrt((fin * a_max) + (dw0 * dw0 / 2.0));
      double t1 = (v_max - dw0) / a_max;
      double t2 = (dw1 - dw0) / a_max;
      c->effort = a_max;
      if (t1 < t2) {
        c->control_case = YO_ACCELERATION2_1;
        c->time = t1;
        c->dw = v_max;
        c->w = (v_max * v_max - dw0 * dw0) / (2.0 * a_max);
      } else {
        c->control_case = YO_ACCELERATION2_2;
        c->time = t2;
        c->dw = dw1;
        c->w = (fin / 2.0) + (dw0 * dw0 / (2.0 * a_max));
      }
      break;
    }
    case YO_CRUISING: /* case22 :276-286 */
      c->control_case = YO_CRUISING;
      c->effort = 0.0;
      c->dw = v_max;
      c->w = fin - (v_max * v_max) / (2 * a_max);
      c->time = (fin / v_max) - (v_max / (2 * a_max));
      break;
    case YO_DECELERATION1: /* case23 :288-298 */
      c->control_case = YO_DECELERATION1;
      c->effort = -a_max;
      c->dw = 0.0;
      c->w = (dw0 * dw0) / (2 * a_max);
      c->time = dw0 / a_max;
      break;
    default: /* case3 :300-310 */
      c->control_case = YO_DECELERATION2;
      c->effort = -a_max;
      c->dw = v_max;
      c->w = (dw0 * dw0 - v_max * v_max) / (2 * a_max);
      c->time = (dw0 - v_max) / a_max;
      break;
  }
}

/* Controller::applyControl (trajectory1d.cpp:203-234): returns 0, or YO_ERR_CASE for "case not match!" */
static int yo_apply(yo_control* cs, double w, double dw, double fin) {
  double wf = fin - w;
  double normal_wf = wf * YO_SIGN(wf);
  double dw0 = dw * YO_SIGN(wf);
  double tmp = dw0 * dw0 / (2.0 * cs->a_max);
  double wf_diff = normal_wf - tmp;
  int which;
  if (dw0 < 0) which = YO_ACCELERATION1;
  else if (dw0 < cs->v_max && wf_diff > 0.00001) which = YO_ACCELERATION2_1;
  else if (dw0 == cs->v_max && wf_diff > 0.00001) which = YO_CRUISING;
  else if (dw0 <= cs->v_max && wf_diff <= 0.00001) which = YO_DECELERATION1;
  else if (dw0 > cs->v_max) which = YO_DECELERATION2;
  else return YO_ERR_CASE;
  yo_segment* c = &cs->seg[cs->n];
  yo_case(which, dw0, normal_wf, cs->v_max, cs->a_max, c);
  cs->n++;
  c->effort *= YO_SIGN(wf);
  c->dw *= YO_SIGN(wf);
  c->w *= YO_SIGN(wf);
  c->w += w;
  cs->distance = c->w;
  if (cs->n > 100) return YO_ERR_SEQUENCE; /* "control sequence error!" :232-233 */
  return 0;
}

/* Controller::optimalControl (trajectory1d.cpp:23-51) */
int yo_optimal_control(double w, double dw, double final_state, double v_max, double a_max, yo_control* cs) {
  memset(cs, 0, sizeof(*cs));
  cs->v_max = v_max;
  cs->a_max = a_max;
  cs->seg[0].control_case = YO_INITIAL_STATE;
  cs->seg[0].w = w;
  cs->seg[0].dw = dw;
  cs->n = 1;
  cs->offset = w < 0.0 ? fabs(w) : 0.0;
  cs->distance = 0.0;
  double t = 0.0;
  for (;;) {
    double w_diff = final_state - w;
    if ((fabs(w_diff) <= 0.00001) && (fabs(dw) <= 0.001)) break;
    int rc = yo_apply(cs, w, dw, final_state);
    if (rc) {
      cs->final_time = t;
      return rc;
    }
    t += fabs(cs->seg[cs->n - 1].time);
    w = cs->seg[cs->n - 1].w;
    dw = cs->seg[cs->n - 1].dw;
  }
  cs->final_time = t;
  return 0;
}

/* AngleController::shortest_path (trajectory1d.cpp:362-394) */
double yo_shortest_path(double from, double to, double* f_offset, double* t_offset) {
  *f_offset = 0.0;
  *t_offset = 0.0;
  double f = from, t = to;
  if (from > M_PI) from = -M_PI + (from - M_PI);
  else if (from < -M_PI) from = M_PI + (from + M_PI);
  if (to > M_PI) to = -M_PI + (to - M_PI);
  else if (to < -M_PI) to = M_PI + (to + M_PI);
  double d = to - from;
  if (fabs(d) > M_PI) {
    int s = (d > 0 ? 1 : -1);
    d = -s * (2 * M_PI - (fabs(from) + fabs(to)));
    *f_offset = -f;
    *t_offset = d - t;
  }
  return d;
}

/* AngleController::angleControl (trajectory1d.cpp:326-332) */
int yo_angle_control(double w, double dw, double final_state, double v_max, double a_max, yo_control* cs) {
  double f_offset, t_offset;
  yo_shortest_path(w, final_state, &f_offset, &t_offset);
  int rc = yo_optimal_control(w + f_offset, dw, final_state + t_offset, v_max, a_max, cs);
  cs->f_offset = f_offset;
  cs->t_offset = t_offset;
  return rc;
}

/* Controller::getState (trajectory1d.cpp:53-81).  *valid = 0 where the reference returns an uninitialised State
 * (time < final_time but past the summed segment times; cannot happen with non-negative segment times). */
void yo_get_state(const yo_control* cs, double time, double* w, double* dw, double* effort, int* valid) {
  double t = 0.0, dt = time;
  *w = 0.0;
  *dw = 0.0;
  *effort = 0.0;
  *valid = 1;
  if (time < cs->final_time) {
    *valid = 0;
    for (int i = 1; i < cs->n; ++i) {
      const yo_segment* c = &cs->seg[i];
      *effort = c->effort;
      t += c->time;
      const yo_segment* s0 = &cs->seg[i - 1];
      if (time < t) {
        *dw = s0->dw + c->effort * dt;
        *w = s0->w + s0->dw * dt + c->effort * dt * dt / 2.0;
        *valid = 1;
        break;
      }
      dt -= c->time;
    }
  } else {
    *w = cs->distance;
    *dw = 0.0;
    *effort = 0.0;
  }
}

/* AngleController::getAngleTrajectory(ctrl, offset, t0, tf, n, inputs) (trajectory1d.cpp:334-346) over
 * Controller::getTrajectory(ctrl, t0, tf, n, inputs) (:83-104).  rows = max(n, 1); returns the row count. */
int yo_angle_trajectory(const yo_control* cs, double t0, double tf, int n, double* states, double* times,
                        double* inputs) {
  double dt = (tf - t0) / n;
  int rows = 0;
  for (int i = 1; i <= n; ++i) {
    double now = (double)i * dt;
    double w, dw, effort;
    int valid;
    yo_get_state(cs, now, &w, &dw, &effort, &valid);
    states[2 * rows] = w;
    states[2 * rows + 1] = dw;
    times[rows] = t0 + now;
    inputs[rows] = effort;
    ++rows;
  }
  if (rows == 0) {
    states[0] = cs->seg[0].w;
    states[1] = cs->seg[0].dw;
    times[0] = t0;
    inputs[0] = 0.0; /* the reference pushes no input here */
    rows = 1;
  }
  for (int k = 0; k < rows; ++k) {
    double a = states[2 * k] - cs->f_offset;
    if (a > M_PI) a = -M_PI + (a - M_PI);
    else if (a < -M_PI) a = M_PI + (a + M_PI);
    states[2 * k] = a;
  }
  return rows;
}

/* the batched operator the CUDA kernel implements: q independent (w, dw, yaw) queries, n samples each */
void yo_yaw_profiles(const double* w, const double* dw, const double* yaw, int q, double v_max, double a_max, int n,
                     double* states, double* times, double* inputs, double* final_time, int32_t* status,
                     int32_t* n_seg) {
  const int rows_per = n > 0 ? n : 1;
  for (int k = 0; k < q; ++k) {
    yo_control cs;
    int rc = yo_angle_control(w[k], dw[k], yaw[k], v_max, a_max, &cs);
    status[k] = rc;
    n_seg[k] = cs.n;
    final_time[k] = cs.final_time;
    double* S = states + (size_t)2 * rows_per * k;
    double* T = times + (size_t)rows_per * k;
    double* U = inputs + (size_t)rows_per * k;
    if (rc) {
      for (int i = 0; i < rows_per; ++i) S[2 * i] = S[2 * i + 1] = T[i] = U[i] = 0.0;
      continue;
    }
    yo_angle_trajectory(&cs, 0.0, cs.final_time, n, S, T, U);
  }
}
