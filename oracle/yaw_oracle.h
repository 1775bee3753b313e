/* yaw_oracle.h — CPU restatement of the reference's yaw (heading) profile generator, Trajectory1D::AngleController
 * (/root/reference/src/robosoccer_motion_planning/src/trajectory1d.cpp:23-51,203-314,326-394; called from
 * robosoccer_motion_planning_node.cpp:306-314 and trajectory_generator_node.cpp:6-53).  SURVEY.md §8(f) row 2.
 *
 * TEST INFRASTRUCTURE ONLY: loaded by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg; the product
 * (libkrrt.so) never links or calls it.
 *
 * PARITY PINNED against the reference's own code: oracle/_ref/libyaw_ref.so is that trajectory1d.cpp compiled where
 * it lies (oracle/Makefile, target `ref`), tests/golden/yaw_ref.npz holds vectors generated from it, and the known
 * answers of test/unit_test.cpp:7-86 are checked too (tests/test_yaw.py).
 */
#ifndef YAW_ORACLE_H
#define YAW_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

enum { /* Trajectory1D::Case, include/trajectory1d.h:14-22 */
  YO_INITIAL_STATE = 0,
  YO_ACCELERATION1,
  YO_ACCELERATION2_1,
  YO_ACCELERATION2_2,
  YO_CRUISING,
  YO_DECELERATION1,
  YO_DECELERATION2
};
enum { YO_OK = 0, YO_ERR_CASE = 1 /* "case not match!" */, YO_ERR_SEQUENCE = 2 /* "control sequence error!" */ };

#define YO_MAX_SEG 102

typedef struct yo_segment { /* Trajectory1D::Control, include/trajectory1d.h:24-29 */
  int32_t control_case;
  int32_t pad;
  double effort, time, w, dw; /* term = (w, dw) */
} yo_segment;

typedef struct yo_control { /* Controller::Control, include/trajectory1d.h:35-43 (+ the AngleOffset of :77-81) */
  int32_t n; /* entries incl. the initial state */
  int32_t pad;
  double final_time, offset, distance, a_max, v_max, f_offset, t_offset;
  yo_segment seg[YO_MAX_SEG];
} yo_control;

int yo_optimal_control(double w, double dw, double final_state, double v_max, double a_max, yo_control* cs);
double yo_shortest_path(double from, double to, double* f_offset, double* t_offset);
int yo_angle_control(double w, double dw, double final_state, double v_max, double a_max, yo_control* cs);
void yo_get_state(const yo_control* cs, double time, double* w, double* dw, double* effort, int* valid);
int yo_angle_trajectory(const yo_control* cs, double t0, double tf, int n, double* states, double* times,
                        double* inputs);
/* q queries, n samples each over [0, final_time]: states q x max(n,1) x 2, times / inputs q x max(n,1) */
void yo_yaw_profiles(const double* w, const double* dw, const double* yaw, int q, double v_max, double a_max, int n,
                     double* states, double* times, double* inputs, double* final_time, int32_t* status,
                     int32_t* n_seg);

#ifdef __cplusplus
}
#endif
#endif
