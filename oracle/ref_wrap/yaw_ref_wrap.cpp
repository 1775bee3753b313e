// yaw_ref_wrap.cpp — extern "C" face of the REFERENCE's own Trajectory1D::AngleController, so the tests can call the
// real thing.  TEST INFRASTRUCTURE ONLY.  This file contains no reference code: it includes the reference's header
// and is linked with the reference's trajectory1d.cpp, both compiled where they lie under /root/reference
// (oracle/Makefile, target `ref`, output oracle/_ref/libyaw_ref.so).  Same batched signature as yo_yaw_profiles.
#include <cstdint>
#include <cstddef>
#include <stdexcept>
#include <string>
#include <vector>

#include "trajectory1d.h"

extern "C" {

double yr_shortest_path(double from, double to, double* f_offset, double* t_offset) {
  Trajectory1D::AngleController c;
  return c.shortest_path(from, to, *f_offset, *t_offset);
}

void yr_yaw_profiles(const double* w, const double* dw, const double* yaw, int q, double v_max, double a_max, int n,
                     double* states, double* times, double* inputs, double* final_time, int32_t* status,
                     int32_t* n_seg) {
  const int rows_per = n > 0 ? n : 1;
  for (int k = 0; k < q; ++k) {
    double* S = states + (size_t)2 * rows_per * k;
    double* T = times + (size_t)rows_per * k;
    double* U = inputs + (size_t)rows_per * k;
    for (int i = 0; i < rows_per; ++i) S[2 * i] = S[2 * i + 1] = T[i] = U[i] = 0.0;
    Trajectory1D::AngleController c;
    c.setLimit(v_max, a_max);
    double t = 0.0;
    try {
      auto ctrl = c.angleControl({w[k], dw[k]}, yaw[k], t);   // robosoccer_motion_planning_node.cpp:306
      status[k] = 0;
      n_seg[k] = (int32_t)ctrl.size();
      final_time[k] = t;
      std::vector<double> u;
      auto trj = Trajectory1D::AngleController::getAngleTrajectory(ctrl, c.getOffset(), 0.0, t, n, &u);  // :314
      for (size_t i = 0; i < trj.first.size() && (int)i < rows_per; ++i) {
        S[2 * i] = trj.first[i].w;
        S[2 * i + 1] = trj.first[i].dw;
        T[i] = trj.second[i];
        if (i < u.size()) U[i] = u[i];
      }
    } catch (const std::runtime_error& e) {
      status[k] = std::string(e.what()) == "case not match!" ? 1 : 2;
      n_seg[k] = 0;
      final_time[k] = 0.0;
    }
  }
}
}
