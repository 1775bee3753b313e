// trajectory_service.hpp — the `compute_trajectory` service body, ROS-free, over the C ABI.
//
// Mirrors trajectory_server() of the reference (src/trajectory_generator_node.cpp:6-53): same request / response
// shape as rrtstar/TrajectoryService.srv as its callers use it (scripts/trajectory_demo.py:313-324: req.model,
// req.params, req.xi.state / .n, req.xf.state / .n; res.model, res.time, res.trajectory[i].state / .t,
// res.inputs[i].state / .t).  A ROS node binds it with
//     nh.advertiseService("compute_trajectory", [&](auto& req, auto& res) { return trajectory_server(session, req, res); });
// after copying the message fields.
//
//   model "angular_trajectory"  — as the reference: params = {v_max, a_max} (optional, AngleController::setLimit),
//       xi.state = {w, dw}, xf.state = {yaw}; 25 rows of getAngleTrajectory(ctrl, offset, 0, t, 25, &u)
//       (trajectory_generator_node.cpp:9-31), computed by k_yaw_profiles; when xi.n <= 1 the reference solves nothing
//       and samples an empty control sequence, which throws inside getAngleTrajectory — here: returns false.
//   model "integrator2d"        — NEW (the reference's server answers "angular_trajectory" only and returns an empty
//       response for anything else): params = {r, dt} (optional; input weight, row spacing, defaults config.yaml:1 /
//       0.1 as in solver.solve(xi, xf, 0.1), …node.cpp:245), xi.state / xf.state = {x, y, vx, vy}; res.time = tau*,
//       rows (x, y, vx, vy) with inputs (ax, ay), computed by k_traj_count / k_traj_fill.
//   any other model             — as the reference: response left empty, returns true.
// No CPU path: every number comes from libkrrt.so kernels; without a device the Session throws krrt::Error.
#pragma once
#include <string>
#include <vector>

#include "rrtstar.hpp"

namespace krrt_srv {

struct StateMsg {            // the srv's state message: n, state[], t (ros::Time in the reference, seconds here)
  int n = 0;
  std::vector<double> state;
  double t = 0.0;
};
struct TrajectoryServiceRequest {
  std::string model;
  std::vector<double> params;
  StateMsg xi, xf;
};
struct TrajectoryServiceResponse {
  std::string model;
  double time = 0.0;
  std::vector<StateMsg> trajectory;
  std::vector<StateMsg> inputs;
};

inline bool trajectory_server(krrt::Session& session, const TrajectoryServiceRequest& req,
                              TrajectoryServiceResponse& res) {
  if (req.model == "angular_trajectory") {
    double v_max = 1.0, a_max = 1.0;  // Controller defaults (trajectory1d.h)
    if (req.params.size() > 1) {
      v_max = req.params.at(0);
      a_max = req.params.at(1);
    }
    if (!(req.xi.n > 1) || req.xi.state.size() < 2 || req.xf.state.empty()) return false;
    constexpr int N = 25;
    double w = req.xi.state.at(0), dw = req.xi.state.at(1), yaw = req.xf.state.at(0);
    double states[N][2], times[N], inputs[N], t = 0.0;
    int32_t status = 0, n_seg = 0;
    session.check(krrt_yaw_profiles(session.ctx(), &w, &dw, &yaw, 1, v_max, a_max, N, &states[0][0], times, inputs, &t,
                                    &status, &n_seg));
    if (status != 0) return false;  // the reference's exceptions ("case not match!", "control sequence error!")
    res.model = "angular_trajectory";
    res.time = t;
    for (int i = 0; i < N; ++i) {
      StateMsg state, input;
      state.n = 2;
      input.n = 1;
      input.state.push_back(inputs[i]);
      input.t = times[i];
      state.state.push_back(states[i][0]);
      state.state.push_back(states[i][1]);
      state.t = times[i];
      res.trajectory.push_back(state);
      res.inputs.push_back(input);
    }
    return true;
  }
  if (req.model == "integrator2d") {
    if (req.xi.state.size() < 4 || req.xf.state.size() < 4) return false;
    const double dt = req.params.size() > 1 ? req.params.at(1) : 0.1;
    if (!req.params.empty()) session.config().r = req.params.at(0);
    if (!(dt > 0.0)) return false;
    krrt_ctx* c = session.ctx();
    int n = 0;
    session.check(krrt_solve_inputs(c, req.xi.state.data(), req.xf.state.data(), dt, 0.0, nullptr, nullptr, 0, &n));
    std::vector<double> rows(5 * (size_t)n), u(2 * (size_t)n);
    session.check(krrt_solve_inputs(c, req.xi.state.data(), req.xf.state.data(), dt, 0.0, rows.data(), u.data(), n, &n));
    res.model = "integrator2d";
    res.time = n > 0 ? rows[5 * (size_t)(n - 1) + 4] : 0.0;
    for (int i = 0; i < n; ++i) {
      StateMsg state, input;
      state.n = 4;
      input.n = 2;
      state.state.assign(rows.begin() + 5 * i, rows.begin() + 5 * i + 4);
      state.t = rows[5 * (size_t)i + 4];
      input.state = {u[2 * (size_t)i], u[2 * (size_t)i + 1]};
      input.t = state.t;
      res.trajectory.push_back(state);
      res.inputs.push_back(input);
    }
    return true;
  }
  return true;  // unknown model: the reference answers with an empty response
}

}  // namespace krrt_srv
