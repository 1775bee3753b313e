// service_check.cpp — compute_trajectory service body (include/trajectory_service.hpp) called the way
// scripts/trajectory_demo.py:313-324 fills the request, checked row for row against the CPU checkers.
// TEST ONLY: links oracle/_build/libyaw_oracle.so and libkrrt_oracle.so.  Exit 0 = identical; 3 = no GPU.
#include <cstdio>
#include <cstring>
#include <random>

#include "../../include/trajectory_service.hpp"
#include "../../oracle/krrt_oracle.h"
#include "../../oracle/yaw_oracle.h"

static int fails = 0;
#define EXPECT(cond)                                                      \
  do {                                                                    \
    if (!(cond)) {                                                        \
      if (fails < 20) std::printf("FAIL %s:%d %s\n", __FILE__, __LINE__, #cond); \
      ++fails;                                                            \
    }                                                                     \
  } while (0)

static bool same(double a, double b) { return std::memcmp(&a, &b, sizeof(double)) == 0 || a == b; }

int main() {
  krrt::Session session;
  try {
    session.ctx();
  } catch (const krrt::Error& e) {
    std::printf("krrt::Error %d: %s (no CPU fallback)\n", e.code, e.what());
    return e.code == KRRT_ERR_NO_DEVICE ? 3 : 4;
  }
  std::mt19937_64 gen(5);
  std::uniform_real_distribution<double> ang(-4.0, 4.0), rate(-2.5, 2.5), px(-9, 9), py(-6, 6), v(-2, 2);
  int served = 0;
  for (int trial = 0; trial < 200; ++trial) {
    krrt_srv::TrajectoryServiceRequest req;
    krrt_srv::TrajectoryServiceResponse res;
    req.model = "angular_trajectory";
    const bool limits = trial & 1;
    if (limits) req.params = {1.5, 0.75};
    req.xi.state = {ang(gen), (trial % 5) ? rate(gen) : 0.0};
    req.xf.state = {ang(gen)};
    req.xi.n = 2;
    req.xf.n = 1;
    const bool ok = krrt_srv::trajectory_server(session, req, res);
    double st[25][2], tm[25], in[25], ft;
    int32_t status, nseg;
    yo_yaw_profiles(&req.xi.state[0], &req.xi.state[1], &req.xf.state[0], 1, limits ? 1.5 : 1.0, limits ? 0.75 : 1.0, 25,
                    &st[0][0], tm, in, &ft, &status, &nseg);
    EXPECT(ok == (status == 0));
    if (!ok) continue;
    ++served;
    EXPECT(res.model == "angular_trajectory" && same(res.time, ft));
    EXPECT(res.trajectory.size() == 25 && res.inputs.size() == 25);
    for (size_t i = 0; i < res.trajectory.size() && i < 25; ++i) {
      EXPECT(res.trajectory[i].n == 2 && res.inputs[i].n == 1);
      EXPECT(same(res.trajectory[i].state.at(0), st[i][0]) && same(res.trajectory[i].state.at(1), st[i][1]));
      EXPECT(same(res.trajectory[i].t, tm[i]) && same(res.inputs[i].state.at(0), in[i]) && same(res.inputs[i].t, tm[i]));
    }
  }
  EXPECT(served > 150);
  {  // xi.n <= 1: nothing to solve (trajectory_generator_node.cpp:19), no response
    krrt_srv::TrajectoryServiceRequest req;
    krrt_srv::TrajectoryServiceResponse res;
    req.model = "angular_trajectory";
    req.xi.state = {0.5};
    req.xi.n = 1;
    req.xf.state = {1.0};
    EXPECT(!krrt_srv::trajectory_server(session, req, res) && res.trajectory.empty());
    req.model = "no_such_model";  // the reference returns true with an empty response
    EXPECT(krrt_srv::trajectory_server(session, req, res) && res.trajectory.empty() && res.model.empty());
  }
  for (int trial = 0; trial < 100; ++trial) {
    krrt_srv::TrajectoryServiceRequest req;
    krrt_srv::TrajectoryServiceResponse res;
    req.model = "integrator2d";
    const double r = (trial & 1) ? 0.5 : 1.25, dt = (trial % 3) ? 0.1 : 0.037;
    req.params = {r, dt};
    req.xi.state = {px(gen), py(gen), v(gen), v(gen)};
    req.xf.state = {px(gen), py(gen), v(gen), v(gen)};
    if (trial == 7) req.xf.state = req.xi.state;  // zero-length edge
    req.xi.n = req.xf.n = 4;
    EXPECT(krrt_srv::trajectory_server(session, req, res));
    double rows[4096 * 5], u[4096 * 2];
    const int n = ko_solve(req.xi.state.data(), req.xf.state.data(), r, dt, 0.0, rows, 4096);
    const int nu = ko_solve_inputs(req.xi.state.data(), req.xf.state.data(), r, dt, u, 4096);
    EXPECT(n == nu && (int)res.trajectory.size() == n && (int)res.inputs.size() == n);
    double tau, cost;
    ko_steer(req.xi.state.data(), req.xf.state.data(), r, &tau, &cost);
    EXPECT(same(res.time, tau));
    for (int i = 0; i < n && i < (int)res.trajectory.size(); ++i) {
      for (int d = 0; d < 4; ++d) EXPECT(same(res.trajectory[i].state.at(d), rows[5 * i + d]));
      EXPECT(same(res.trajectory[i].t, rows[5 * i + 4]));
      EXPECT(same(res.inputs[i].state.at(0), u[2 * i]) && same(res.inputs[i].state.at(1), u[2 * i + 1]));
    }
    // the inputs ARE the trajectory's second derivative: integrate u over the rows (trapezoid is exact for linear u)
    if (n > 2 && tau > 0.0) {
      double vx = req.xi.state[2], vy = req.xi.state[3];
      for (int i = 1; i < n; ++i) {
        const double h = rows[5 * i + 4] - rows[5 * (i - 1) + 4];
        vx += 0.5 * h * (u[2 * i] + u[2 * (i - 1)]);
        vy += 0.5 * h * (u[2 * i + 1] + u[2 * (i - 1) + 1]);
      }
      EXPECT(std::fabs(vx - req.xf.state[2]) < 1e-9 * (1.0 + std::fabs(vx)) * 1e3);
      EXPECT(std::fabs(vy - req.xf.state[3]) < 1e-9 * (1.0 + std::fabs(vy)) * 1e3);
    }
  }
  std::printf("service_check: %d angular requests served; %d check failure(s)\n", served, fails);
  return fails ? 1 : 0;
}
