// rrtstar_custom_benchmark — the reference's `rosrun rrtstar rrtstar_custom_benchmark` (README.md:39-42: "custom
// settings (report min, max, mean also prints individual test results)") over this repository's component surface:
// computation time and solution cost against node size, with and without direct goal sampling — the quantities of the
// README's time_plot / cost_plot / box plots (README.md:44-53).  The reference's own benchmark source sits in the
// absent `src/motion_planning` submodule; scene (C1 of SURVEY.md 8d), goal region and trial count are this tool's.
//
// Plain C++17 + libkrrt.so (no ROS, no google-benchmark).  One planner singleton set as the node does
// (robosoccer_motion_planning_node.cpp:43-61,84-92,131-141), one `grow(xg, n)` per plan.
//   rrtstar_custom_benchmark [trials=20] [-v]        -v prints every individual result
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <vector>

#include "../../include/integrator2drrt.hpp"

struct ObState : std::array<double, 4> {
  const double& operator()(size_t i) const { return (*this)[i]; }
};

int main(int argc, char** argv) {
  int trials = 20;
  bool verbose = false;
  for (int a = 1; a < argc; ++a) {
    if (!std::strcmp(argv[a], "-v")) verbose = true;
    else trials = std::max(1, std::atoi(argv[a]));
  }
  try {
    Models::init_integrator2d();
    auto& rrt = Kinodynamic::Wrapper::get_rrtstar_int2d();
    auto& env = Kinodynamic::Wrapper::get_robosoccer_env();
    auto& goal = Kinodynamic::Wrapper::get_goal();
    auto& sampler = Kinodynamic::Wrapper::get_sampler();
    auto& solver = Kinodynamic::Wrapper::get_trajectory_solver();
    auto& tree = Kinodynamic::Wrapper::get_tree_int2d();
    env.collision_radius = 1.35;   // config/config.yaml:5
    solver.set_weight(0.5);        // config/config.yaml:1
    tree.capacity = 256;
    const double sp[9][2] = {{-2, 1}, {-2, -1}, {-5, 2}, {-5, -2}, {8.5, 0}, {2, 1}, {2, -1}, {5, 2}, {5, -2}};
    std::vector<ObState> obstacles(9);
    for (int j = 0; j < 9; ++j) obstacles[j] = ObState{{sp[j][0], sp[j][1], 0.0, 0.0}};
    env.setObstacles(obstacles);
    auto xg = goal.randomGoal();
    xg(0) = 7.0; xg(1) = 0.0; xg(2) = 0.0; xg(3) = 0.0;
    Kinodynamic::TreeInt2D::State start;
    start(0) = -8.5; start(1) = 0.0; start(2) = 0.0; start(3) = 0.0;

    const int sizes[] = {10, 25, 50, 75, 100, 150, 200};
    const double probs[] = {0.0, 0.1, 0.25, 0.5};
    std::printf("# rrtstar_custom_benchmark: %d trials per point, scene C1 (9 static robots), goal connected (O8)\n", trials);
    std::printf("# %-10s %-9s %10s %10s %10s %8s %10s %10s %10s\n", "p_direct", "node_size", "t_min_ms", "t_mean_ms",
                "t_max_ms", "solved", "cost_min", "cost_mean", "cost_max");
    for (double p : probs) {
      sampler.set_direct_sample(p > 0.0, p);
      sampler.target = xg;
      for (int n : sizes) {
        std::vector<double> times, costs;
        for (int trial = -1; trial < trials; ++trial) {  // trial -1: warm-up (context creation, first launch)
          sampler.query_id = (uint32_t)(trial < 0 ? 0 : trial);   // the seed of this trial's sample stream
          const auto t0 = std::chrono::steady_clock::now();
          rrt.setStart(start);
          rrt.setIteration(0);
          const bool solved = rrt.grow(&xg, n);
          const int gi = rrt.goalIndex();
          const double cost = solved ? tree.cost()[gi] : 0.0;   // reads the tree back, as the node's visualiser does
          const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
          if (trial < 0) continue;
          times.push_back(ms);
          if (solved) costs.push_back(cost);
          if (verbose) std::printf("  p=%.2f n=%d trial=%d time=%.3f ms solved=%d cost=%.4f\n", p, n, trial, ms, (int)solved, cost);
        }
        const double tm = std::accumulate(times.begin(), times.end(), 0.0) / times.size();
        const double cm = costs.empty() ? 0.0 : std::accumulate(costs.begin(), costs.end(), 0.0) / costs.size();
        std::printf("  %-10.2f %-9d %10.3f %10.3f %10.3f %5zu/%-3d %10.4f %10.4f %10.4f\n", p, n,
                    *std::min_element(times.begin(), times.end()), tm, *std::max_element(times.begin(), times.end()),
                    costs.size(), trials, costs.empty() ? 0.0 : *std::min_element(costs.begin(), costs.end()), cm,
                    costs.empty() ? 0.0 : *std::max_element(costs.begin(), costs.end()));
      }
    }
  } catch (const krrt::Error& e) {
    std::printf("krrt::Error %d: %s (no CPU fallback)\n", e.code, e.what());
    return e.code == KRRT_ERR_NO_DEVICE ? 3 : 4;
  }
  return 0;
}
