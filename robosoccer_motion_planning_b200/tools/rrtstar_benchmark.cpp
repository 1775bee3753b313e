// rrtstar_benchmark — the reference's `rosrun rrtstar rrtstar_benchmark` (README.md:34-37, "use google benchmark
// settings") over this repository's component surface.  The reference's own benchmark source sits in the absent
// `src/motion_planning` submodule and google-benchmark is not in this image, so this is a plain C++17 program that
// keeps google-benchmark's conventions: one named case per argument, repeated until `--benchmark_min_time` seconds
// have been spent (at least 3 runs), a `Benchmark / Time / CPU / Iterations` table on stdout, or
// `--benchmark_format=json` (context + benchmarks[] with real_time / cpu_time / iterations / counters).
//
// Cases (BASELINE.json config 1: 2-D double integrator, 1 robot, 9 static obstacles, fixed seed):
//   BM_RRTStarInt2D/<n>        one plan of n iterations through `rrt.grow(xg, n)`, n in {10,25,50,75,100,150,200}
//                              (node sizes of the README's time_plot / cost_plot) and 10000 (the config's own size)
//   BM_RRTStarInt2D_PerCall/75 the caller's own loop, `for (i < 75) rrt.grow(xg)` (…node.cpp:176-185; config.yaml:3)
// Counters per case: expansions/s, nodes, solved, cost.  Plain C++17 + libkrrt.so, no ROS; a missing GPU is an error
// (exit 3), there is no CPU path.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <string>
#include <vector>

#include "../../include/integrator2drrt.hpp"

struct ObState : std::array<double, 4> {
  const double& operator()(size_t i) const { return (*this)[i]; }
};

struct Result {
  std::string name;
  double real_ns = 0, cpu_ns = 0;
  long iterations = 0;
  double exp_per_s = 0, nodes = 0, solved = 0, cost = 0;
};

static double cpu_now() { return (double)std::clock() / CLOCKS_PER_SEC; }

int main(int argc, char** argv) {
  double min_time = 0.5;
  bool json = false;
  std::string filter;
  for (int a = 1; a < argc; ++a) {
    if (!std::strncmp(argv[a], "--benchmark_min_time=", 21)) min_time = std::atof(argv[a] + 21);
    else if (!std::strcmp(argv[a], "--benchmark_format=json")) json = true;
    else if (!std::strncmp(argv[a], "--benchmark_filter=", 19)) filter = argv[a] + 19;
  }
  try {
    Models::init_integrator2d();
    auto& rrt = Kinodynamic::Wrapper::get_rrtstar_int2d();
    auto& env = Kinodynamic::Wrapper::get_robosoccer_env();
    auto& sampler = Kinodynamic::Wrapper::get_sampler();
    auto& solver = Kinodynamic::Wrapper::get_trajectory_solver();
    auto& tree = Kinodynamic::Wrapper::get_tree_int2d();
    env.collision_radius = 1.35;  // config/config.yaml:5
    solver.set_weight(0.5);       // config/config.yaml:1
    tree.capacity = 16384;
    sampler.seed = 0;             // "fixed seed"
    sampler.set_direct_sample(false, 0.0);
    const double sp[9][2] = {{-2, 1}, {-2, -1}, {-5, 2}, {-5, -2}, {8.5, 0}, {2, 1}, {2, -1}, {5, 2}, {5, -2}};
    std::vector<ObState> obstacles(9);
    for (int j = 0; j < 9; ++j) obstacles[j] = ObState{{sp[j][0], sp[j][1], 0.0, 0.0}};
    env.setObstacles(obstacles);
    Kinodynamic::TreeInt2D::State xs, xg;
    xs(0) = -8.5, xs(1) = 0.0, xs(2) = 0.0, xs(3) = 0.0;  // nubot1 spawn pose (robot_up.sh:11-12)
    xg(0) = 7.0, xg(1) = 0.0, xg(2) = 0.0, xg(3) = 0.0;

    struct Case { const char* base; int n; bool per_call; };
    const Case cases[] = {{"BM_RRTStarInt2D", 10, false},  {"BM_RRTStarInt2D", 25, false},  {"BM_RRTStarInt2D", 50, false},
                          {"BM_RRTStarInt2D", 75, false},  {"BM_RRTStarInt2D", 100, false}, {"BM_RRTStarInt2D", 150, false},
                          {"BM_RRTStarInt2D", 200, false}, {"BM_RRTStarInt2D", 10000, false},
                          {"BM_RRTStarInt2D_PerCall", 75, true}};
    std::vector<Result> results;
    for (const Case& c : cases) {
      Result r;
      r.name = std::string(c.base) + "/" + std::to_string(c.n);
      if (!filter.empty() && r.name.find(filter) == std::string::npos) continue;
      auto plan = [&]() {
        rrt.setStart(xs);
        rrt.setIteration(0);
        bool solved = false;
        if (c.per_call)
          for (int i = 0; i < c.n; ++i) solved = rrt.grow(&xg);
        else
          solved = rrt.grow(&xg, c.n);
        return solved;
      };
      plan();  // warm-up: context creation, first launch
      double real = 0, cpu = 0;
      long it = 0;
      bool solved = false;
      while (it < 3 || real < min_time) {
        const double c0 = cpu_now();
        const auto t0 = std::chrono::steady_clock::now();
        solved = plan();
        real += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        cpu += cpu_now() - c0;
        ++it;
      }
      r.iterations = it;
      r.real_ns = real / it * 1e9;
      r.cpu_ns = cpu / it * 1e9;
      r.exp_per_s = c.n / (real / it);
      r.nodes = (double)tree.tree.size();
      r.solved = solved ? 1 : 0;
      r.cost = solved ? tree.cost().at((size_t)rrt.goalIndex()) : 0.0;
      results.push_back(r);
    }
    if (json) {
      std::printf("{\n  \"context\": {\"executable\": \"rrtstar_benchmark\", \"library\": \"libkrrt.so (sm_100a)\", \"scene\": "
                  "\"C1: 9 static robots, start (-8.5,0), goal (7,0), seed 0, p_direct 0\"},\n  \"benchmarks\": [\n");
      for (size_t i = 0; i < results.size(); ++i) {
        const Result& r = results[i];
        std::printf("    {\"name\": \"%s\", \"iterations\": %ld, \"real_time\": %.1f, \"cpu_time\": %.1f, \"time_unit\": \"ns\", "
                    "\"expansions_per_second\": %.1f, \"nodes\": %.0f, \"solved\": %.0f, \"cost\": %.9g}%s\n",
                    r.name.c_str(), r.iterations, r.real_ns, r.cpu_ns, r.exp_per_s, r.nodes, r.solved, r.cost,
                    i + 1 < results.size() ? "," : "");
      }
      std::printf("  ]\n}\n");
    } else {
      std::printf("%-34s %15s %15s %12s\n", "Benchmark", "Time", "CPU", "Iterations");
      std::printf("--------------------------------------------------------------------------------------------------------\n");
      for (const Result& r : results)
        std::printf("%-34s %12.0f ns %12.0f ns %12ld expansions/s=%.4gM nodes=%.0f solved=%.0f cost=%.4f\n", r.name.c_str(),
                    r.real_ns, r.cpu_ns, r.iterations, r.exp_per_s * 1e-6, r.nodes, r.solved, r.cost);
    }
    return 0;
  } catch (const krrt::Error& e) {
    std::printf("rrtstar_benchmark: %s\n", e.what());
    return e.code == KRRT_ERR_NO_DEVICE ? 3 : 1;
  }
}
