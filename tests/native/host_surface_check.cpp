// host_surface_check.cpp — a ROS-free caller written the way the reference's node uses the planner
// (/root/reference/src/robosoccer_motion_planning/src/robosoccer_motion_planning_node.cpp:43-61 singletons,
// :118-141 set_start / set_obstacles, :167-196 solve_rrt loop with one grow(xg) per iteration and the forced goal
// samples, :232-249 get_trajectory; include/robosoccer_visual.hpp:13-48 tree read-out), compiled by g++ -std=c++17
// against include/rrtstar.hpp + include/integrator2drrt.hpp and linked with libkrrt.so.
//
// TEST ONLY: also links oracle/_build/libkrrt_oracle.so and checks the tree grown through the C++ surface
// bit-for-bit against the CPU oracle fed the same calls.  Usage: host_surface_check [static|dynamic] [iterations]
// Exit 0 = identical; 3 = no GPU (the surface threw krrt::Error with KRRT_ERR_NO_DEVICE, as it must).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <type_traits>
#include <vector>

#include "../../include/integrator2drrt.hpp"
#include "../../oracle/krrt_oracle.h"

struct ObState : std::array<double, 4> {  // shaped like RobotSubscriber::State (robotsubscriber.h:15-25)
  const double& operator()(size_t i) const { return (*this)[i]; }
};

static int fails = 0;
#define EXPECT(cond)                                                      \
  do {                                                                    \
    if (!(cond)) {                                                        \
      if (fails < 20) std::printf("FAIL %s:%d %s\n", __FILE__, __LINE__, #cond); \
      ++fails;                                                            \
    }                                                                     \
  } while (0)

template <typename RRT, typename Env, typename Tree, typename State, typename Sampler>
static bool solve_rrt(RRT& rrt, Env& env, Tree& tree, const std::vector<ObState>& obstacles, const State& start,
                      State* xg, Sampler& sampler, size_t iteration, bool direct_sampling_en, ko_planner* oracle,
                      int* oracle_goal) {
  Kinodynamic::TreeInt2D::State xs;
  xs(0) = start(0); xs(1) = start(1); xs(2) = start(2); xs(3) = start(3);
  rrt.setStart(xs);
  rrt.setIteration(0);
  env.setObstacles(obstacles);
  ko_set_obstacles(oracle, obstacles[0].data(), (int)obstacles.size(), Env::dynamic ? 1 : 0);
  ko_set_start(oracle, xs.data());
  bool solved = false;
  auto ts = tree.tree.size();
  bool forced = false;
  for (size_t i = 0; i < iteration; i++) {
    solved = rrt.grow(xg);
    // the oracle gets the same call: an explicit sample when the node forced one, else its own Philox draw
    double nan_row[4] = {std::nan(""), 0, 0, 0};
    const double* row = forced ? xg->data() : nan_row;
    int o_solved = ko_grow(oracle, xg->data(), 1, row, oracle_goal);
    forced = false;
    auto s = sampler.last_sample();
    const double* os = ko_last_sample(oracle);
    EXPECT(s(0) == os[0] && s(1) == os[1] && s(2) == os[2] && s(3) == os[3]);
    EXPECT((tree.tree.size() > ts) == (ko_size(oracle) > (int)ts));
    EXPECT(solved == (o_solved != 0));
    ts = tree.tree.size();
    if (direct_sampling_en && ((i == (iteration - 2)) || (i == (iteration / 2)))) {
      sampler.set_next_sample(*xg);
      forced = true;
    }
  }
  return solved;
}

int main(int argc, char** argv) {
  const bool dynamic = argc > 1 && std::string(argv[1]) == "dynamic";
  const size_t target_size = argc > 2 ? (size_t)std::atoi(argv[2]) : 100;
  try {
    Models::init_integrator2d();
    auto& rrt = Kinodynamic::Wrapper::get_rrtstar_int2d();
    auto& env = Kinodynamic::Wrapper::get_robosoccer_env();
    auto& goal = Kinodynamic::Wrapper::get_goal();
    auto& sampler = Kinodynamic::Wrapper::get_sampler();
    auto& rrt_dyn = Kinodynamic::Wrapper::get_rrtstar_int2d_timespace_obs();
    auto& env_dyn = Kinodynamic::Wrapper::get_dynamic_soccer_env();
    auto& goal_dyn = Kinodynamic::Wrapper::get_goal_dynamic_env();
    auto& sampler_dyn = Kinodynamic::Wrapper::get_sampler_dynamic_env();
    auto& solver = Kinodynamic::Wrapper::get_trajectory_solver();
    auto& tree = Kinodynamic::Wrapper::get_tree_int2d();
    (void)goal_dyn;
    static_assert(!std::is_same<std::decay_t<decltype(env)>, std::decay_t<decltype(env_dyn)>>::value,
                  "robosoccer_visual.hpp dispatches on the environment type");

    const double cr = 1.35, r = 0.5, ds_prob = 0.1;
    env.collision_radius = cr;
    env_dyn.collision_radius = cr;
    solver.set_weight(r);
    tree.capacity = (int)target_size + 2;

    // nine robots at their spawn poses (robot_up.sh:11-14); moving ones for the dynamic environment
    const double sp[9][2] = {{-2, 1}, {-2, -1}, {-5, 2}, {-5, -2}, {8.5, 0}, {2, 1}, {2, -1}, {5, 2}, {5, -2}};
    std::vector<ObState> obstacles(9);
    for (int j = 0; j < 9; ++j) {
      obstacles[j][0] = sp[j][0]; obstacles[j][1] = sp[j][1];
      obstacles[j][2] = dynamic ? 0.1 * (j - 4) : 0.0;
      obstacles[j][3] = dynamic ? 0.05 * (4 - j) : 0.0;
    }
    env.setObstacles(obstacles);
    env_dyn.setObstacles(obstacles);

    auto xg = goal.randomGoal();
    xg(0) = 7.0; xg(1) = 0.0; xg(2) = 0.0; xg(3) = 0.0;  // goal_subs.getGoal(xg): set_xyz writes x, y and vx = z = 0
    sampler.set_direct_sample(true, ds_prob);
    sampler_dyn.set_direct_sample(true, ds_prob);
    sampler.target = xg;
    sampler_dyn.target = xg;
    EXPECT(sampler_dyn.direct_sampler->p[0] == ds_prob);

    ko_config oc;
    ko_default_config(&oc);
    oc.r = r; oc.R = cr; oc.p_direct = ds_prob; oc.capacity = (int)target_size + 2;
    ko_planner* oracle = ko_create(&oc);
    int ogi = -1;

    Kinodynamic::TreeInt2D::State start;
    start(0) = -8.5; start(1) = 0.0; start(2) = 0.0; start(3) = 0.0;
    bool solved;
    int gi;
    if (dynamic) {
      solved = solve_rrt(rrt_dyn, env_dyn, tree, obstacles, start, &xg, sampler_dyn, target_size, true, oracle, &ogi);
      gi = rrt_dyn.goalIndex();
    } else {
      solved = solve_rrt(rrt, env, tree, obstacles, start, &xg, sampler, target_size, true, oracle, &ogi);
      gi = rrt.goalIndex();
    }
    EXPECT(gi == ogi);
    EXPECT(solved == (gi > 0));

    // robosoccer_visual.hpp:18: tree.tree.cloud.states, tree.trajectories, tree.parent, tree.tree.size()
    const size_t n = tree.tree.size();
    EXPECT((int)n == ko_size(oracle));
    const std::vector<Kinodynamic::State4>& cloud = tree.tree.cloud.states;
    const std::vector<int>& par = tree.parent;
    // the mirrors carry the connected goal node behind the tree while rrt.goalIndex() == tree.tree.size()
    // (robosoccer_visual.hpp:45: tree(rrt.goalIndex()))
    const size_t n_all = n + ((size_t)gi == n ? 1 : 0);
    EXPECT(cloud.size() == n_all && par.size() == n_all && tree.trajectories.size() == n_all);
    if (gi > 0) EXPECT(std::memcmp(tree((size_t)gi).data(), xg.data(), 32) == 0);
    for (size_t i = 0; i < n_all && i < cloud.size(); ++i) {
      EXPECT(std::memcmp(cloud[i].data(), ko_states(oracle) + 4 * i, 32) == 0);
      EXPECT(par[i] == ko_parents(oracle)[i]);
      EXPECT(tree.cost()[i] == ko_ctc(oracle)[i] && tree.time()[i] == ko_times(oracle)[i]);
      if (i > 0) {  // stored edge = S samples of the optimal connection parent -> i, time since the root
        const auto& e = tree.trajectories[i];
        const double* a = ko_states(oracle) + 4 * par[i];
        const double* b = ko_states(oracle) + 4 * i;
        double tau, c, st[4];
        ko_steer(a, b, r, &tau, &c);
        EXPECT((int)e.size() == tree.edge_samples);
        for (int k = 0; k < (int)e.size(); ++k) {
          const double h = (double)k / (double)(e.size() - 1);
          ko_edge_state(a, b, tau, h, st);
          EXPECT(e[k](0) == st[0] && e[k](1) == st[1] && e[k](2) == st[2] && e[k](3) == st[3]);
          EXPECT(e[k](4) == ko_times(oracle)[par[i]] + tau * h);
        }
        EXPECT(e.path().size() == e.size());
      }
    }
    auto xs0 = tree(0);
    EXPECT(xs0(0) == start(0) && xs0(1) == start(1));
    auto obs_t = env_dyn.at(2.0);
    EXPECT(obs_t.size() == 9 && obs_t[0][0] == std::fma(obstacles[0][2], 2.0, obstacles[0][0]));
    for (const auto& o : env.obs) EXPECT(std::get<0>(o) == o[0] && std::get<1>(o) == o[1]);

    if (solved) {
      // get_trajectory (…node.cpp:232-249)
      auto states = tree.states(gi);
      std::vector<double> ost(4 * n), otm(n);
      int olen = ko_path(oracle, ogi, ost.data(), otm.data(), (int)n);
      EXPECT((int)states.size() == olen);
      auto dt = 0.1;
      Kinodynamic::VTrajectory ret;
      std::vector<double> rows(5 * 100000);
      size_t expect_rows = 0;
      for (size_t i = 1; i < states.size(); i++) {
        auto tc = solver.solver.cost(states[i - 1], states[i]);
        auto t = std::get<0>(tc);
        auto c = std::get<1>(tc);
        double ot, ocst;
        ko_steer(ost.data() + 4 * (i - 1), ost.data() + 4 * i, r, &ot, &ocst);
        EXPECT(t == ot && c == ocst);
        auto trajectory = solver.solver.solve(states[i - 1], states[i], dt);
        int nr = ko_solve(ost.data() + 4 * (i - 1), ost.data() + 4 * i, r, dt, 0.0, rows.data(), 100000);
        EXPECT((int)trajectory.size() == nr);
        for (int k = 0; k < nr && k < (int)trajectory.size(); ++k)
          EXPECT(std::memcmp(trajectory[k].data(), rows.data() + 5 * k, 40) == 0);
        ret.insert(trajectory);
        expect_rows += nr;
      }
      EXPECT(ret.size() == expect_rows);
      for (size_t k = 1; k < ret.size(); ++k) EXPECT(ret.at(k)(4) >= ret.at(k - 1)(4));
      auto goal_trj = tree.get_trajectory(gi);
      EXPECT(goal_trj.size() + 1 == states.size());
    }
    // one multi-expansion call gives the same tree as single calls
    {
      auto& rr = Kinodynamic::Wrapper::get_rrtstar_int2d();
      rr.setStart(start);
      rr.setIteration(0);
      rr.grow(&xg, (int)target_size / 2);
      ko_set_obstacles(oracle, obstacles[0].data(), 9, 0);
      ko_set_start(oracle, start.data());
      ko_grow(oracle, xg.data(), (int)target_size / 2, nullptr, &ogi);
      EXPECT((int)tree.tree.size() == ko_size(oracle) && rr.goalIndex() == ogi);
      const std::vector<int>& p2 = tree.parent;
      for (size_t i = 0; i < p2.size(); ++i) EXPECT(p2[i] == ko_parents(oracle)[i]);
    }
    // The run-ahead of the per-iteration loop (krrt::RunAhead, include/rrtstar.hpp): whatever the caller does — the same
    // cycle again, a longer one, a shorter one, a tree read-out in the middle of the loop, another goal from some
    // iteration on, a forced sample where none was the cycle before — every single call answers as the CPU checker does.
    if (target_size >= 60) {
      auto& rr = Kinodynamic::Wrapper::get_rrtstar_int2d();
      krrt::RunAhead& ra = Kinodynamic::Wrapper::get_session().run_ahead();
      // (change_at: the obstacles move at that iteration of the cycle — env.setObstacles in the middle of the loop)
      std::vector<ObState> moved = obstacles;
      for (auto& o : moved) { o[0] += 0.6; o[1] -= 0.9; }
      auto run_cycle = [&](int iters, int read_at, int goal_change_at, int force_at, int change_at = -1) {
        rr.setStart(start);
        rr.setIteration(0);
        env.setObstacles(obstacles);
        ko_set_obstacles(oracle, obstacles[0].data(), 9, 0);
        ko_set_start(oracle, start.data());
        auto g = xg;
        bool forced = false;
        Kinodynamic::TreeInt2D::State frow = g;
        for (int i = 0; i < iters; ++i) {
          if (i == goal_change_at) {
            g(0) = 6.0;
            g(1) = 2.5;
          }
          if (i == change_at) {
            env.setObstacles(moved);
            ko_set_obstacles(oracle, moved[0].data(), 9, 0);
          }
          if (i == force_at) {
            Kinodynamic::TreeInt2D::State x;
            x(0) = 1.0; x(1) = -2.0; x(2) = 0.5; x(3) = 0.25;
            sampler.set_next_sample(x);
            forced = true;
            frow = x;
          }
          const bool sv = rr.grow(&g);
          const double nan_row[4] = {std::nan(""), 0, 0, 0};
          const int osv = ko_grow(oracle, g.data(), 1, forced ? frow.data() : nan_row, &ogi);
          forced = false;
          EXPECT(sv == (osv != 0));
          EXPECT(rr.goalIndex() == ogi);
          const auto ls = sampler.last_sample();
          const double* os = ko_last_sample(oracle);
          EXPECT(ls(0) == os[0] && ls(1) == os[1] && ls(2) == os[2] && ls(3) == os[3]);
          EXPECT((int)tree.tree.size() == ko_size(oracle));
          if (i == read_at) {
            const std::vector<int>& pp = tree.parent;
            for (size_t k = 0; k < (size_t)ko_size(oracle) && k < pp.size(); ++k) EXPECT(pp[k] == ko_parents(oracle)[k]);
          }
          if (i == iters - 2 || i == iters / 2) {
            sampler.set_next_sample(g);
            forced = true;
            frow = g;
          }
        }
        double dummy[4];
        sampler.take_next_sample(dummy);  // (a sample forced behind the last iteration is not carried into the next cycle)
        const std::vector<int>& pp = tree.parent;
        EXPECT((int)tree.tree.size() == ko_size(oracle));
        for (size_t k = 0; k < (size_t)ko_size(oracle) && k < pp.size(); ++k) {
          EXPECT(pp[k] == ko_parents(oracle)[k]);
          EXPECT(tree.cost()[k] == ko_ctc(oracle)[k]);
        }
      };
      const int N = (int)target_size - 13;
      run_cycle(N, -1, -1, -1);
      const long c0 = ra.device_calls, r0 = ra.replays;
      run_cycle(N, -1, -1, -1);
      // the steady state: ONE device call per cycle (per 256 iterations: what one logged call records), nothing replayed
      if (ra.enabled) EXPECT(ra.device_calls - c0 == (N + 255) / 256 && ra.replays == r0);
      run_cycle((int)target_size, 7, 20, 31);
      run_cycle(N / 2, -1, -1, -1);
      run_cycle(N / 2, 3, -1, 5);
      // obstacles that change in the middle of a cycle: the iterations before the change keep the tree they built with the old
      // discs whatever happens later — a tree read-out, a forced sample nobody assumed, another goal (each used to replay the
      // WHOLE cycle with the new discs)
      run_cycle(N, 45, 50, 55, 30);  // (the CPU checker's tree after 45 iterations with the moved discs throughout differs: 26 nodes against 20)
      if (ra.enabled) EXPECT(ra.frozen);
      const long c1 = ra.device_calls, r1 = ra.replays;
      run_cycle(N, -1, -1, -1);           // the next cycle runs ahead again
      if (ra.enabled) EXPECT(!ra.frozen && ra.replays - r1 <= 2 && ra.device_calls - c1 <= 8);
      run_cycle(N, 5, -1, -1, 1);
      run_cycle(N, -1, -1, -1);
      std::printf("run-ahead: %ld device calls, %ld replays over 9 cycles\n", ra.device_calls, ra.replays);
    }
    ko_destroy(oracle);
    std::printf("%s: %s, %zu nodes, goalIndex %d, %d check failure(s)\n", dynamic ? "dynamic" : "static",
                solved ? "solved" : "not solved", n, gi, fails);
    return fails ? 1 : 0;
  } catch (const krrt::Error& e) {
    std::printf("krrt::Error %d: %s\n", e.code, e.what());
    return e.code == KRRT_ERR_NO_DEVICE ? 3 : 2;
  }
}
