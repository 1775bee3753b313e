// rrtstar.hpp — host-side C++17 mirror of the reference's RRT* driver template
//
//   RRTStar<Tree, Cost, Sampler, NeighborRadius, Environment, GoalChecker, Connector>
//
// (template slots as named by BASELINE.json north_star; the reference's own header of this name lives in
// the absent submodule src/motion_planning, so the members below are exactly the ones its in-tree callers
// use — file:line under /root/reference/src/robosoccer_motion_planning/):
//   rrt.setStart(xs); rrt.setIteration(0);     src/robosoccer_motion_planning_node.cpp:135-136
//   solved = rrt.grow(xg);                     src/robosoccer_motion_planning_node.cpp:177
//   rrt.goalIndex()  (> 0 <=> solved)          include/robosoccer_visual.hpp:19,40,45 ; …node.cpp:274,284
//
// The driver does no arithmetic of its own: every expansion runs in the sm_100a kernels of libkrrt.so
// through the C ABI of krrt.h.  There is no CPU implementation behind this template; a component set that
// is not device-backed does not compile (static_assert), and a missing GPU surfaces as krrt::Error.
#ifndef KRRT_RRTSTAR_HPP
#define KRRT_RRTSTAR_HPP

#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <type_traits>

#include "krrt.h"

namespace krrt {

struct Error : std::runtime_error {
  int code;
  Error(int c, const std::string& what) : std::runtime_error("krrt[" + std::to_string(c) + "]: " + what), code(c) {}
};

// One device context shared by the components of a planner (and by the static / dynamic planner pair, which
// share one tree in the reference: …node.cpp:60-61).  Created lazily so that constructing the singletons
// does not touch the GPU; re-created only when a create-time field (capacity, device) changes.
class Session {
 public:
  Session() { krrt_default_config(&cfg_); }
  ~Session() { krrt_destroy(ctx_); }
  Session(const Session&) = delete;
  Session& operator=(const Session&) = delete;

  krrt_config& config() { return cfg_; }
  const krrt_config& config() const { return cfg_; }

  // push the current config: new ctx when needed, krrt_set_params otherwise
  krrt_ctx* ctx() {
    if (ctx_ && (cfg_.capacity != live_.capacity || cfg_.device != live_.device || cfg_.batch != live_.batch ||
                 cfg_.near_cap != live_.near_cap || cfg_.precision != live_.precision)) {
      krrt_destroy(ctx_);
      ctx_ = nullptr;
    }
    if (!ctx_) {
      int rc = krrt_create(&cfg_, &ctx_);
      if (rc != KRRT_OK) {
        ctx_ = nullptr;
        throw Error(rc, krrt_last_error(nullptr));
      }
      live_ = cfg_;
      ++epoch_;
    } else if (std::memcmp(&cfg_, &live_, sizeof(cfg_)) != 0) {
      check(krrt_set_params(ctx_, &cfg_));
      live_ = cfg_;
    }
    return ctx_;
  }
  void check(int rc) const {
    if (rc != KRRT_OK) throw Error(rc, krrt_last_error(ctx_));
  }
  // bumped whenever the device tree may have changed; host mirrors compare against it
  unsigned long version() const { return version_; }
  void touch() { ++version_; }
  bool started = false;
  const void* obstacle_owner = nullptr;  // which Environment's discs are resident on the device
  unsigned long obstacle_version = 0;

 private:
  krrt_config cfg_{}, live_{};
  krrt_ctx* ctx_ = nullptr;
  unsigned long version_ = 1, epoch_ = 0;
};

template <typename T, typename = void>
struct is_device_backed : std::false_type {};
template <typename T>
struct is_device_backed<T, std::void_t<decltype(T::device_backed)>> : std::bool_constant<T::device_backed> {};

}  // namespace krrt

template <typename Tree, typename Cost, typename Sampler, typename NeighborRadius, typename Environment,
          typename GoalChecker, typename Connector>
class RRTStar {
  static_assert(krrt::is_device_backed<Tree>::value && krrt::is_device_backed<Cost>::value &&
                    krrt::is_device_backed<Sampler>::value && krrt::is_device_backed<NeighborRadius>::value &&
                    krrt::is_device_backed<Environment>::value && krrt::is_device_backed<GoalChecker>::value &&
                    krrt::is_device_backed<Connector>::value,
                "RRTStar: every component must be one of the CUDA-backed set (no CPU path exists)");

 public:
  using State = typename Tree::State;

  RRTStar(Tree& tree, Cost& cost, Sampler& sampler, NeighborRadius& radius, Environment& env, GoalChecker& goal,
          Connector& connector)
      : tree_(tree), cost_(cost), sampler_(sampler), radius_(radius), env_(env), goal_(goal), connector_(connector) {}

  // == rrt.setStart(xs): fresh tree holding only the root (…node.cpp:135)
  void setStart(const State& xs) {
    krrt::Session& s = tree_.session();
    push_config(s);
    const double x0[4] = {xs(0), xs(1), xs(2), xs(3)};
    s.check(krrt_set_start(s.ctx(), x0));
    s.started = true;
    s.touch();
    goal_index_ = -1;
    iteration_ = 0;
  }
  // == rrt.setIteration(0): restart of the shrinking-ball schedule and of the sample counter (…node.cpp:136).
  // The device keeps both in the tree's control block, which setStart has just reset; other values are rejected.
  void setIteration(int i) {
    if (i != 0) throw krrt::Error(KRRT_ERR_UNSUPPORTED, "setIteration(i != 0): only the restart the caller uses is supported");
    iteration_ = 0;
  }

  // == solved = rrt.grow(xg) (…node.cpp:177): ONE exact RRT* expansion
  bool grow(State* xg) { return grow(xg, 1); }
  // extension: n_iter expansions in one device call (identical tree to n_iter single calls)
  bool grow(State* xg, int n_iter) {
    krrt::Session& s = tree_.session();
    if (!s.started) throw krrt::Error(KRRT_ERR_STATE, "grow before setStart");
    push_config(s);
    krrt_ctx* c = s.ctx();
    env_.upload(s);
    double g[4] = {0, 0, 0, 0};
    if (xg) {
      for (int k = 0; k < 4; ++k) g[k] = (*xg)(k);
    }
    double forced[4];
    const double* samples = nullptr;
    if (n_iter >= 1 && sampler_.take_next_sample(forced)) samples = forced;  // sampler.set_next_sample(x) (…node.cpp:184)
    int solved = 0, gi = -1;
    if (samples && n_iter > 1) {
      s.check(krrt_grow(c, xg ? g : nullptr, 1, samples, &solved, &gi));
      s.check(krrt_grow(c, xg ? g : nullptr, n_iter - 1, nullptr, &solved, &gi));
    } else {
      s.check(krrt_grow(c, xg ? g : nullptr, n_iter, samples, &solved, &gi));
    }
    {  // the iterations the device really ran (a tree that reached `capacity` stops early: full() tells)
      int64_t done = 0;
      int full = 0;
      s.check(krrt_grow_done(c, &done, &full));
      iteration_ += (samples && n_iter > 1) ? done + 1 : done;
      full_ = full != 0;
    }
    s.touch();
    goal_index_ = gi;
    return solved != 0;
  }
  // the tree holds Tree::capacity nodes: the last grow call stopped before its iterations were used up
  bool full() const { return full_; }

  // == rrt.goalIndex(): index of the goal node, > 0 iff a solution exists (robosoccer_visual.hpp:19)
  int goalIndex() const { return goal_index_; }
  long iteration() const { return iteration_; }

 private:
  void push_config(krrt::Session& s) {
    krrt_config& c = s.config();
    tree_.configure(c);
    cost_.configure(c);
    connector_.configure(c);
    sampler_.configure(c);
    radius_.configure(c);
    env_.configure(c);
    goal_.configure(c);
  }

  Tree& tree_;
  Cost& cost_;
  Sampler& sampler_;
  NeighborRadius& radius_;
  Environment& env_;
  GoalChecker& goal_;
  Connector& connector_;
  int goal_index_ = -1;
  long iteration_ = 0;
  bool full_ = false;
};

#endif  // KRRT_RRTSTAR_HPP
