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

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

#include "krrt.h"

namespace krrt {

// KRRT_PRECISION=f32 in the environment: the component set plans in binary32 unless told otherwise (krrt.h: KRRT_F32)
inline int default_precision() {
  const char* e = std::getenv("KRRT_PRECISION");
  return (e && (e[0] == 'f' || e[0] == 'F') && e[1] == '3') ? KRRT_F32 : KRRT_F64;
}

struct Error : std::runtime_error {
  int code;
  Error(int c, const std::string& what) : std::runtime_error("krrt[" + std::to_string(c) + "]: " + what), code(c) {}
};

// One device context shared by the components of a planner (and by the static / dynamic planner pair, which
// share one tree in the reference: …node.cpp:60-61).  Created lazily so that constructing the singletons
// does not touch the GPU; re-created only when a create-time field (capacity, device) changes.
// Run-ahead of the caller's per-iteration loop.  The node calls rrt.grow(xg) once per iteration (…node.cpp:176-185) and
// reads tree.tree.size() and sampler.last_sample() after every call; one device call per iteration costs ~60 us of launch
// and read-back latency.  The planner is deterministic (counter-based samples), so the first grow(xg) of a cycle runs the
// iterations that are LIKELY to follow — as many as the cycle before had, with the goal forced as the sample where the
// caller forced it last time (…node.cpp:183-184) — in ONE logged device call (krrt_grow_logged), and the following
// calls are answered from its per-iteration records.  Every call is checked against what was assumed (same goal, same forced
// sample or none); on a mismatch, or when the tree is read while the device is ahead of the caller, the device tree is
// rebuilt at the caller's iteration by replaying the cycle from its start (same samples -> the same tree, bit for bit).
// What the caller sees is therefore exactly what one device call per iteration gives (tests/native/host_surface_check.cpp
// runs the node's loop against the CPU checker call by call).  A replay uses the settings that are live, so a cycle whose
// obstacles or parameters change after its first iteration stops running ahead (RunAhead::frozen) until the next setStart.
// KRRT_RUN_AHEAD=0 in the environment switches it off.
struct RunAhead {
  struct Forced { long at; double x[4]; };
  struct GoalSeg { long from; bool has; double g[4]; };
  bool enabled = true;
  std::vector<krrt_iter_record> log;  // record of iteration i of the current cycle (n_nodes < 0: not recorded)
  long served = 0;                    // iterations handed to the caller since setStart
  long ahead = 0;                     // iterations the device tree has run since setStart (>= served)
  double start[4] = {0, 0, 0, 0};
  std::vector<Forced> forced;         // explicit sample rows the device used in [0, ahead): the caller's and the assumed ones
  std::vector<GoalSeg> goals;         // goal handed to grow() from iteration `from` on
  long last_len = 0;                  // iterations of the cycle before
  std::vector<long> last_goal_forced_at, goal_forced_at;  // where the caller forced the goal as the sample (before / now)
  long replays = 0, device_calls = 0;  // diagnostics
  // The cycle's settings (discs, r, R, radius schedule, sampler) changed after its first iteration.  A replay rebuilds the
  // cycle from its start with the settings that are live NOW, so it would redo the earlier iterations with settings they
  // never saw.  For the rest of such a cycle nothing is run ahead — one iteration per device call, the device tree is
  // always the caller's tree and nothing is ever replayed.  The next setStart lifts it.
  bool frozen = false;

  const Forced* forced_at(long i) const {
    for (const Forced& f : forced)
      if (f.at == i) return &f;
    return nullptr;
  }
  const GoalSeg* goal_at(long i) const {
    const GoalSeg* r = nullptr;
    for (const GoalSeg& g : goals)
      if (g.from <= i) r = &g;
    return r;
  }
  static bool same4(const double* a, const double* b) { return a[0] == b[0] && a[1] == b[1] && a[2] == b[2] && a[3] == b[3]; }
  void note_goal(long i, bool has, const double* g) {
    const GoalSeg* cur = goal_at(i);
    if (cur && cur->has == has && (!has || same4(cur->g, g))) return;
    GoalSeg n{i, has, {0, 0, 0, 0}};
    if (has) std::memcpy(n.g, g, sizeof(n.g));
    goals.push_back(n);
  }
};

class Session {
 public:
  Session() {
    krrt_default_config(&cfg_);
    if (const char* e = std::getenv("KRRT_RUN_AHEAD")) ra_.enabled = std::atoi(e) != 0;
  }
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
      settings_change();  // iterations run ahead were run with the settings that are live now: settle the tree before they change
      check(krrt_set_params(ctx_, &cfg_));
      live_ = cfg_;
    }
    return ctx_;
  }
  RunAhead& run_ahead() { return ra_; }
  // a new cycle (rrt.setStart): remember how the last one went, forget its records
  void begin_cycle(const double x0[4]) {
    if (ra_.served > 0) {
      ra_.last_len = ra_.served;
      ra_.last_goal_forced_at = ra_.goal_forced_at;
    }
    ra_.log.clear();
    ra_.forced.clear();
    ra_.goals.clear();
    ra_.goal_forced_at.clear();
    ra_.served = ra_.ahead = 0;
    ra_.frozen = false;
    std::memcpy(ra_.start, x0, sizeof(ra_.start));
  }
  // about to change what the tree is grown with (krrt_set_params, krrt_set_obstacles): the device tree goes back to the
  // caller's iteration while the old settings are still live, and a cycle that is already under way stops running ahead
  void settings_change() {
    sync_tree();
    if (started && ra_.served > 0) ra_.frozen = true;
  }
  // the device tree at the caller's iteration: anything that reads the tree, or changes what it is grown with, calls this first
  void sync_tree() {
    if (ctx_ && started && ra_.ahead > ra_.served) rewind_to(ra_.served);
  }
  // rebuild the device tree as it stood after iteration i - 1 of the cycle: restart + replay (same samples, same tree)
  void rewind_to(long i) {
    check(krrt_set_start(ctx_, ra_.start));
    std::vector<double> rows;
    for (size_t k = 0; k < ra_.goals.size() && ra_.goals[k].from < i; ++k) {
      const RunAhead::GoalSeg& seg = ra_.goals[k];
      const long lo = seg.from, hi = (k + 1 < ra_.goals.size() && ra_.goals[k + 1].from < i) ? ra_.goals[k + 1].from : i;
      if (hi <= lo) continue;
      rows.assign((size_t)(hi - lo) * 4, std::nan(""));
      bool any = false;
      for (const RunAhead::Forced& f : ra_.forced)
        if (f.at >= lo && f.at < hi) {
          std::memcpy(&rows[(size_t)(f.at - lo) * 4], f.x, sizeof(f.x));
          any = true;
        }
      int solved = 0, gi = -1;
      check(krrt_grow(ctx_, seg.has ? seg.g : nullptr, (int)(hi - lo), any ? rows.data() : nullptr, &solved, &gi));
      ++ra_.device_calls;
    }
    ra_.ahead = i;
    if ((long)ra_.log.size() > i) ra_.log.resize((size_t)i);
    for (size_t k = ra_.forced.size(); k-- > 0;)
      if (ra_.forced[k].at >= i) ra_.forced.erase(ra_.forced.begin() + (long)k);
    while (!ra_.goals.empty() && ra_.goals.back().from >= i && i > 0) ra_.goals.pop_back();
    if (i == 0) ra_.goals.clear();
    ++ra_.replays;
    touch();
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
  RunAhead ra_;
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
    s.begin_cycle(x0);  // (first: whatever was run ahead of the last cycle is dropped, not replayed)
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

  // == solved = rrt.grow(xg) (…node.cpp:177): ONE exact RRT* expansion — answered from the run-ahead records when the
  // call is the one that was assumed (krrt::RunAhead above), by a device call otherwise
  bool grow(State* xg) {
    krrt::Session& s = tree_.session();
    if (!s.started) throw krrt::Error(KRRT_ERR_STATE, "grow before setStart");
    krrt::RunAhead& ra = s.run_ahead();
    if (!ra.enabled) return grow(xg, 1);
    push_config(s);
    krrt_ctx* c = s.ctx();
    if (env_.stale(s)) {
      s.sync_tree();  // (the iterations run ahead saw the discs that are resident now)
      env_.upload(s);
    }
    double g[4] = {0, 0, 0, 0};
    const bool has = xg != nullptr;
    if (has)
      for (int k = 0; k < 4; ++k) g[k] = (*xg)(k);
    double fr[4];
    const bool has_forced = sampler_.take_next_sample(fr);  // sampler.set_next_sample(x) (…node.cpp:184)
    const long i = ra.served;
    if (i < ra.ahead) {  // this iteration has been run ahead: was it run with what the caller hands over now?
      const krrt::RunAhead::GoalSeg* gs = ra.goal_at(i);
      const krrt::RunAhead::Forced* f = ra.forced_at(i);
      const bool ok = gs && gs->has == has && (!has || krrt::RunAhead::same4(gs->g, g)) && ((f != nullptr) == has_forced) &&
                      (!f || krrt::RunAhead::same4(f->x, fr)) && ra.log[(size_t)i].n_nodes >= 0;
      if (!ok) s.rewind_to(i);
    }
    if (i == ra.ahead) {  // run this iteration and the ones that are likely to follow
      ra.note_goal(i, has, g);
      const int n_now = i > 0 && ra.log[(size_t)i - 1].n_nodes >= 0 ? ra.log[(size_t)i - 1].n_nodes : (i > 0 ? (int)tree_.device_size() : 1);
      long L = ra.last_len > i ? ra.last_len - i : (i < 1 ? 1 : (i < 64 ? i : 64));
      if (L > 256) L = 256;
      if (n_now + L > 1024) L = 1024 - n_now;
      if (ra.frozen && L > 1) L = 1;  // (settings changed mid-cycle: no replay could be exact any more, so nothing is run ahead)
      std::vector<krrt_iter_record> recs;
      int solved = 0, gi = -1, rc = KRRT_ERR_UNSUPPORTED;
      if (L >= 1) {
        std::vector<double> rows((size_t)L * 4, std::nan(""));
        std::vector<krrt::RunAhead::Forced> assumed;
        if (has_forced) {
          krrt::RunAhead::Forced f{i, {fr[0], fr[1], fr[2], fr[3]}};
          assumed.push_back(f);
        }
        if (has)
          for (long p : ra.last_goal_forced_at)
            if (p > i && p < i + L) {
              krrt::RunAhead::Forced f{p, {g[0], g[1], g[2], g[3]}};
              assumed.push_back(f);
            }
        for (const auto& f : assumed) std::memcpy(&rows[(size_t)(f.at - i) * 4], f.x, sizeof(f.x));
        recs.resize((size_t)L);
        rc = krrt_grow_logged(c, has ? g : nullptr, (int)L, assumed.empty() ? nullptr : rows.data(), recs.data(), &solved, &gi);
        if (rc == KRRT_OK) ra.forced.insert(ra.forced.end(), assumed.begin(), assumed.end());
      }
      if (rc == KRRT_ERR_UNSUPPORTED) {  // a tree too large for per-iteration records: one plain device call for this iteration
        s.check(krrt_grow(c, has ? g : nullptr, 1, has_forced ? fr : nullptr, &solved, &gi));
        if (has_forced) {
          krrt::RunAhead::Forced f{i, {fr[0], fr[1], fr[2], fr[3]}};
          ra.forced.push_back(f);
        }
        krrt_iter_record r{};
        int n = 0;
        s.check(krrt_tree_size(c, &n));
        s.check(krrt_last_sample(c, r.sample));
        r.inserted = n > n_now ? 1 : 0;
        r.n_nodes = n;
        r.solved = solved;
        r.goal_index = gi;
        recs.assign(1, r);
        L = 1;
      } else {
        s.check(rc);
      }
      ++ra.device_calls;
      ra.log.resize((size_t)i);
      ra.log.insert(ra.log.end(), recs.begin(), recs.end());
      ra.ahead = i + L;
    }
    const krrt_iter_record& r = ra.log[(size_t)i];
    if (has_forced && has && krrt::RunAhead::same4(fr, g)) ra.goal_forced_at.push_back(i);
    ra.served = i + 1;
    ++iteration_;
    s.touch();
    goal_index_ = r.goal_index;
    return r.solved != 0;
  }
  // extension: n_iter expansions in one device call (identical tree to n_iter single calls)
  bool grow(State* xg, int n_iter) {
    krrt::Session& s = tree_.session();
    if (!s.started) throw krrt::Error(KRRT_ERR_STATE, "grow before setStart");
    push_config(s);
    krrt_ctx* c = s.ctx();
    s.sync_tree();
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
    {  // the cycle's history, should a replay be needed later: this call's goal and explicit first sample
      krrt::RunAhead& ra = s.run_ahead();
      ra.note_goal(ra.served, xg != nullptr, g);
      if (samples) {
        krrt::RunAhead::Forced f{ra.served, {forced[0], forced[1], forced[2], forced[3]}};
        ra.forced.push_back(f);
      }
      krrt_iter_record none{};
      none.n_nodes = -1;  // no per-iteration record: the getters ask the device
      ra.log.resize((size_t)ra.served);
      ra.log.resize((size_t)(ra.served + n_iter), none);
      ra.served += n_iter;
      ra.ahead = ra.served;
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
