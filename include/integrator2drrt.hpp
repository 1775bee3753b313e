// integrator2drrt.hpp — the CUDA-backed component set of the 2-D double-integrator kinodynamic RRT*
// (namespaces Kinodynamic::, Kinodynamic::Wrapper::, Models:: as the reference's callers spell them).
//
// The reference's header of this name is part of the absent submodule src/motion_planning; what is mirrored
// here is the surface its in-tree callers use (file:line under /root/reference/src/robosoccer_motion_planning/):
//   Kinodynamic::TreeInt2D::State xs; xs(0)=…            src/robosoccer_motion_planning_node.cpp:131-133
//   tree.tree.size()                                      …node.cpp:175,179   (tree nodes; the connected goal node is
//                                                         entry size() of the mirrors below while rrt.goalIndex() == size())
//   tree.tree.cloud.states, tree.trajectories, tree.parent include/robosoccer_visual.hpp:18
//   tree(i), tree.get_trajectory(idx) -> edges with .path() and (x,y,vx,vy,t) samples  robosoccer_visual.hpp:20-22,37-45
//   tree.states(idx)                                      …node.cpp:237
//   solver.set_weight(r); solver.solver.cost(a,b) -> (t,c); solver.solver.solve(a,b,dt)   …node.cpp:91-92,242-245
//   Kinodynamic::VTrajectory ret; ret.insert(trajectory)  …node.cpp:239-246
//   sampler.last_sample(), set_next_sample(x), set_direct_sample(bool,p), .target, .direct_sampler->p[0]   …node.cpp:178,184,225-229
//   env.collision_radius, env.setObstacles(v), env.obs, env.at(t)   …node.cpp:84-85,140 ; robosoccer_visual.hpp:15-17,35
//   goal.randomGoal()                                     …node.cpp:221
//   Kinodynamic::Wrapper::get_*()                         …node.cpp:46-61
//   Models::init_integrator2d()                           …node.cpp:43
// Every numeric result comes from libkrrt.so (krrt.h); nothing here computes steering, distances or collisions
// on the host.
#ifndef KRRT_INTEGRATOR2DRRT_HPP
#define KRRT_INTEGRATOR2DRRT_HPP

#include <array>
#include <cmath>
#include <cstdint>
#include <memory>
#include <random>
#include <tuple>
#include <vector>

#include "rrtstar.hpp"

namespace Kinodynamic {

// s(i) and s[i] both work (elements.hpp:15-29 uses [i], the node (i)).  On an rvalue s(i) is a VALUE: the caller forms
// std::array<decltype(std::declval<State>()(0)), 2> (…node.cpp:300-301), which must be an array of double, not of double&.
template <size_t N>
struct StateN : std::array<double, N> {
  const double& operator()(size_t i) const& { return (*this)[i]; }
  double& operator()(size_t i) & { return (*this)[i]; }
  double operator()(size_t i) && { return (*this)[i]; }
};
using State4 = StateN<4>;     // (x, y, vx, vy)
using StateTime = StateN<5>;  // (x, y, vx, vy, t)

// one tree edge as the Connector stores it: S samples of (x,y,vx,vy,t), both end points included
struct Edge : std::vector<StateTime> {
  std::vector<std::array<double, 2>> path() const {
    std::vector<std::array<double, 2>> p;
    p.reserve(size());
    for (const auto& s : *this) p.push_back({s[0], s[1]});
    return p;
  }
};

// flat (x,y,vx,vy,t) trajectory; insert(rows) appends rows with their time shifted so that index 4 stays
// "time since the root" over the concatenation (…node.cpp:316-328 reads it as an absolute offset)
struct VTrajectory : std::vector<StateTime> {
  using std::vector<StateTime>::insert;
  template <typename Rows>
  void insert(const Rows& rows) {
    const double t0 = empty() ? 0.0 : back()[4];
    for (const auto& r : rows) {
      StateTime s;
      for (int k = 0; k < 4; ++k) s[k] = r[k];
      s[4] = t0 + r[4];
      push_back(s);
    }
  }
};

// host mirror of a device array, refreshed on first access after the tree changed
template <typename T, typename Owner>
class Mirror {
 public:
  explicit Mirror(Owner* o) : owner_(o) {}
  const std::vector<T>& get() const {
    owner_->refresh();
    return v_;
  }
  operator const std::vector<T>&() const { return get(); }
  auto begin() const { return get().begin(); }
  auto end() const { return get().end(); }
  size_t size() const { return get().size(); }
  bool empty() const { return get().empty(); }
  const T& operator[](size_t i) const { return get()[i]; }
  const T& at(size_t i) const { return get().at(i); }
  std::vector<T>& raw() { return v_; }

 private:
  Owner* owner_;
  std::vector<T> v_;
};

// ---------------------------------------------------------------------------------------------- solver
// == Kinodynamic::Wrapper::get_trajectory_solver(): .set_weight(r), .solver.cost(), .solver.solve()
class TrajectorySolver {
 public:
  explicit TrajectorySolver(krrt::Session* s) : solver{s, this} {}
  void set_weight(double r) { weight = r; }
  double weight = 0.5;  // config/config.yaml:1

  struct Solver {
    krrt::Session* session;
    TrajectorySolver* owner;
    // (t*, c*) of the optimal connection xi -> xf  (…node.cpp:242-244)
    template <typename A, typename B>
    std::tuple<double, double> cost(const A& xi, const B& xf) const {
      session->config().r = owner->weight;
      const double a[4] = {xi[0], xi[1], xi[2], xi[3]}, b[4] = {xf[0], xf[1], xf[2], xf[3]};
      double t = 0, c = 0;
      session->check(krrt_steer_cost(session->ctx(), a, b, 1, &t, &c));
      return {t, c};
    }
    // rows (x,y,vx,vy,t) at t = 0, dt, 2dt, … < t*, then t*  (…node.cpp:245)
    template <typename A, typename B>
    std::vector<StateTime> solve(const A& xi, const B& xf, double dt) const {
      session->config().r = owner->weight;
      const double a[4] = {xi[0], xi[1], xi[2], xi[3]}, b[4] = {xf[0], xf[1], xf[2], xf[3]};
      int n = 0;
      krrt_ctx* c = session->ctx();
      session->check(krrt_solve(c, a, b, dt, 0.0, nullptr, 0, &n));
      std::vector<StateTime> rows(n);
      if (n > 0) session->check(krrt_solve(c, a, b, dt, 0.0, rows[0].data(), n, &n));
      return rows;
    }
  } solver;
};

// ------------------------------------------------------------------------------------------------ tree
class TreeInt2D {
 public:
  static constexpr bool device_backed = true;
  using State = State4;

  explicit TreeInt2D(krrt::Session* s) : tree(this), trajectories(this), parent(this), session_(s) {}
  krrt::Session& session() { return *session_; }
  void configure(krrt_config& c) const {
    c.capacity = capacity;
    c.S = edge_samples;
    c.precision = precision;
  }
  // KRRT_F32: the whole planner in binary32 (krrt.h); the reference is double throughout.  (Environment KRRT_PRECISION=f32
  // makes it the default: the reference's unmodified node can be run either way.)
  int precision = krrt::default_precision();
  int capacity = 1 << 16;
  int edge_samples = 10;

  struct Cloud {
    explicit Cloud(TreeInt2D* t) : states(t) {}
    Mirror<State, TreeInt2D> states;
  };
  struct Index {
    explicit Index(TreeInt2D* t) : cloud(t), owner(t) {}
    Cloud cloud;
    TreeInt2D* owner;
    size_t size() const { return owner->node_count(); }
  } tree;
  Mirror<Edge, TreeInt2D> trajectories;  // trajectories[i] = edge parent[i] -> i  (entry 0: the root, S copies)
  Mirror<int, TreeInt2D> parent;

  size_t node_count() const {
    if (!session_->started) return 0;
    // tree.tree.size() right after a grow(xg) that was answered from the run-ahead records: the record has it (rrtstar.hpp)
    const krrt::RunAhead& ra = session_->run_ahead();
    if (ra.served > 0 && (size_t)ra.served <= ra.log.size() && ra.log[(size_t)ra.served - 1].n_nodes >= 0)
      return (size_t)ra.log[(size_t)ra.served - 1].n_nodes;
    return device_size();
  }
  // the device's own count, with the device tree settled at the caller's iteration
  size_t device_size() const {
    if (!session_->started) return 0;
    krrt_ctx* c = session_->ctx();
    session_->sync_tree();
    int n = 0;
    session_->check(krrt_tree_size(c, &n));
    return (size_t)n;
  }
  State operator()(size_t i) const { return tree.cloud.states.at(i); }
  const std::vector<double>& cost() const {
    refresh();
    return cost_;
  }
  const std::vector<double>& time() const {
    refresh();
    return time_;
  }
  // root -> idx chain of states (…node.cpp:237)
  std::vector<State> states(int idx) const {
    std::vector<State> out;
    if (idx < 0 || !session_->started) return out;
    int len = 0;
    krrt_ctx* c = session_->ctx();
    session_->sync_tree();
    session_->check(krrt_path(c, idx, nullptr, nullptr, 0, &len));
    out.resize(len);
    if (len > 0) session_->check(krrt_path(c, idx, out[0].data(), nullptr, len, &len));
    return out;
  }
  // root -> idx chain of stored edges (robosoccer_visual.hpp:20,37)
  std::vector<Edge> get_trajectory(int idx) const {
    std::vector<Edge> out;
    const auto& par = parent.get();
    if (idx < 0 || (size_t)idx >= par.size()) return out;
    for (int i = idx; i > 0; i = par[i]) out.push_back(trajectories[i]);
    return std::vector<Edge>(out.rbegin(), out.rend());
  }

  void refresh() const {
    if (seen_ == session_->version() || !session_->started) return;
    auto* self = const_cast<TreeInt2D*>(this);
    krrt_ctx* c = session_->ctx();
    session_->sync_tree();  // (iterations run ahead of the caller are taken back first: rrtstar.hpp)
    int n = 0, gi = -1;
    session_->check(krrt_tree_size(c, &n));
    session_->check(krrt_goal_index(c, &gi));
    if (gi == n) ++n;  // the goal node connected behind the tree: tree(goalIndex()), get_trajectory(goalIndex()), parent[goalIndex()]
    auto& st = self->tree.cloud.states.raw();
    auto& pa = self->parent.raw();
    st.resize(n);
    pa.resize(n);
    self->cost_.resize(n);
    self->time_.resize(n);
    const int S = edge_samples;
    std::vector<double> rows((size_t)n * S * 5);
    session_->check(krrt_tree_read(c, 0, n, st[0].data(), pa.data(), self->cost_.data(), self->time_.data()));
    session_->check(krrt_tree_edges(c, S, rows.data(), n));
    auto& tr = self->trajectories.raw();
    tr.assign(n, Edge());
    for (int i = 0; i < n; ++i) {
      tr[i].resize(S);
      for (int k = 0; k < S; ++k)
        for (int d = 0; d < 5; ++d) tr[i][k][d] = rows[((size_t)i * S + k) * 5 + d];
    }
    seen_ = session_->version();
  }

 private:
  krrt::Session* session_;
  mutable unsigned long seen_ = 0;
  std::vector<double> cost_, time_;
};

// ---------------------------------------------------------------------------- Cost / Connector / Radius
// Cost slot: c*(xi -> xf) and the arrival time, same call as solver.cost
class CostInt2D {
 public:
  static constexpr bool device_backed = true;
  explicit CostInt2D(TrajectorySolver* s) : solver_(s) {}
  void configure(krrt_config& c) const { c.r = solver_->weight; }
  template <typename A, typename B>
  std::tuple<double, double> operator()(const A& xi, const B& xf) const { return solver_->solver.cost(xi, xf); }

 private:
  TrajectorySolver* solver_;
};

// Connector slot: the optimal edge xi -> xf as S samples (x,y,vx,vy,t0+t), what the tree stores per node
class ConnectorInt2D {
 public:
  static constexpr bool device_backed = true;
  ConnectorInt2D(TrajectorySolver* s, TreeInt2D* t) : solver_(s), tree_(t) {}
  void configure(krrt_config& c) const { c.r = solver_->weight; }
  template <typename A, typename B>
  Edge operator()(const A& xi, const B& xf, double t0 = 0.0) const {
    krrt::Session& s = tree_->session();
    s.config().r = solver_->weight;
    const double a[4] = {xi[0], xi[1], xi[2], xi[3]}, b[4] = {xf[0], xf[1], xf[2], xf[3]};
    const int S = tree_->edge_samples;
    Edge e;
    e.resize(S);
    s.check(krrt_edge_samples(s.ctx(), a, b, &t0, 1, S, e[0].data()));
    return e;
  }

 private:
  TrajectorySolver* solver_;
  TreeInt2D* tree_;
};

// NeighborRadius slot: rho(n) = min(rho_max, gamma (ln n / n)^(1/4)) in the 4-D state space
class NeighborRadiusInt2D {
 public:
  static constexpr bool device_backed = true;
  double gamma = 7.0, rho_max = 4.0;
  void configure(krrt_config& c) const {
    c.gamma = gamma;
    c.rho_max = rho_max;
  }
};

// ---------------------------------------------------------------------------------------------- sampler
class SamplerInt2D {
 public:
  static constexpr bool device_backed = true;
  using State = State4;
  explicit SamplerInt2D(krrt::Session* s) : direct_sampler(new Direct), session_(s) {}

  State target{};  // sampler.target = xg (…node.cpp:227-228,263-264): the goal handed to grow() is the direct sample
  struct Direct {
    std::array<double, 2> p{{0.0, 1.0}};  // p[0] = probability of returning the target (…node.cpp:229)
  };
  std::unique_ptr<Direct> direct_sampler;
  std::array<double, 4> lo{{-9.0, -6.0, -2.0, -2.0}}, hi{{9.0, 6.0, 2.0, 2.0}};  // field 18 x 12 m, |v| <= 2
  uint64_t seed = 0;
  uint32_t query_id = 0;

  void set_direct_sample(bool enable, double prob) {
    direct_sampler->p = {enable ? prob : 0.0, enable ? 1.0 - prob : 1.0};
  }
  template <typename X>
  void set_next_sample(const X& x) {
    for (int k = 0; k < 4; ++k) next_[k] = x[k];
    has_next_ = true;
  }
  bool take_next_sample(double out[4]) {
    if (!has_next_) return false;
    for (int k = 0; k < 4; ++k) out[k] = next_[k];
    has_next_ = false;
    return true;
  }
  // the sample the last expansion used (…node.cpp:178)
  State last_sample() const {
    State s{};
    if (!session_->started) return s;
    const krrt::RunAhead& ra = session_->run_ahead();
    if (ra.served > 0 && (size_t)ra.served <= ra.log.size() && ra.log[(size_t)ra.served - 1].n_nodes >= 0) {
      for (int k = 0; k < 4; ++k) s[k] = ra.log[(size_t)ra.served - 1].sample[k];
      return s;
    }
    krrt_ctx* c = session_->ctx();
    session_->sync_tree();
    session_->check(krrt_last_sample(c, s.data()));
    return s;
  }
  void configure(krrt_config& c) const {
    for (int k = 0; k < 4; ++k) {
      c.lo[k] = lo[k];
      c.hi[k] = hi[k];
    }
    c.p_direct = direct_sampler->p[0];
    c.seed = seed;
    c.query_id = query_id;
  }

 private:
  krrt::Session* session_;
  double next_[4] = {0, 0, 0, 0};
  bool has_next_ = false;
};

// ----------------------------------------------------------------------------------------- environments
// Dynamic = false: static discs (get_robosoccer_env); true: discs moving at constant velocity, tested in
// time-space (get_dynamic_soccer_env).  Distinct types, as robosoccer_visual.hpp:13,28 dispatches on them.
template <bool Dynamic>
class SoccerEnv {
 public:
  static constexpr bool device_backed = true;
  static constexpr bool dynamic = Dynamic;
  explicit SoccerEnv(krrt::Session* s) : session_(s) {}

  double collision_radius = 1.35;           // config/config.yaml:5
  std::vector<std::array<double, 4>> obs;   // (x, y, vx, vy), metres

  template <typename Container>
  void setObstacles(const Container& c) {   // env.setObstacles(subs.getObstacles()) (…node.cpp:140)
    obs.clear();
    for (const auto& o : c) {
      if (obs.size() == KRRT_MAX_OBS) break;
      obs.push_back({o[0], o[1], o[2], o[3]});
    }
    ++version_;
  }
  // obstacle positions at time t (robosoccer_visual.hpp:35)
  std::vector<std::array<double, 2>> at(double t) const {
    std::vector<std::array<double, 2>> p;
    for (const auto& o : obs) p.push_back({Dynamic ? std::fma(o[2], t, o[0]) : o[0], Dynamic ? std::fma(o[3], t, o[1]) : o[1]});
    return p;
  }
  // Environment predicate on its own: does the optimal edge xi -> xf, started at absolute time t0, hit a disc?
  template <typename A, typename B>
  bool collide(const A& xi, const B& xf, double t0 = 0.0) {
    krrt_ctx* c = session_->ctx();
    upload(*session_);
    const double a[4] = {xi[0], xi[1], xi[2], xi[3]}, b[4] = {xf[0], xf[1], xf[2], xf[3]};
    double tau = 0, cost = 0;
    session_->check(krrt_steer_cost(c, a, b, 1, &tau, &cost));
    unsigned char hit = 0;
    session_->check(krrt_edge_collide(c, a, b, &tau, &t0, 1, &hit));
    return hit != 0;
  }
  void configure(krrt_config& c) const { c.R = collision_radius; }
  // the discs resident on the device are not (the current version of) this environment's
  bool stale(const krrt::Session& s) const { return !(s.obstacle_owner == this && s.obstacle_version == version_); }
  // make this environment's discs the ones resident on the device (two environments share one tree/session)
  void upload(krrt::Session& s) {
    static_assert(sizeof(std::array<double, 4>) == 4 * sizeof(double), "obstacle layout");
    if (s.obstacle_owner == this && s.obstacle_version == version_) return;
    s.settings_change();  // (the iterations run ahead saw the discs that are resident now; a cycle under way stops running ahead)
    s.config().R = collision_radius;
    s.check(krrt_set_obstacles(s.ctx(), obs.empty() ? nullptr : obs[0].data(), (int)obs.size(), Dynamic ? 1 : 0));
    s.obstacle_owner = this;
    s.obstacle_version = version_;
  }

 private:
  krrt::Session* session_;
  unsigned long version_ = 1;
};
using RobosoccerEnv = SoccerEnv<false>;
using DynamicRobosoccerEnv = SoccerEnv<true>;

// ------------------------------------------------------------------------------------------ goal checker
template <typename Env>
class GoalCheckerInt2D {
 public:
  static constexpr bool device_backed = true;
  using State = State4;
  GoalCheckerInt2D(Env* env, SamplerInt2D* sampler) : env_(env), sampler_(sampler), rng_(12345u) {}
  // KRRT_GOAL_CONNECT (default): the goal is connected to the cheapest collision-free node -> goal edge the tree offers
  // when grow() returns; KRRT_GOAL_TOLERANCE: a node is a goal node iff it lies within goal_tol of the goal (krrt.h)
  int goal_mode = KRRT_GOAL_CONNECT;
  double goal_tol = 0.0;
  void configure(krrt_config& c) const {
    c.goal_mode = goal_mode;
    c.goal_tol = goal_tol;
  }
  // a resting goal on the field, outside every disc at t = 0 (…node.cpp:221; host-side, called once at start-up)
  State randomGoal() {
    State g{};
    for (int tries = 0; tries < 1000; ++tries) {
      g[0] = std::uniform_real_distribution<double>(sampler_->lo[0], sampler_->hi[0])(rng_);
      g[1] = std::uniform_real_distribution<double>(sampler_->lo[1], sampler_->hi[1])(rng_);
      bool free = true;
      for (const auto& o : env_->obs) free = free && std::hypot(g[0] - o[0], g[1] - o[1]) >= env_->collision_radius;
      if (free) break;
    }
    return g;
  }

 private:
  Env* env_;
  SamplerInt2D* sampler_;
  std::mt19937_64 rng_;
};

using RRTStarInt2D = RRTStar<TreeInt2D, CostInt2D, SamplerInt2D, NeighborRadiusInt2D, RobosoccerEnv,
                             GoalCheckerInt2D<RobosoccerEnv>, ConnectorInt2D>;
using RRTStarInt2DTimeSpaceObs = RRTStar<TreeInt2D, CostInt2D, SamplerInt2D, NeighborRadiusInt2D, DynamicRobosoccerEnv,
                                         GoalCheckerInt2D<DynamicRobosoccerEnv>, ConnectorInt2D>;

// --------------------------------------------------------------------------------------------- singletons
namespace Wrapper {
inline krrt::Session& get_session() { static krrt::Session s; return s; }
inline TreeInt2D& get_tree_int2d() { static TreeInt2D t(&get_session()); return t; }
inline TrajectorySolver& get_trajectory_solver() { static TrajectorySolver s(&get_session()); return s; }
inline CostInt2D& get_cost_int2d() { static CostInt2D c(&get_trajectory_solver()); return c; }
inline ConnectorInt2D& get_connector_int2d() { static ConnectorInt2D c(&get_trajectory_solver(), &get_tree_int2d()); return c; }
inline NeighborRadiusInt2D& get_radius_int2d() { static NeighborRadiusInt2D r; return r; }
inline SamplerInt2D& get_sampler() { static SamplerInt2D s(&get_session()); return s; }
inline SamplerInt2D& get_sampler_dynamic_env() { static SamplerInt2D s(&get_session()); return s; }
inline RobosoccerEnv& get_robosoccer_env() { static RobosoccerEnv e(&get_session()); return e; }
inline DynamicRobosoccerEnv& get_dynamic_soccer_env() { static DynamicRobosoccerEnv e(&get_session()); return e; }
inline GoalCheckerInt2D<RobosoccerEnv>& get_goal() {
  static GoalCheckerInt2D<RobosoccerEnv> g(&get_robosoccer_env(), &get_sampler());
  return g;
}
inline GoalCheckerInt2D<DynamicRobosoccerEnv>& get_goal_dynamic_env() {
  static GoalCheckerInt2D<DynamicRobosoccerEnv> g(&get_dynamic_soccer_env(), &get_sampler_dynamic_env());
  return g;
}
inline RRTStarInt2D& get_rrtstar_int2d() {
  static RRTStarInt2D r(get_tree_int2d(), get_cost_int2d(), get_sampler(), get_radius_int2d(), get_robosoccer_env(),
                        get_goal(), get_connector_int2d());
  return r;
}
inline RRTStarInt2DTimeSpaceObs& get_rrtstar_int2d_timespace_obs() {
  static RRTStarInt2DTimeSpaceObs r(get_tree_int2d(), get_cost_int2d(), get_sampler_dynamic_env(), get_radius_int2d(),
                                    get_dynamic_soccer_env(), get_goal_dynamic_env(), get_connector_int2d());
  return r;
}
}  // namespace Wrapper
}  // namespace Kinodynamic

namespace Models {
inline void init_integrator2d() {}  // "actually this is no longer needed" (…node.cpp:42-43)
}

#endif  // KRRT_INTEGRATOR2DRRT_HPP
