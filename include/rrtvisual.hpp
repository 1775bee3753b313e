// rrtvisual.hpp — headless stand-in for the reference's RRTVisual (marker publisher of the absent submodule
// src/motion_planning; included by /root/reference/src/robosoccer_motion_planning/include/robosoccer_visual.hpp:6 and
// constructed at src/robosoccer_motion_planning_node.cpp:33).
//
// The reference's class turns the tree into visualization_msgs markers for RViz.  RViz / any graphics stack is
// unavailable on this box, so this one RECORDS what would have been published: every call appends
// primitives (discs, polylines, points; colour, namespace, z) to an in-memory frame, publish() closes the frame
// (and appends it to a CSV file when one is set: set_csv(path) or the environment variable KRRT_RRTVISUAL_CSV), clear()
// starts the next one.  Members are exactly the ones the two in-tree callers use (file:line under
// /root/reference/src/robosoccer_motion_planning/):
//   RRTVisual vis(node);                                                         src/robosoccer_motion_planning_node.cpp:33
//   vis.add_obstacles(x, y, r);                                                  include/robosoccer_visual.hpp:17
//   vis.set_trajectories(states, trajectories, parent, dim, size);               robosoccer_visual.hpp:18
//   vis.set_trajectory(path, dim);                                               robosoccer_visual.hpp:22
//   vis.add_circles(centres, r, dz, z, red, green, blue, alpha, ns);             robosoccer_visual.hpp:35
//   vis.add_trajectories<3,0,1,4>(edges, red, green, blue, alpha, ns[, w0, w1]); robosoccer_visual.hpp:39,42
//   vis.add_point<2,0,1>(state, red, green, blue, alpha, ns);                    robosoccer_visual.hpp:47-48
//   vis.delete_all(); vis.publish(); vis.clear();                                …node.cpp:146-148
// No arithmetic of the planner happens here; it only reads host mirrors the component set already holds.
#ifndef KRRT_RRTVISUAL_HPP
#define KRRT_RRTVISUAL_HPP

#include <array>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

class RRTVisual {
 public:
  enum Kind { DISC = 0, LINE_STRIP = 1, POINT = 2, DELETE_ALL = 3 };
  struct Primitive {
    Kind kind = DISC;
    std::string ns;
    float r = 0, g = 0, b = 0, a = 1;
    double radius = 0, height = 0;          // DISC: radius, extrusion in z; LINE_STRIP: line widths in radius / height
    std::vector<std::array<double, 3>> xyz;  // DISC / POINT: one entry; LINE_STRIP: the vertices
  };
  struct Frame {
    std::vector<Primitive> prims;
    size_t vertices() const {
      size_t n = 0;
      for (const auto& p : prims) n += p.xyz.size();
      return n;
    }
  };

  RRTVisual() { env_csv(); }
  // RRTVisual vis(node): the node handle is only used by the reference to advertise marker topics
  template <typename Node>
  explicit RRTVisual(Node&) { env_csv(); }

  void set_csv(const std::string& path) { csv_ = path; }
  const Frame& frame() const { return cur_; }              // what has been added since the last clear()
  const std::vector<Frame>& published() const { return out_; }
  size_t publish_count() const { return n_published_; }
  size_t keep_frames = 4;                                  // published frames kept in memory (the CSV keeps all)

  // one disc per obstacle at z = 0
  void add_obstacles(double x, double y, double r) { disc(x, y, 0.0, r, 0.0, 0.f, 0.f, 0.f, 1.f, "_obstacles"); }

  // the whole tree: edge i = trajectories[i] (samples of the edge parent[i] -> i); `dim` position components are drawn
  template <typename States, typename Trajectories, typename Parents>
  void set_trajectories(const States& states, const Trajectories& trajectories, const Parents& parent, int dim, size_t n) {
    (void)states;
    size_t i = 0;
    for (const auto& e : trajectories) {
      if (i >= n) break;
      if (i > 0 && parent[i] >= 0) strip(e, dim, 0.f, 1.f, 0.f, 0.4f, "_tree", 0.02f, 0.02f);
      ++i;
    }
  }
  // one polyline of points with `dim` components (the solution's edges: Edge::path())
  template <typename Path>
  void set_trajectory(const Path& path, int dim) {
    Primitive p;
    p.kind = LINE_STRIP;
    p.ns = "_solution";
    p.r = 0.f, p.g = 0.f, p.b = 1.f, p.a = 1.f;
    p.radius = p.height = 0.05;
    for (const auto& q : path) p.xyz.push_back({q[0], q[1], dim > 2 ? q[2] : 0.0});
    cur_.prims.push_back(std::move(p));
  }
  // moving obstacles drawn as a stack of discs over time: centres at height z, each dz thick
  template <typename Centres>
  void add_circles(const Centres& centres, double r, double dz, double z, double red, double green, double blue, double alpha,
                   const std::string& ns) {
    for (const auto& c : centres) disc(c[0], c[1], z, r, dz, (float)red, (float)green, (float)blue, (float)alpha, ns);
  }
  // edges as polylines in DIM dimensions: component IX -> x, IY -> y, IZ -> z (4 = time since the root)
  template <int DIM, int IX, int IY, int IZ, typename Edges>
  void add_trajectories(const Edges& edges, double red, double green, double blue, double alpha, const std::string& ns,
                        float w0 = 0.02f, float w1 = 0.02f) {
    for (const auto& e : edges) {
      Primitive p;
      p.kind = LINE_STRIP;
      p.ns = ns;
      p.r = (float)red, p.g = (float)green, p.b = (float)blue, p.a = (float)alpha;
      p.radius = w0, p.height = w1;
      for (const auto& s : e) p.xyz.push_back({s[IX], s[IY], DIM > 2 ? s[IZ] : 0.0});
      cur_.prims.push_back(std::move(p));
    }
  }
  template <int DIM, int IX, int IY, typename State>
  void add_point(const State& s, float red, float green, float blue, float alpha, const std::string& ns) {
    static_assert(DIM == 2, "add_point<2,ix,iy>");
    Primitive p;
    p.kind = POINT;
    p.ns = ns;
    p.r = red, p.g = green, p.b = blue, p.a = alpha;
    p.xyz.push_back({s[IX], s[IY], 0.0});
    cur_.prims.push_back(std::move(p));
  }

  // "clear all before re-drawing" (…node.cpp:145-146): a DELETEALL marker in front of the frame
  void delete_all() {
    Primitive p;
    p.kind = DELETE_ALL;
    cur_.prims.insert(cur_.prims.begin(), std::move(p));
  }
  void publish() {
    ++n_published_;
    if (!csv_.empty()) append_csv();
    out_.push_back(cur_);
    if (out_.size() > keep_frames) out_.erase(out_.begin());
  }
  void clear() { cur_.prims.clear(); }

 private:
  void env_csv() {
    if (const char* e = std::getenv("KRRT_RRTVISUAL_CSV")) csv_ = e;
  }
  void disc(double x, double y, double z, double r, double h, float cr, float cg, float cb, float ca, const std::string& ns) {
    Primitive p;
    p.kind = DISC;
    p.ns = ns;
    p.r = cr, p.g = cg, p.b = cb, p.a = ca;
    p.radius = r, p.height = h;
    p.xyz.push_back({x, y, z});
    cur_.prims.push_back(std::move(p));
  }
  template <typename Edge>
  void strip(const Edge& e, int dim, float cr, float cg, float cb, float ca, const char* ns, float w0, float w1) {
    Primitive p;
    p.kind = LINE_STRIP;
    p.ns = ns;
    p.r = cr, p.g = cg, p.b = cb, p.a = ca;
    p.radius = w0, p.height = w1;
    for (const auto& s : e) p.xyz.push_back({s[0], s[1], dim > 2 ? s[2] : 0.0});
    cur_.prims.push_back(std::move(p));
  }
  // frame,primitive,kind,ns,r,g,b,a,radius,height,vertex,x,y,z — one row per vertex
  void append_csv() const {
    std::FILE* f = std::fopen(csv_.c_str(), n_published_ == 1 ? "w" : "a");
    if (!f) return;
    if (n_published_ == 1) std::fputs("frame,primitive,kind,ns,r,g,b,a,radius,height,vertex,x,y,z\n", f);
    size_t k = 0;
    for (const auto& p : cur_.prims) {
      if (p.xyz.empty())
        std::fprintf(f, "%zu,%zu,%d,%s,%g,%g,%g,%g,%g,%g,-1,0,0,0\n", n_published_, k, (int)p.kind, p.ns.c_str(), p.r, p.g, p.b,
                     p.a, p.radius, p.height);
      for (size_t v = 0; v < p.xyz.size(); ++v)
        std::fprintf(f, "%zu,%zu,%d,%s,%g,%g,%g,%g,%g,%g,%zu,%.9g,%.9g,%.9g\n", n_published_, k, (int)p.kind, p.ns.c_str(), p.r,
                     p.g, p.b, p.a, p.radius, p.height, v, p.xyz[v][0], p.xyz[v][1], p.xyz[v][2]);
      ++k;
    }
    std::fclose(f);
  }

  Frame cur_;
  std::vector<Frame> out_;
  std::string csv_;
  size_t n_published_ = 0;
};

#endif  // KRRT_RRTVISUAL_HPP
