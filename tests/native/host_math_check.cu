// Host-side bit-exactness check of csrc/krrt_math.cuh (compiled for the host by nvcc) against the CPU oracle.
// TEST ONLY: links oracle/_build/libkrrt_oracle.so.  Usage: host_math_check <n_pairs> ; exit 0 when every value is
// bit-identical.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>

#include "../../oracle/krrt_oracle.h"
#include "../../robosoccer_motion_planning_b200/csrc/krrt_math.cuh"

static bool same(double a, double b) { return std::memcmp(&a, &b, sizeof(double)) == 0; }

int main(int argc, char** argv) {
  int n = argc > 1 ? std::atoi(argv[1]) : 200000;
  std::mt19937_64 rng(12345);
  std::uniform_real_distribution<double> ux(-9, 9), uy(-6, 6), uv(-2, 2), u01(0, 1);
  long bad = 0;
  for (int i = 0; i < n; ++i) {
    double a[4] = {ux(rng), uy(rng), uv(rng), uv(rng)}, b[4] = {ux(rng), uy(rng), uv(rng), uv(rng)};
    if (i % 7 == 0) { b[0] = a[0] + 1e-3 * uv(rng); b[1] = a[1] + 1e-3 * uv(rng); }   // near-coincident
    if (i % 11 == 0) { a[2] = a[3] = b[2] = b[3] = 0.0; }                                 // rest-to-rest
    if (i % 13 == 0) { b[0] = a[0]; b[1] = a[1]; }                                        // same position
    if (i % 17 == 0) { a[2] *= 4; a[3] *= 4; b[2] *= 4; b[3] *= 4; b[0] = a[0] + 0.3 * uv(rng); b[1] = a[1] + 0.3 * uv(rng); }  // multi-root zone
    double r = (i % 5 == 0) ? 0.1 + 2.0 * u01(rng) : 0.5;
    double t0, c0, t1, c1;
    ko_steer(a, b, r, &t0, &c0);
    krrt::State4 x0 = {a[0], a[1], a[2], a[3]}, x1 = {b[0], b[1], b[2], b[3]};
    krrt::steer(x0, x1, r, t1, c1);
    if (!same(t0, t1) || !same(c0, c1)) { if (bad < 5) std::printf("steer mismatch %d: %.17g %.17g vs %.17g %.17g\n", i, t0, c0, t1, c1); ++bad; }
    // edge states + collision points
    double h = u01(rng), e0[4];
    ko_edge_state(a, b, t0, h, e0);
    krrt::Edge ed = krrt::make_edge(x0, x1, t1, 0.25);
    double e1[4] = {krrt::axis_pos(ed.kx, h), krrt::axis_pos(ed.ky, h), krrt::axis_vel(ed.kx, h, x0.vx, t1), krrt::axis_vel(ed.ky, h, x0.vy, t1)};
    for (int k = 0; k < 4; ++k) if (!same(e0[k], e1[k])) { if (bad < 5) std::printf("edge mismatch %d/%d\n", i, k); ++bad; }
    double obs[8] = {ux(rng), uy(rng), uv(rng) / 2, uv(rng) / 2, ux(rng), uy(rng), uv(rng) / 2, uv(rng) / 2};
    for (int dyn = 0; dyn < 2; ++dyn) {
      int hit0 = ko_edge_collide(a, b, t0, 0.25, obs, 2, dyn, 1.35, 10);
      bool hit1 = false;
      for (int k = 0; k < 10; ++k) for (int j = 0; j < 2; ++j) {
        double4 o = make_double4(obs[4 * j], obs[4 * j + 1], obs[4 * j + 2], obs[4 * j + 3]);
        hit1 |= krrt::edge_point_hits(ed, k, 10, o, dyn, 1.35 * 1.35);
      }
      if ((hit0 != 0) != hit1) { if (bad < 5) std::printf("collide mismatch %d dyn=%d\n", i, dyn); ++bad; }
    }
  }
  // sampler
  ko_config c;
  ko_default_config(&c);
  c.seed = 0x123456789abcdefULL; c.query_id = 77; c.p_direct = 0.1;
  krrt::SamplerParams sp;
  for (int k = 0; k < 4; ++k) { sp.lo[k] = c.lo[k]; sp.hi[k] = c.hi[k]; }
  sp.p_direct = c.p_direct; sp.seed = c.seed;
  double goal[4] = {7, 0, 0, 0.5};
  krrt::State4 g = {7, 0, 0, 0.5};
  for (uint64_t it = 0; it < 50000; ++it) {
    uint64_t iter = it < 40000 ? it : (it << 20);
    double s0[4];
    ko_sample(&c, iter, goal, s0);
    krrt::State4 s1 = krrt::draw_sample(sp, c.query_id, iter, true, g);
    if (!same(s0[0], s1.x) || !same(s0[1], s1.y) || !same(s0[2], s1.vx) || !same(s0[3], s1.vy)) { if (bad < 5) std::printf("sample mismatch %llu\n", (unsigned long long)iter); ++bad; }
    ko_sample(&c, iter, nullptr, s0);
    s1 = krrt::draw_sample(sp, c.query_id, iter, false, g);
    if (!same(s0[0], s1.x) || !same(s0[3], s1.vy)) ++bad;
  }
  std::printf("host_math_check: %d pairs, mismatches=%ld\n", n, bad);
  return bad ? 1 : 0;
}
