// krrt_math.cuh — per-thread double-integrator steering, optimal-edge and Philox sampler math.
//
// Arithmetic contract (what makes tree indices comparable bit-for-bit with the CPU oracle):
// IEEE-754 binary64, only + - * / sqrt and fma(); every fused multiply-add is written
// explicitly and the translation unit is compiled with -fmad=false so nvcc contracts nothing
// else.  div.rn.f64 / sqrt.rn.f64 are correctly rounded on sm_100a, as on the host.
//
// Stands in for the reference's Cost / Connector / Sampler slots:
//   solver.cost(xi,xf) -> (t,c)   robosoccer_motion_planning_node.cpp:242-244
//   solver.solve(xi,xf,dt)        robosoccer_motion_planning_node.cpp:245
//   sampler (direct sampling p)   robosoccer_motion_planning_node.cpp:225-229
// Math: SURVEY.md §8-M (fixed-final-state, free-final-time LQR for p'' = u, J = int 1 + r|u|^2).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include "krrt_real.cuh"

#define KRRT_HD __host__ __device__ __forceinline__
#define KRRT_NEWTON_MAX 100
#define KRRT_STATIONARY_MAX 48  // == KO_MAX_STATIONARY: linear convergence (ratio 1/2) at a near-double root of f'
// relative Newton step that counts as converged (a few ulp; the CPU checker uses the same constant): without it the
// iteration dithers +-1 ulp around the root for ~20 more steps and every lane of the warp waits for the slowest
#ifdef KRRT_F32_BUILD
#define KRRT_XTOL 5e-7f   // binary32 build (krrt_real.cuh): the same few ulp (== KO_XTOL of the CPU checker's float build)
#else
#define KRRT_XTOL 1e-15
#endif
#ifndef KRRT_ITER_HOOK
#define KRRT_ITER_HOOK(which)  // test instrumentation point (tests/native/newton_iters.cu counts iterations here)
#endif

namespace krrt {

struct State4 {
  double x, y, vx, vy;
};

// quartic f(t) = t^4 - a t^2 + b t - c0 and derivatives (Horner with explicit fma)
KRRT_HD double quartic(double t, double a, double b, double c0) { return fma(fma(fma(t, t, -a), t, b), t, -c0); }
KRRT_HD double quartic_d1(double t, double a, double b) { return fma(fma(KR(4.0) * t, t, -KR(2.0) * a), t, b); }
KRRT_HD double quartic_d2(double t, double a) { return fma(KR(12.0) * t, t, -KR(2.0) * a); }
// J(t) = t + a/t - (b/2)/t^2 + (c0/3)/t^3
KRRT_HD double arrival_cost(double t, double a, double b, double c0) {
  double w = KR(1.0) / t;
  return fma(w, fma(w, fma(w, c0 / KR(3.0), -KR(0.5) * b), a), t);
}

// Newton safeguarded by a sign bracket [lo,hi] with f(lo) <= 0 <= f(hi)
KRRT_HD double bracketed_root(double lo, double hi, double x, double a, double b, double c0) {
  for (int it = 0; it < KRRT_NEWTON_MAX; ++it) {
    KRRT_ITER_HOOK(0)
    double f = quartic(x, a, b, c0);
    if (f == KR(0.0)) return x;
    if (f < KR(0.0)) lo = x; else hi = x;
    double xn = x - f / quartic_d1(x, a, b);
    if (!(xn >= lo && xn <= hi)) xn = lo + KR(0.5) * (hi - lo);  // Newton left the bracket: bisect
    if (fabs(xn - x) <= KRRT_XTOL * x) return xn;
    x = xn;
  }
  return x;
}

// stationary points of f: monotone Newton on f' from the left (start 0) and from the right (start sqrt(a/2))
KRRT_HD double stationary_left(double a, double b) {
  double x = KR(0.0);
  for (int it = 0; it < KRRT_STATIONARY_MAX; ++it) {
    KRRT_ITER_HOOK(1)
    double g = quartic_d1(x, a, b);
    if (!(g > KR(0.0))) break;
    double xn = x - g / quartic_d2(x, a);
    if (!(xn > x)) break;
    if (xn - x <= KRRT_XTOL * xn) {
      x = xn;
      break;
    }
    x = xn;
  }
  return x;
}
KRRT_HD double stationary_right(double a, double b) {
  double x = sqrt(KR(0.5) * a);
  for (int it = 0; it < KRRT_STATIONARY_MAX; ++it) {
    KRRT_ITER_HOOK(2)
    double g = quartic_d1(x, a, b);
    if (!(g > KR(0.0))) break;
    double xn = x - g / quartic_d2(x, a);
    if (!(xn < x)) break;
    if (x - xn <= KRRT_XTOL * x) {
      x = xn;
      break;
    }
    x = xn;
  }
  return x;
}

// optimal arrival time tau* and cost c* for x0 -> x1: every positive root of the quartic is
// visited and the cheapest kept (1 or 3 positive roots).
//
// Written for SIMT convergence: every lane of a warp walks the SAME instruction sequence — one bracket
// set-up, ONE bracketed-Newton loop for the primary root, and a rare (0.07 % of field-scale pairs) second pass
// for the three-root zone — instead of one inlined Newton loop per case.  The operations each lane performs,
// and their order, are exactly those of the oracle's case tree (the CPU checker restates the same case tree), so results are
// bit-identical; only the instruction layout differs.
KRRT_HD void steer(const State4& x0, const State4& x1, double r, double& tau_out, double& cost_out) {
  double dpx = x1.x - x0.x, dpy = x1.y - x0.y;
  double sx = x0.vx + x1.vx, sy = x0.vy + x1.vy;
  double dd = fma(dpy, dpy, dpx * dpx);
  double ds = fma(dpy, sy, dpx * sx);
  double v00 = fma(x0.vy, x0.vy, x0.vx * x0.vx);
  double v01 = fma(x0.vy, x1.vy, x0.vx * x1.vx);
  double v11 = fma(x1.vy, x1.vy, x1.vx * x1.vx);
  double vv = (v00 + v01) + v11;
  double a = (KR(4.0) * r) * vv;
  double b = (KR(24.0) * r) * ds;
  double c0 = (KR(36.0) * r) * dd;
  if (a < KR(0.0)) a = KR(0.0);
  if (c0 == KR(0.0)) {  // coincident positions: f = t^2 (t^2 - a)
    double t = sqrt(a);
    tau_out = t;
    cost_out = t + t;
    return;
  }
  double disc = sqrt(fma(a, a, KR(4.0) * c0));
  double T2 = KR(0.5) * (a + disc);
  double T = sqrt(T2);  // root of the b = 0 quartic: upper bound when b >= 0, lower bound when b < 0
  // bracket [lo, hi] and start x of the primary root
  double lo, hi, x;
  double s1 = KR(0.0);
  bool three = false;  // three-root zone: a second root right of the right stationary point may be cheaper
  if (b < KR(0.0)) {
    lo = T;
    hi = T + (-b) / T2;
    x = hi;
  } else {
    double ti = sqrt(a / KR(6.0));  // inflection of f
    double bcrit = (KR(4.0) * a / KR(3.0)) * ti;
    hi = T;
    x = T;
    if (b >= bcrit) {  // f monotone for t > 0
      lo = KR(0.0);
    } else if (b * ti < c0) {  // f < 0 on (0, ti]
      lo = ti;
    } else {
      s1 = stationary_left(a, b);
      if (quartic(s1, a, b, c0) < KR(0.0)) {
        lo = ti;
      } else {
        three = true;
        lo = KR(0.0);
        hi = s1;
        x = c0 / b;
        if (!(x < s1)) x = KR(0.5) * s1;
      }
    }
  }
  double tau = bracketed_root(lo, hi, x, a, b, c0);
  double cost = arrival_cost(tau, a, b, c0);
  if (three) {
    double s2 = stationary_right(a, b);
    if (quartic(s2, a, b, c0) <= KR(0.0)) {
      double t3 = bracketed_root(s2, T, T, a, b, c0);
      double c3 = arrival_cost(t3, a, b, c0);
      if (c3 < cost) {
        tau = t3;
        cost = c3;
      }
    }
  }
  tau_out = tau;
  cost_out = cost;
}

// One axis of the optimal edge as a cubic in h = t / tau: p(h) = p0 + A h + B h^2 + C h^3
struct AxisCubic {
  double p0, A, B, C;
};
KRRT_HD AxisCubic axis_cubic(double p0, double v0, double p1, double v1, double tau) {
  AxisCubic k;
  double dp = p1 - p0;
  k.p0 = p0;
  k.A = v0 * tau;
  k.B = fma(-(KR(2.0) * v0 + v1), tau, KR(3.0) * dp);
  k.C = fma(v0 + v1, tau, -KR(2.0) * dp);
  return k;
}
KRRT_HD double axis_pos(const AxisCubic& k, double h) { return fma(h, fma(h, fma(h, k.C, k.B), k.A), k.p0); }
KRRT_HD double axis_vel(const AxisCubic& k, double h, double v0, double tau) {
  if (tau == KR(0.0)) return v0;
  return v0 + (h * fma(KR(3.0) * h, k.C, KR(2.0) * k.B)) / tau;
}

// control input u = p'' of the edge at h (the service response's `inputs` for the 2-D integrator model)
KRRT_HD double axis_acc(const AxisCubic& k, double h, double tau) {
  if (tau == KR(0.0)) return KR(0.0);
  return fma(KR(6.0) * h, k.C, KR(2.0) * k.B) / (tau * tau);
}

struct Edge {
  AxisCubic kx, ky;
  double tau, t0, vx0, vy0;
};
KRRT_HD Edge make_edge(const State4& a, const State4& b, double tau, double t0) {
  Edge e;
  e.kx = axis_cubic(a.x, a.vx, b.x, b.vx, tau);
  e.ky = axis_cubic(a.y, a.vy, b.y, b.vy, tau);
  e.tau = tau;
  e.t0 = t0;
  e.vx0 = a.vx;
  e.vy0 = a.vy;
  return e;
}

// disc test of edge sample k (of S) against obstacle o = (x,y,vx,vy); R2 = R*R
KRRT_HD bool edge_point_hits(const Edge& e, int k, int S, const double4& o, int dynamic, double R2) {
  double h = (S > 1) ? (double)k / (double)(S - 1) : KR(0.0);
  double px = axis_pos(e.kx, h), py = axis_pos(e.ky, h);
  double ox = o.x, oy = o.y;
  if (dynamic) {
    double t = e.t0 + e.tau * h;
    ox = fma(o.z, t, ox);
    oy = fma(o.w, t, oy);
  }
  double dx = px - ox, dy = py - oy;
  return fma(dy, dy, dx * dx) < R2;
}

// the same test against all m obstacles for the sample point at h = k / (S - 1) (position evaluated once)
KRRT_HD bool edge_point_hits_any(const Edge& e, double h, const double4* obs, int m, int dynamic, double R2) {
  const double px = axis_pos(e.kx, h), py = axis_pos(e.ky, h);
  const double t = e.t0 + e.tau * h;
  bool hit = false;
  for (int j = 0; j < m; ++j) {
    const double4 o = obs[j];
    double ox = o.x, oy = o.y;
    if (dynamic) {
      ox = fma(o.z, t, ox);
      oy = fma(o.w, t, oy);
    }
    const double dx = px - ox, dy = py - oy;
    hit |= fma(dy, dy, dx * dx) < R2;
  }
  return hit;
}

KRRT_HD double dist2(const State4& a, const State4& b) {
  double d0 = a.x - b.x, d1 = a.y - b.y, d2 = a.vx - b.vx, d3 = a.vy - b.vy;
  return fma(d3, d3, fma(d2, d2, fma(d1, d1, d0 * d0)));
}

// Philox4x32-10 (Salmon et al., SC'11): key = seed, counter = (c0,c1,c2,c3)
struct U4 {
  uint32_t a, b, c, d;
};
KRRT_HD U4 philox4x32_10(uint64_t seed, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3) {
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int round = 0; round < 10; ++round) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    uint32_t n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    uint32_t n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  U4 o = {c0, c1, c2, c3};
  return o;
}
KRRT_HD double unit53(uint32_t hi, uint32_t lo) {
  uint64_t u = ((uint64_t)hi << 32) | lo;
  return (double)(u >> 11) * KR(0x1.0p-53);
}

struct SamplerParams {
  double lo[4], hi[4];
  double p_direct;
  uint64_t seed;
};
// counter = (iter lo, query_id, word, iter hi); word 2 = direct-goal Bernoulli, 0 = position, 1 = velocity
KRRT_HD State4 draw_sample(const SamplerParams& sp, uint32_t query_id, uint64_t iter, bool has_goal,
                           const State4& goal) {
  uint32_t i0 = (uint32_t)iter, i3 = (uint32_t)(iter >> 32);
  if (has_goal && sp.p_direct > KR(0.0)) {
    U4 w = philox4x32_10(sp.seed, i0, query_id, 2u, i3);
    if (unit53(w.a, w.b) < sp.p_direct) return goal;
  }
  State4 s;
  U4 w = philox4x32_10(sp.seed, i0, query_id, 0u, i3);
  s.x = sp.lo[0] + unit53(w.a, w.b) * (sp.hi[0] - sp.lo[0]);
  s.y = sp.lo[1] + unit53(w.c, w.d) * (sp.hi[1] - sp.lo[1]);
  w = philox4x32_10(sp.seed, i0, query_id, 1u, i3);
  s.vx = sp.lo[2] + unit53(w.a, w.b) * (sp.hi[2] - sp.lo[2]);
  s.vy = sp.lo[3] + unit53(w.c, w.d) * (sp.hi[3] - sp.lo[3]);
  return s;
}

}  // namespace krrt
