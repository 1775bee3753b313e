/* krrt_oracle.c — CPU restatement of the kinodynamic RRT* hot path.
 * TEST INFRASTRUCTURE ONLY; PARITY UNPINNED — see krrt_oracle.h for why, and
 * for the reference call sites (file:line) each entry point stands in for.
 *
 * Arithmetic contract (shared with the CUDA kernels so that tree indices can
 * be compared bit-for-bit): IEEE-754 binary64, only + - * / sqrt and fma(),
 * every fused multiply-add written explicitly, everything else evaluated in
 * the written order.  Build with -ffp-contract=off (see oracle/Makefile).
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* The binary32 build (oracle/Makefile: libkrrt_oracle_f32.so, the checker of KRRT_F32 contexts) is THIS source with
 * every `double` replaced by `float`: -DKO_F32 below, gcc -fsingle-precision-constant for the literals, <tgmath.h> so
 * that fma / sqrt / fabs / log on float operands are the float functions.  Its interface is float throughout
 * (tests/oracle_binding.py builds a second set of ctypes structures for it). */
#ifdef KO_F32
#include <tgmath.h>
#define double float
#endif

#include "krrt_oracle.h"

#define KO_MAX_NEWTON 100
/* Newton on f' converges only linearly (ratio 1/2) when its two positive roots nearly coincide: 48 steps = 2^-48 */
#define KO_MAX_STATIONARY 48
/* relative step below which a Newton iteration has converged: a few ulp.  Without it the iteration dithers by
 * +-1 ulp around the root for ~20 more steps (the rounding noise of f is of the order of one step). */
#ifdef KO_F32
#define KO_XTOL 5e-7 /* == KRRT_XTOL of the kernels' binary32 build */
#else
#define KO_XTOL 1e-15
#endif

/* ------------------------------------------------------------------ */
/* Steering: tau^4 - a tau^2 + b tau - c0 = 0, SURVEY.md §8-M          */
/* stands in for solver.cost(xi,xf) (…node.cpp:242-244)               */
/* ------------------------------------------------------------------ */

static inline double q_f(double t, double a, double b, double c0) {
  return fma(fma(fma(t, t, -a), t, b), t, -c0);
}
static inline double q_df(double t, double a, double b) {
  return fma(fma(4.0 * t, t, -2.0 * a), t, b);
}
static inline double q_ddf(double t, double a) { return fma(12.0 * t, t, -2.0 * a); }

/* c(tau) = tau + a/tau - (b/2)/tau^2 + (c0/3)/tau^3 */
static inline double q_cost(double t, double a, double b, double c0) {
  double inv = 1.0 / t;
  return fma(inv, fma(inv, fma(inv, c0 / 3.0, -0.5 * b), a), t);
}

/* bracketed Newton: f(lo) <= 0 <= f(hi), start x in [lo,hi] */
static double q_root(double lo, double hi, double x, double a, double b, double c0) {
  for (int it = 0; it < KO_MAX_NEWTON; ++it) {
    double f = q_f(x, a, b, c0);
    if (f == 0.0) return x;
    if (f < 0.0) lo = x; else hi = x;
    double xn = x - f / q_df(x, a, b);
    if (!(xn >= lo && xn <= hi)) xn = lo + 0.5 * (hi - lo); /* Newton left the bracket: bisect */
    if (fabs(xn - x) <= KO_XTOL * x) return xn;
    x = xn;
  }
  return x;
}

/* smallest positive critical point of f (root of f' on (0, sqrt(a/6))), b > 0 */
static double q_crit_left(double a, double b) {
  double x = 0.0;
  for (int it = 0; it < KO_MAX_STATIONARY; ++it) {
    double g = q_df(x, a, b);
    if (!(g > 0.0)) break;
    double xn = x - g / q_ddf(x, a);
    if (!(xn > x)) break;
    if (xn - x <= KO_XTOL * xn) { x = xn; break; }
    x = xn;
  }
  return x;
}
/* largest critical point of f (root of f' right of sqrt(a/6)), 0 <= b */
static double q_crit_right(double a, double b) {
  double x = sqrt(0.5 * a);
  for (int it = 0; it < KO_MAX_STATIONARY; ++it) {
    double g = q_df(x, a, b);
    if (!(g > 0.0)) break;
    double xn = x - g / q_ddf(x, a);
    if (!(xn < x)) break;
    if (x - xn <= KO_XTOL * x) { x = xn; break; }
    x = xn;
  }
  return x;
}

void ko_steer(const double x0[4], const double x1[4], double r, double* tau_out, double* cost_out) {
  double dpx = x1[0] - x0[0], dpy = x1[1] - x0[1];
  double sx = x0[2] + x1[2], sy = x0[3] + x1[3];
  double dd = fma(dpy, dpy, dpx * dpx);
  double ds = fma(dpy, sy, dpx * sx);
  double v00 = fma(x0[3], x0[3], x0[2] * x0[2]);
  double v01 = fma(x0[3], x1[3], x0[2] * x1[2]);
  double v11 = fma(x1[3], x1[3], x1[2] * x1[2]);
  double vv = (v00 + v01) + v11;
  double a = (4.0 * r) * vv;
  double b = (24.0 * r) * ds;
  double c0 = (36.0 * r) * dd;
  if (a < 0.0) a = 0.0;

  if (c0 == 0.0) { /* same position: f = tau^2 (tau^2 - a) */
    double t = sqrt(a);
    *tau_out = t;
    *cost_out = t + t;
    return;
  }
  double D = sqrt(fma(a, a, 4.0 * c0));
  double T2 = 0.5 * (a + D);
  double T = sqrt(T2); /* root of tau^4 - a tau^2 - c0 */
  double tau, cost;
  if (b < 0.0) {
    double U = T + (-b) / T2;
    tau = q_root(T, U, U, a, b, c0);
    cost = q_cost(tau, a, b, c0);
  } else {
    double ti = sqrt(a / 6.0); /* inflection of f */
    double bcrit = (4.0 * a / 3.0) * ti;
    if (b >= bcrit) { /* f monotone on tau > 0 */
      tau = q_root(0.0, T, T, a, b, c0);
      cost = q_cost(tau, a, b, c0);
    } else if (b * ti < c0) { /* f < 0 on (0, ti]: only the large root */
      tau = q_root(ti, T, T, a, b, c0);
      cost = q_cost(tau, a, b, c0);
    } else {
      double s1 = q_crit_left(a, b);
      if (q_f(s1, a, b, c0) < 0.0) {
        tau = q_root(ti, T, T, a, b, c0);
        cost = q_cost(tau, a, b, c0);
      } else {
        double x = c0 / b;
        if (!(x < s1)) x = 0.5 * s1;
        tau = q_root(0.0, s1, x, a, b, c0);
        cost = q_cost(tau, a, b, c0);
        double s2 = q_crit_right(a, b);
        if (q_f(s2, a, b, c0) <= 0.0) {
          double t3 = q_root(s2, T, T, a, b, c0);
          double c3 = q_cost(t3, a, b, c0);
          if (c3 < cost) { tau = t3; cost = c3; }
        }
      }
    }
  }
  *tau_out = tau;
  *cost_out = cost;
}

void ko_steer_batch(const double* x0, const double* x1, int n, double r, double* tau, double* cost) {
  for (int i = 0; i < n; ++i) ko_steer(x0 + 4 * i, x1 + 4 * i, r, tau + i, cost + i);
}

/* ------------------------------------------------------------------ */
/* Optimal edge (cubic Hermite in position) — Connector / solver.solve */
/* ------------------------------------------------------------------ */

typedef struct { double p0, A, B, C; } ko_axis;

static inline void ko_axis_coef(double p0, double v0, double p1, double v1, double tau, ko_axis* k) {
  double dp = p1 - p0;
  k->p0 = p0;
  k->A = v0 * tau;
  k->B = fma(-(2.0 * v0 + v1), tau, 3.0 * dp);
  k->C = fma(v0 + v1, tau, -2.0 * dp);
}
static inline double ko_axis_pos(const ko_axis* k, double h) {
  return fma(h, fma(h, fma(h, k->C, k->B), k->A), k->p0);
}
static inline double ko_axis_vel(const ko_axis* k, double h, double v0, double tau) {
  if (tau == 0.0) return v0;
  return v0 + (h * fma(3.0 * h, k->C, 2.0 * k->B)) / tau;
}

void ko_edge_state(const double x0[4], const double x1[4], double tau, double h, double out[4]) {
  ko_axis kx, ky;
  ko_axis_coef(x0[0], x0[2], x1[0], x1[2], tau, &kx);
  ko_axis_coef(x0[1], x0[3], x1[1], x1[3], tau, &ky);
  out[0] = ko_axis_pos(&kx, h);
  out[1] = ko_axis_pos(&ky, h);
  out[2] = ko_axis_vel(&kx, h, x0[2], tau);
  out[3] = ko_axis_vel(&ky, h, x0[3], tau);
}

int ko_solve(const double x0[4], const double x1[4], double r, double dt, double t0, double* rows, int cap) {
  double tau, cost;
  ko_steer(x0, x1, r, &tau, &cost);
  int n = 0;
  for (int k = 0;; ++k) {
    double t = (double)k * dt;
    int last = !(t < tau);
    if (last) t = tau;
    if (n < cap) {
      double h = (tau == 0.0) ? 0.0 : t / tau;
      ko_edge_state(x0, x1, tau, h, rows + 5 * n);
      rows[5 * n + 4] = t0 + t;
    }
    ++n;
    if (last) break;
  }
  return n;
}

/* control input of the optimal edge: u(t) = p''(t) = (2B + 6C h) / tau^2 per axis, at the rows ko_solve emits
 * (TrajectoryService response `inputs`, trajectory_generator_node.cpp:33-47, for the 2-D integrator model) */
static inline double ko_axis_acc(const ko_axis* k, double h, double tau) {
  if (tau == 0.0) return 0.0;
  return fma(6.0 * h, k->C, 2.0 * k->B) / (tau * tau);
}
int ko_solve_inputs(const double x0[4], const double x1[4], double r, double dt, double* rows, int cap) {
  double tau, cost;
  ko_steer(x0, x1, r, &tau, &cost);
  ko_axis kx, ky;
  ko_axis_coef(x0[0], x0[2], x1[0], x1[2], tau, &kx);
  ko_axis_coef(x0[1], x0[3], x1[1], x1[3], tau, &ky);
  int n = 0;
  for (int k = 0;; ++k) {
    double t = (double)k * dt;
    int last = !(t < tau);
    if (last) t = tau;
    if (n < cap) {
      double h = (tau == 0.0) ? 0.0 : t / tau;
      rows[2 * n + 0] = ko_axis_acc(&kx, h, tau);
      rows[2 * n + 1] = ko_axis_acc(&ky, h, tau);
    }
    ++n;
    if (last) break;
  }
  return n;
}

/* Environment::collide for one edge — static (dynamic=0) or time-space (dynamic=1);
 * stands in for the Environment slot (…node.cpp:84-85,140; robosoccer_visual.hpp:35) */
int ko_edge_collide(const double x0[4], const double x1[4], double tau, double t0, const double* obs, int m,
                    int dynamic, double R, int S) {
  ko_axis kx, ky;
  ko_axis_coef(x0[0], x0[2], x1[0], x1[2], tau, &kx);
  ko_axis_coef(x0[1], x0[3], x1[1], x1[3], tau, &ky);
  double R2 = R * R;
  for (int k = 0; k < S; ++k) {
    double h = (S > 1) ? (double)k / (double)(S - 1) : 0.0;
    double px = ko_axis_pos(&kx, h), py = ko_axis_pos(&ky, h);
    double t = t0 + tau * h;
    for (int j = 0; j < m; ++j) {
      const double* o = obs + 4 * j;
      double ox = o[0], oy = o[1];
      if (dynamic) { ox = fma(o[2], t, ox); oy = fma(o[3], t, oy); }
      double dx = px - ox, dy = py - oy;
      if (fma(dy, dy, dx * dx) < R2) return 1;
    }
  }
  return 0;
}

/* ------------------------------------------------------------------ */
/* Sampler — Philox4x32-10 (Salmon et al. 2011)                        */
/* ------------------------------------------------------------------ */

void ko_philox(uint64_t seed, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t out[4]) {
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
  for (int r = 0; r < 10; ++r) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    uint32_t n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    uint32_t n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

static inline double ko_u01(uint32_t hi, uint32_t lo) {
  uint64_t u = ((uint64_t)hi << 32) | lo;
  return (double)(u >> 11) * 0x1.0p-53;
}

void ko_sample(const ko_config* c, uint64_t iter, const double* goal, double out[4]) {
  uint32_t w[4];
  uint32_t i0 = (uint32_t)iter, i3 = (uint32_t)(iter >> 32);
  if (goal && c->p_direct > 0.0) {
    ko_philox(c->seed, i0, c->query_id, 2u, i3, w);
    if (ko_u01(w[0], w[1]) < c->p_direct) {
      out[0] = goal[0]; out[1] = goal[1]; out[2] = goal[2]; out[3] = goal[3];
      return;
    }
  }
  ko_philox(c->seed, i0, c->query_id, 0u, i3, w);
  out[0] = c->lo[0] + ko_u01(w[0], w[1]) * (c->hi[0] - c->lo[0]);
  out[1] = c->lo[1] + ko_u01(w[2], w[3]) * (c->hi[1] - c->lo[1]);
  ko_philox(c->seed, i0, c->query_id, 1u, i3, w);
  out[2] = c->lo[2] + ko_u01(w[0], w[1]) * (c->hi[2] - c->lo[2]);
  out[3] = c->lo[3] + ko_u01(w[2], w[3]) * (c->hi[3] - c->lo[3]);
}

double ko_rho2(const ko_config* c, int n) {
  if (n < 2) return 0.0;
  double q = log((double)n) / (double)n;
  double rho = c->gamma * sqrt(sqrt(q));
  if (rho > c->rho_max) rho = c->rho_max;
  return rho * rho;
}

void ko_default_config(ko_config* c) {
  memset(c, 0, sizeof(*c));
  c->r = 0.5; c->R = 1.35; c->S = 10; c->capacity = 1 << 16;
  c->lo[0] = -9.0; c->hi[0] = 9.0; c->lo[1] = -6.0; c->hi[1] = 6.0;
  c->lo[2] = -2.0; c->hi[2] = 2.0; c->lo[3] = -2.0; c->hi[3] = 2.0;
  c->gamma = 7.0; c->rho_max = 4.0; c->p_direct = 0.0; c->goal_tol = 0.0;
  c->seed = 0; c->query_id = 0; c->use_kdtree = 1; c->goal_mode = KO_GOAL_CONNECT;
}

/* ------------------------------------------------------------------ */
/* Planner                                                             */
/* ------------------------------------------------------------------ */

struct ko_planner {
  ko_config cfg;
  int n;
  int64_t iter;
  double* state; /* n x 4 */
  int32_t* parent;
  double *ctc, *time, *ecost, *etime;
  int32_t *head, *next, *prev; /* child lists */
  double obs[KO_MAX_OBS * 4];
  int m, dynamic;
  double* rho2;
  /* kd-tree over node indices */
  int32_t *kl, *kr;
  uint8_t* kdim;
  /* scratch */
  int32_t* near;
  int near_cap;
  double *cf, *tf;
  uint8_t* tried;
  int32_t* stack;
  double* stack_d;
  double last_sample[4];
  ko_stats st;
  void* gsort; /* connect_goal: (cost through node, node) pairs */
};

typedef struct { double c; int32_t i; } ko_gcand;

static inline double d2_4(const double* a, const double* b) {
  double d0 = a[0] - b[0], d1 = a[1] - b[1], d2 = a[2] - b[2], d3 = a[3] - b[3];
  return fma(d3, d3, fma(d2, d2, fma(d1, d1, d0 * d0)));
}

ko_planner* ko_create(const ko_config* c) {
  ko_planner* p = (ko_planner*)calloc(1, sizeof(ko_planner));
  p->cfg = *c;
  int cap = c->capacity + 1; /* + the goal node's slot right behind the last tree node (connect_goal) */
  p->state = (double*)malloc(sizeof(double) * 4 * cap);
  p->parent = (int32_t*)malloc(sizeof(int32_t) * cap);
  p->ctc = (double*)malloc(sizeof(double) * cap);
  p->time = (double*)malloc(sizeof(double) * cap);
  p->ecost = (double*)malloc(sizeof(double) * cap);
  p->etime = (double*)malloc(sizeof(double) * cap);
  p->head = (int32_t*)malloc(sizeof(int32_t) * cap);
  p->next = (int32_t*)malloc(sizeof(int32_t) * cap);
  p->prev = (int32_t*)malloc(sizeof(int32_t) * cap);
  p->kl = (int32_t*)malloc(sizeof(int32_t) * cap);
  p->kr = (int32_t*)malloc(sizeof(int32_t) * cap);
  p->kdim = (uint8_t*)malloc(cap);
  p->rho2 = (double*)malloc(sizeof(double) * (cap + 1));
  for (int n = 0; n <= cap; ++n) p->rho2[n] = ko_rho2(c, n);
  p->gsort = malloc(sizeof(ko_gcand) * cap);
  p->near_cap = cap;
  p->near = (int32_t*)malloc(sizeof(int32_t) * cap);
  p->cf = (double*)malloc(sizeof(double) * cap);
  p->tf = (double*)malloc(sizeof(double) * cap);
  p->tried = (uint8_t*)malloc(cap);
  p->stack = (int32_t*)malloc(sizeof(int32_t) * cap);
  p->stack_d = (double*)malloc(sizeof(double) * cap);
  return p;
}

void ko_destroy(ko_planner* p) {
  if (!p) return;
  free(p->state); free(p->parent); free(p->ctc); free(p->time); free(p->ecost); free(p->etime);
  free(p->head); free(p->next); free(p->prev); free(p->kl); free(p->kr); free(p->kdim); free(p->rho2);
  free(p->gsort); free(p->near); free(p->cf); free(p->tf); free(p->tried); free(p->stack); free(p->stack_d);
  free(p);
}

/* == env.setObstacles(subs.getObstacles()) (…node.cpp:139-141) */
void ko_set_obstacles(ko_planner* p, const double* o, int m, int dynamic) {
  if (m > KO_MAX_OBS) m = KO_MAX_OBS;
  p->m = m;
  p->dynamic = dynamic;
  if (m > 0) memcpy(p->obs, o, sizeof(double) * 4 * m);
}

static void kd_insert(ko_planner* p, int idx) {
  p->kl[idx] = -1; p->kr[idx] = -1;
  if (idx == 0) { p->kdim[0] = 0; return; }
  const double* x = p->state + 4 * idx;
  int cur = 0;
  for (;;) {
    int d = p->kdim[cur];
    int right = !(x[d] < p->state[4 * cur + d]);
    int32_t* slot = right ? &p->kr[cur] : &p->kl[cur];
    if (*slot < 0) { *slot = idx; p->kdim[idx] = (uint8_t)((d + 1) & 3); return; }
    cur = *slot;
  }
}

static void link_child(ko_planner* p, int par, int c) {
  int h = p->head[par];
  p->next[c] = h; p->prev[c] = -1;
  if (h >= 0) p->prev[h] = c;
  p->head[par] = c;
}
static void unlink_child(ko_planner* p, int c) {
  int par = p->parent[c], nx = p->next[c], pv = p->prev[c];
  if (pv >= 0) p->next[pv] = nx; else p->head[par] = nx;
  if (nx >= 0) p->prev[nx] = pv;
}

/* == rrt.setStart(xs); rrt.setIteration(0) (…node.cpp:135-136) */
/* NeighborRadius parameters changed on a live planner (what krrt_set_params does for gamma / rho_max) */
void ko_set_near_radius(ko_planner* p, double gamma, double rho_max) {
  p->cfg.gamma = gamma;
  p->cfg.rho_max = rho_max;
  for (int n = 0; n <= p->cfg.capacity; ++n) p->rho2[n] = ko_rho2(&p->cfg, n);
}

void ko_set_start(ko_planner* p, const double x0[4]) {
  p->n = 1; p->iter = 0;
  memcpy(p->state, x0, sizeof(double) * 4);
  p->parent[0] = -1; p->ctc[0] = 0.0; p->time[0] = 0.0; p->ecost[0] = 0.0; p->etime[0] = 0.0;
  p->head[0] = -1; p->next[0] = -1; p->prev[0] = -1;
  kd_insert(p, 0);
  memset(&p->st, 0, sizeof(p->st));
}

static int cmp_i32(const void* a, const void* b) {
  int32_t x = *(const int32_t*)a, y = *(const int32_t*)b;
  return (x > y) - (x < y);
}

/* near set {i : d2 <= radius2} ascending + nearest (lowest index on ties) */
static int near_query(ko_planner* p, const double* x, double radius2, int32_t* out, int cap, int32_t* nearest,
                      double* nearest_d2) {
  int k = 0, best = -1;
  double bd = INFINITY;
  if (!p->cfg.use_kdtree) {
    for (int i = 0; i < p->n; ++i) {
      double d = d2_4(p->state + 4 * i, x);
      if (d < bd) { bd = d; best = i; }
      if (d <= radius2) { if (k < cap) out[k] = i; ++k; }
    }
  } else {
    int sp = 0;
    int32_t* st = p->stack;
    double* sd = p->stack_d;
    st[sp] = 0; sd[sp++] = 0.0;
    while (sp > 0) {
      int cur = st[--sp];
      /* re-check the plane distance recorded at push time against the current bound */
      if (sd[sp] > (radius2 > bd ? radius2 : bd)) continue;
      const double* s = p->state + 4 * cur;
      double d = d2_4(s, x);
      if (d < bd || (d == bd && cur < best)) { bd = d; best = cur; }
      if (d <= radius2) { if (k < cap) out[k] = cur; ++k; }
      int dim = p->kdim[cur];
      double diff = x[dim] - s[dim];
      int nearc = diff < 0.0 ? p->kl[cur] : p->kr[cur];
      int farc = diff < 0.0 ? p->kr[cur] : p->kl[cur];
      double lim = radius2 > bd ? radius2 : bd; /* must serve both queries */
      if (farc >= 0 && diff * diff <= lim) { st[sp] = farc; sd[sp++] = diff * diff; }
      if (nearc >= 0) { st[sp] = nearc; sd[sp++] = 0.0; }
    }
    int kk = k < cap ? k : cap;
    qsort(out, kk, sizeof(int32_t), cmp_i32);
  }
  *nearest = best;
  *nearest_d2 = bd;
  return k;
}

int ko_near(ko_planner* p, const double x[4], double radius2, int32_t* idx, int cap, int32_t* nearest,
            double* nearest_d2) {
  return near_query(p, x, radius2, idx, cap, nearest, nearest_d2);
}

static int sample_in_static_collision(const ko_planner* p, const double* x) {
  double R2 = p->cfg.R * p->cfg.R;
  for (int j = 0; j < p->m; ++j) {
    double dx = x[0] - p->obs[4 * j], dy = x[1] - p->obs[4 * j + 1];
    if (fma(dy, dy, dx * dx) < R2) return 1;
  }
  return 0;
}

/* recompute cost-to-come / arrival time of every descendant of j from its parent (O7) */
static void propagate(ko_planner* p, int j) {
  int sp = 0;
  int32_t* st = p->stack;
  for (int c = p->head[j]; c >= 0; c = p->next[c]) st[sp++] = c;
  while (sp > 0) {
    int d = st[--sp];
    int par = p->parent[d];
    p->ctc[d] = p->ctc[par] + p->ecost[d];
    p->time[d] = p->time[par] + p->etime[d];
    p->st.propagated++;
    for (int c = p->head[d]; c >= 0; c = p->next[c]) st[sp++] = c;
  }
}

/* one RRT* expansion == RRTStar::grow (…node.cpp:177).  1 inserted, 0 failed, -1 full */
static int grow_once(ko_planner* p, const double* xs) {
  const ko_config* c = &p->cfg;
  p->st.iters++;
  memcpy(p->last_sample, xs, sizeof(double) * 4);
  if (!p->dynamic && sample_in_static_collision(p, xs)) { p->st.fail_reject++; return 0; }
  int n = p->n;
  if (n >= c->capacity) return -1;
  int32_t nearest;
  double nd2;
  int K = near_query(p, xs, p->rho2[n], p->near, p->near_cap, &nearest, &nd2);
  if (nd2 == 0.0) { p->st.fail_dup++; return 0; }
  if (K == 0) { p->near[0] = nearest; K = 1; }
  p->st.sum_near += K;
  /* forward costs near -> new */
  for (int k = 0; k < K; ++k) {
    int i = p->near[k];
    ko_steer(p->state + 4 * i, xs, c->r, &p->tf[k], &p->cf[k]);
    p->tried[k] = 0;
  }
  p->st.steer_calls += K;
  /* parent = argmin over collision-free of ctc[i] + c(i->new), lowest index on ties */
  int bestk = -1;
  for (int round = 0; round < K; ++round) {
    int kk = -1;
    double bc = INFINITY;
    for (int k = 0; k < K; ++k) {
      if (p->tried[k]) continue;
      double cand = p->ctc[p->near[k]] + p->cf[k];
      if (cand < bc) { bc = cand; kk = k; }
    }
    if (kk < 0) break;
    p->tried[kk] = 1;
    int i = p->near[kk];
    p->st.collide_calls++;
    if (!ko_edge_collide(p->state + 4 * i, xs, p->tf[kk], p->time[i], p->obs, p->m, p->dynamic, c->R, c->S)) {
      bestk = kk;
      break;
    }
  }
  if (bestk < 0) { p->st.fail_nofree++; return 0; }
  int par = p->near[bestk];
  int nw = n;
  memcpy(p->state + 4 * nw, xs, sizeof(double) * 4);
  p->parent[nw] = par;
  p->ecost[nw] = p->cf[bestk];
  p->etime[nw] = p->tf[bestk];
  p->ctc[nw] = p->ctc[par] + p->cf[bestk];
  p->time[nw] = p->time[par] + p->tf[bestk];
  p->head[nw] = -1;
  link_child(p, par, nw);
  kd_insert(p, nw);
  p->n = n + 1;
  p->st.inserted++;
  /* rewire, ascending index, immediate propagation */
  for (int k = 0; k < K; ++k) {
    int j = p->near[k];
    if (j == par) continue;
    double tb, cb;
    ko_steer(xs, p->state + 4 * j, c->r, &tb, &cb);
    p->st.steer_calls++;
    double cand = p->ctc[nw] + cb;
    if (cand < p->ctc[j]) {
      p->st.collide_calls++;
      if (!ko_edge_collide(xs, p->state + 4 * j, tb, p->time[nw], p->obs, p->m, p->dynamic, c->R, c->S)) {
        unlink_child(p, j);
        p->parent[j] = nw;
        link_child(p, nw, j);
        p->ecost[j] = cb; p->etime[j] = tb;
        p->ctc[j] = cand;
        p->time[j] = p->time[nw] + tb;
        p->st.rewired++;
        propagate(p, j);
      }
    }
  }
  return 1;
}

/* O8, goal_mode KO_GOAL_TOLERANCE: the goal node is the cheapest tree node within goal_tol of the goal (4-D) */
static int find_goal(const ko_planner* p, const double* goal) {
  if (!goal) return -1;
  double tol2 = p->cfg.goal_tol * p->cfg.goal_tol;
  int best = -1;
  double bc = INFINITY;
  for (int i = 0; i < p->n; ++i) {
    if (d2_4(p->state + 4 * i, goal) <= tol2 && p->ctc[i] < bc) { bc = p->ctc[i]; best = i; }
  }
  return best;
}

static int cmp_gcand(const void* a, const void* b) {
  const ko_gcand *x = (const ko_gcand*)a, *y = (const ko_gcand*)b;
  if (x->c != y->c) return x->c < y->c ? -1 : 1;
  return (x->i > y->i) - (x->i < y->i);
}

/* O8, goal_mode KO_GOAL_CONNECT (default): the goal is CONNECTED to the tree.  Every node the planner inserted
 * (i >= 1; the start itself is not a candidate: a direct start -> goal edge is what direct sampling is for,
 * ...node.cpp:225-229) offers the edge i -> goal: optimal steering, cost ctc[i] + c*(i -> goal), collision-tested as
 * any other edge, departing at the node's arrival time.  The goal's parent is the cheapest collision-free offer on the
 * tree AS IT STANDS when grow() returns (lowest index among equal costs), so rewires that make a node cheaper make the
 * goal cheaper, and the answer does not depend on how the iterations were split over grow() calls.  The goal node is
 * stored right behind the last tree node (index n = ko_size): rrt.goalIndex() > 0 (robosoccer_visual.hpp:19,40-45),
 * tree(goalIndex) is the goal state and tree.states(goalIndex) the chain root -> ... -> goal; it is not part of the
 * tree the next grow() call searches (no near query finds it, it has no children) and is rebuilt at the end of that
 * call.  A node that coincides with the goal (a direct sample) offers a zero-length edge and, if it wins, IS the goal
 * node.  Returns the goal index or -1. */
static int connect_goal(ko_planner* p, const double* goal) {
  if (!goal) return -1;
  const ko_config* c = &p->cfg;
  ko_gcand* g = (ko_gcand*)p->gsort;
  int n = p->n, k = 0;
  for (int i = 1; i < n; ++i) {
    const double* x = p->state + 4 * i;
    double tg = 0.0, cg = 0.0;
    if (d2_4(x, goal) != 0.0) ko_steer(x, goal, c->r, &tg, &cg);
    p->tf[i] = tg; p->cf[i] = cg;
    g[k].c = p->ctc[i] + cg; g[k].i = i; ++k;
  }
  qsort(g, k, sizeof(ko_gcand), cmp_gcand);
  for (int j = 0; j < k; ++j) {
    int i = g[j].i;
    const double* x = p->state + 4 * i;
    if (d2_4(x, goal) == 0.0) return i;
    if (ko_edge_collide(x, goal, p->tf[i], p->time[i], p->obs, p->m, p->dynamic, c->R, c->S)) continue;
    memcpy(p->state + 4 * n, goal, sizeof(double) * 4);
    p->parent[n] = i;
    p->ecost[n] = p->cf[i]; p->etime[n] = p->tf[i];
    p->ctc[n] = g[j].c;
    p->time[n] = p->time[i] + p->tf[i];
    return n;
  }
  return -1;
}

int ko_grow(ko_planner* p, const double* goal, int n_iter, const double* samples, int* goal_index) {
  for (int it = 0; it < n_iter; ++it) {
    double xs[4];
    int64_t gi = p->iter++;
    if (samples && !(samples[4 * it] != samples[4 * it])) memcpy(xs, samples + 4 * it, sizeof(xs));
    else ko_sample(&p->cfg, (uint64_t)gi, goal, xs);
    if (grow_once(p, xs) < 0) break;
  }
  int g = p->cfg.goal_mode == KO_GOAL_TOLERANCE ? find_goal(p, goal) : connect_goal(p, goal);
  if (goal_index) *goal_index = g;
  return g >= 0;
}

int ko_size(const ko_planner* p) { return p->n; }
int64_t ko_iteration(const ko_planner* p) { return p->iter; }
const double* ko_states(const ko_planner* p) { return p->state; }
const int32_t* ko_parents(const ko_planner* p) { return p->parent; }
const double* ko_ctc(const ko_planner* p) { return p->ctc; }
const double* ko_times(const ko_planner* p) { return p->time; }
const double* ko_last_sample(const ko_planner* p) { return p->last_sample; }
void ko_get_stats(const ko_planner* p, ko_stats* s) { *s = p->st; }

/* == tree.states(idx) (…node.cpp:237) */
int ko_path(const ko_planner* p, int idx, double* states, double* times, int cap) {
  int len = 0;
  for (int i = idx; i >= 0; i = p->parent[i]) ++len;
  int pos = len - 1;
  for (int i = idx; i >= 0; i = p->parent[i], --pos) {
    if (pos < cap) {
      if (states) memcpy(states + 4 * pos, p->state + 4 * i, sizeof(double) * 4);
      if (times) times[pos] = p->time[i];
    }
  }
  return len;
}

void ko_fill_result(const ko_planner* p, int solved, int goal_index, ko_result* out) {
  memset(out, 0, sizeof(*out));
  out->solved = solved;
  out->goal_index = goal_index;
  out->n_nodes = p->n;
  if (!solved) return;
  out->cost = p->ctc[goal_index];
  out->time = p->time[goal_index];
  int len = 0;
  for (int i = goal_index; i >= 0; i = p->parent[i]) ++len;
  out->path_len = len;
  int pos = len - 1;
  for (int i = goal_index; i >= 0; i = p->parent[i], --pos) {
    if (pos < KO_MAX_WAYPOINTS) {
      memcpy(out->waypoints[pos], p->state + 4 * i, sizeof(double) * 4);
      out->waypoints[pos][4] = p->time[i];
    }
  }
}

int ko_check_invariants(const ko_planner* p, double tol) {
  const ko_config* c = &p->cfg;
  if (p->n < 1 || p->parent[0] != -1 || p->ctc[0] != 0.0 || p->time[0] != 0.0) return 1;
  for (int i = 1; i < p->n; ++i) {
    int par = p->parent[i];
    if (par < 0 || par >= p->n || par == i) return 2;
    /* acyclic: walk to the root in < n steps */
    int steps = 0, a = i;
    while (a > 0 && steps <= p->n) { a = p->parent[a]; ++steps; }
    if (a != 0) return 3;
    if (p->ctc[i] != p->ctc[par] + p->ecost[i]) return 4;
    if (p->time[i] != p->time[par] + p->etime[i]) return 5;
    double t, cst;
    ko_steer(p->state + 4 * par, p->state + 4 * i, c->r, &t, &cst);
    if (t != p->etime[i] || cst != p->ecost[i]) return 6;
    /* endpoint hit and cost identity  c = int (1 + r u^2) dt */
    double e[4];
    ko_edge_state(p->state + 4 * par, p->state + 4 * i, t, 1.0, e);
    for (int k = 0; k < 4; ++k)
      if (fabs(e[k] - p->state[4 * i + k]) > tol * (1.0 + fabs(p->state[4 * i + k]))) return 7;
    /* child list membership */
    int found = 0;
    for (int ch = p->head[par]; ch >= 0; ch = p->next[ch]) if (ch == i) { found = 1; break; }
    if (!found) return 8;
  }
  return 0;
}

/* ------------------------------------------------------------------ */
/* Batched independent queries (configs C4/C5)                         */
/* ------------------------------------------------------------------ */

typedef struct {
  const ko_config* c;
  const double *starts, *goals, *obstacles;
  int q, m, dynamic, n_iter;
  uint32_t query_base;
  ko_result* out;
  int tid, nthreads;
} batch_arg;

static void* batch_worker(void* vp) {
  batch_arg* a = (batch_arg*)vp;
  ko_config c = *a->c;
  if (c.capacity > a->n_iter + 2) c.capacity = a->n_iter + 2;
  ko_planner* p = ko_create(&c);
  for (int i = a->tid; i < a->q; i += a->nthreads) {
    p->cfg.query_id = a->query_base + (uint32_t)i;
    ko_set_obstacles(p, a->obstacles + (size_t)i * a->m * 4, a->m, a->dynamic);
    ko_set_start(p, a->starts + 4 * (size_t)i);
    int gi = -1;
    int solved = ko_grow(p, a->goals + 4 * (size_t)i, a->n_iter, NULL, &gi);
    ko_fill_result(p, solved, gi, &a->out[i]);
  }
  ko_destroy(p);
  return NULL;
}

void ko_plan_batch(const ko_config* c, const double* starts, const double* goals, const double* obstacles, int q,
                   int m, int dynamic, int n_iter, uint32_t query_base, ko_result* out, int nthreads) {
  if (nthreads < 1) nthreads = 1;
  if (nthreads > q) nthreads = q > 0 ? q : 1;
  pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * nthreads);
  batch_arg* args = (batch_arg*)malloc(sizeof(batch_arg) * nthreads);
  for (int t = 0; t < nthreads; ++t) {
    args[t] = (batch_arg){c, starts, goals, obstacles, q, m, dynamic, n_iter, query_base, out, t, nthreads};
    if (t > 0) pthread_create(&th[t], NULL, batch_worker, &args[t]);
  }
  batch_worker(&args[0]);
  for (int t = 1; t < nthreads; ++t) pthread_join(th[t], NULL);
  free(th);
  free(args);
}
