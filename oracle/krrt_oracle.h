/* krrt_oracle.h — CPU restatement of the kinodynamic RRT* hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load it.  The CUDA library (libkrrt.so) never links,
 * includes or calls anything in this directory.
 *
 * PARITY UNPINNED.  The reference's own implementation of this path lives in
 * the git submodule src/motion_planning (alifahrri/motion_planning @ 3b3183e,
 * /root/reference/.gitmodules:1-3) which is absent from the mounted tree, and
 * no reference test pins any RRT-star, cost, KD-tree or collision value
 * (src/robosoccer_motion_planning/test/unit_test.cpp:7-86 covers Trajectory1D
 * only).  This file therefore restates the *published* fixed-final-state /
 * free-final-time double-integrator steering (Webb & van den Berg 2013,
 * SURVEY.md §8-M) and the canonical RRT* loop behind the call-site contract
 * visible in the reference:
 *   rrt.setStart / setIteration(0)   robosoccer_motion_planning_node.cpp:135-136
 *   env.setObstacles                 robosoccer_motion_planning_node.cpp:140
 *   rrt.grow(xg) -> bool             robosoccer_motion_planning_node.cpp:177
 *   sampler.last_sample/set_next_sample/set_direct_sample/target   :178,184,225-228
 *   rrt.goalIndex()                  robosoccer_visual.hpp:19
 *   tree.states(idx)                 robosoccer_motion_planning_node.cpp:237
 *   solver.cost(xi,xf)->(t,c), solver.solve(xi,xf,dt)   :242-245
 *   env.at(t), env.collision_radius  robosoccer_visual.hpp:15,35
 * Every free choice is a named field of ko_config.
 */
#ifndef KRRT_ORACLE_H
#define KRRT_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KO_MAX_OBS 64
#define KO_MAX_WAYPOINTS 32
enum { KO_GOAL_CONNECT = 0, KO_GOAL_TOLERANCE = 1 };

typedef struct ko_config {
  double r;            /* input weight of J = int (1 + r|u|^2) dt  (config.yaml:1 -> 0.5) */
  double R;            /* collision radius (config.yaml:5 -> 1.35)                         */
  int32_t S;           /* samples per edge incl. both end points (O5)                      */
  int32_t capacity;    /* max nodes                                                        */
  double lo[4];        /* sampler lower bounds x,y,vx,vy (O1)                              */
  double hi[4];        /* sampler upper bounds                                             */
  double gamma;        /* near radius rho(n) = min(rho_max, gamma (ln n / n)^(1/4)) (O3)   */
  double rho_max;
  double p_direct;     /* probability of sampling the goal itself (O2, config.yaml:7)      */
  double goal_tol;     /* KO_GOAL_TOLERANCE: node is a goal node iff d2(x, goal) <= goal_tol^2 */
  uint64_t seed;       /* Philox key                                                       */
  uint32_t query_id;   /* Philox counter word 1                                            */
  int32_t use_kdtree;  /* 1: incremental kd-tree near/nearest, 0: brute force              */
  int32_t goal_mode;   /* O8: KO_GOAL_CONNECT (default) or KO_GOAL_TOLERANCE, see connect_goal / find_goal */
} ko_config;

typedef struct ko_result {
  int32_t solved;
  int32_t goal_index;
  int32_t n_nodes;
  int32_t path_len;    /* true length of root->goal chain (may exceed KO_MAX_WAYPOINTS) */
  double cost;
  double time;
  double waypoints[KO_MAX_WAYPOINTS][5]; /* x,y,vx,vy,t_arrival */
} ko_result;

typedef struct ko_stats {
  int64_t iters, inserted, fail_reject, fail_dup, fail_nofree;
  int64_t rewired, propagated, sum_near, steer_calls, collide_calls;
} ko_stats;

typedef struct ko_planner ko_planner;

void ko_default_config(ko_config* c);

/* steering: optimal arrival time and cost == solver.cost(xi,xf) */
void ko_steer(const double x0[4], const double x1[4], double r, double* tau, double* cost);
void ko_steer_batch(const double* x0, const double* x1, int n, double r, double* tau, double* cost);
/* state on the optimal edge at h = t/tau in [0,1]; out = x,y,vx,vy */
void ko_edge_state(const double x0[4], const double x1[4], double tau, double h, double out[4]);
/* == solver.solve(xi,xf,dt): rows (x,y,vx,vy,t0+t) at t = 0,dt,2dt,.. < tau, then tau. returns rows */
int ko_solve(const double x0[4], const double x1[4], double r, double dt, double t0, double* rows, int cap);
/* accelerations (ax, ay) at the same rows: the service response's `inputs` for the 2-D integrator model */
int ko_solve_inputs(const double x0[4], const double x1[4], double r, double dt, double* rows, int cap);
/* time-space disc collision of the optimal edge started at absolute time t0 */
int ko_edge_collide(const double x0[4], const double x1[4], double tau, double t0,
                    const double* obs, int m, int dynamic, double R, int S);
/* sampler: Philox4x32-10 keyed (seed), counter (iter, query_id, word, 0) */
void ko_sample(const ko_config* c, uint64_t iter, const double* goal, double out[4]);
void ko_philox(uint64_t seed, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t out[4]);
double ko_rho2(const ko_config* c, int n);

ko_planner* ko_create(const ko_config* c);
void ko_destroy(ko_planner* p);
void ko_set_obstacles(ko_planner* p, const double* xyvxvy, int m, int dynamic);
void ko_set_start(ko_planner* p, const double x0[4]);
void ko_set_near_radius(ko_planner* p, double gamma, double rho_max);
/* n_iter expansions. samples: NULL or n_iter x 4 (a row whose first entry is NaN
 * means "draw internally").  returns 1 when a goal node exists at the end; *goal_index is its index: a tree node, or
 * (KO_GOAL_CONNECT) ko_size(p), the goal node stored right behind the tree — ko_states / ko_parents / ko_ctc /
 * ko_times / ko_path / ko_fill_result all accept it.                                                          */
int ko_grow(ko_planner* p, const double* goal, int n_iter, const double* samples, int* goal_index);
int ko_size(const ko_planner* p);
int64_t ko_iteration(const ko_planner* p);
const double* ko_states(const ko_planner* p);   /* n x 4 AoS */
const int32_t* ko_parents(const ko_planner* p);
const double* ko_ctc(const ko_planner* p);
const double* ko_times(const ko_planner* p);
const double* ko_last_sample(const ko_planner* p);
void ko_get_stats(const ko_planner* p, ko_stats* s);
/* brute-force / kd-tree near query on the current tree (ascending indices) */
int ko_near(ko_planner* p, const double x[4], double radius2, int32_t* idx, int cap, int32_t* nearest,
            double* nearest_d2);
/* root -> idx chain: states (len x 4) ; returns len */
int ko_path(const ko_planner* p, int idx, double* states, double* times, int cap);
void ko_fill_result(const ko_planner* p, int solved, int goal_index, ko_result* out);
/* check tree invariants; returns 0 if OK else a code */
int ko_check_invariants(const ko_planner* p, double tol);

/* q independent queries: starts q x 4, goals q x 4, obstacles q x m x 4.  Philox query_id = base + i */
void ko_plan_batch(const ko_config* c, const double* starts, const double* goals, const double* obstacles,
                   int q, int m, int dynamic, int n_iter, uint32_t query_base, ko_result* out, int nthreads);

#ifdef __cplusplus
}
#endif
#endif
