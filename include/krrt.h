/* krrt.h — C ABI of libkrrt.so, the sm_100a kinodynamic RRT* hot path.
 *
 * This is the drop-in boundary: plain C, opaque handles, caller-owned HOST
 * buffers, the library owns device memory, every call returns an int status (0 = OK) and
 * never throws or longjmps.  One krrt_ctx = one CUDA device + one stream; a ctx is NOT
 * thread-safe, distinct ctxs may be used from distinct threads / GPUs.  There is no CPU
 * fallback: every entry point runs CUDA kernels or fails with a status.
 *
 * What each entry point replaces in the reference (all paths under /root/reference/src/;
 * the RRT* sources themselves are the absent submodule src/motion_planning, so the citations
 * are the in-tree call sites of the RRTStar<Tree,Cost,Sampler,NeighborRadius,Environment,
 * GoalChecker,Connector> component surface):
 *
 *   krrt_set_start      rrt.setStart(xs); rrt.setIteration(0)   robosoccer_motion_planning/src/robosoccer_motion_planning_node.cpp:135-136
 *   krrt_set_obstacles  env.setObstacles(subs.getObstacles())   …node.cpp:139-141
 *   krrt_grow           for(i<iteration) rrt.grow(xg)           …node.cpp:176-185   (+ sampler.set_next_sample :184 via `samples`)
 *   goal_index out      rrt.goalIndex()                         robosoccer_motion_planning/include/robosoccer_visual.hpp:19
 *   krrt_steer_cost     solver.solver.cost(xi,xf) -> (t,c)      …node.cpp:242-244   (Cost slot)
 *   krrt_solve          solver.solver.solve(xi,xf,dt)           …node.cpp:245       (Connector slot)
 *   krrt_near           Tree nearest / radius query (KD-tree in the reference)      (Tree slot)
 *   krrt_edge_collide   Environment collision predicate, static and env.at(t) time-space variants
 *                                                               robosoccer_visual.hpp:16-17,35 ; …node.cpp:84-85
 *   krrt_tree_download  tree.tree.cloud.states / tree.parent / tree.trajectories    robosoccer_visual.hpp:18
 *   krrt_path           tree.states(idx) root->idx chain        …node.cpp:237
 *   krrt_trajectory     get_trajectory(): per edge solver.solve(a,b,0.1) appended   …node.cpp:232-249
 *   krrt_yaw_profiles   Trajectory1D::AngleController::angleControl + getAngleTrajectory for q queries
 *                                                               robosoccer_motion_planning/src/trajectory1d.cpp:23-51,203-394 ;
 *                                                               …node.cpp:306-314 ; trajectory_generator_node.cpp:6-53
 *   krrt_plan_batch     solve_rrt() (…node.cpp:167-196) for q independent queries at once
 */
#ifndef KRRT_H
#define KRRT_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KRRT_MAX_OBS 64
#define KRRT_MAX_WAYPOINTS 32

enum {
  KRRT_OK = 0,
  KRRT_ERR_INVALID = 1,       /* bad argument / config                                      */
  KRRT_ERR_CUDA = 2,          /* a CUDA runtime call failed (text in krrt_last_error)       */
  KRRT_ERR_NO_DEVICE = 3,     /* no usable sm_100 device: there is NO CPU fallback          */
  KRRT_ERR_NEAR_OVERFLOW = 4, /* a near set exceeded near_cap; raise near_cap and re-plan   */
  KRRT_ERR_STATE = 5,         /* call order (e.g. grow before set_start)                    */
  KRRT_ERR_UNSUPPORTED = 6
};

enum { KRRT_F64 = 0, KRRT_F32 = 1 };
/* Goal rule (GoalChecker slot; `solved` of rrt.grow(xg) …node.cpp:177, rrt.goalIndex() robosoccer_visual.hpp:19,40-45).
 * KRRT_GOAL_CONNECT (default): when a grow call returns, every node the planner inserted (index >= 1) offers the edge
 *   node -> goal (optimal steering, cost-to-come + c*, collision-tested like any other edge from the node's arrival
 *   time); the goal's parent is the cheapest collision-free offer on the tree as it stands, lowest index among equal
 *   costs.  The goal node is stored right behind the tree: goal_index == krrt_tree_size (> 0), valid for krrt_path /
 *   krrt_trajectory / krrt_result_of / krrt_tree_edges until the next call that changes the tree; a tree node that
 *   coincides with the goal (a direct sample) IS the goal node when it is the cheapest offer.
 * KRRT_GOAL_TOLERANCE: the goal node is the cheapest tree node within goal_tol of the goal (4-D Euclidean).       */
enum { KRRT_GOAL_CONNECT = 0, KRRT_GOAL_TOLERANCE = 1 };

typedef struct krrt_config {
  int32_t device;     /* CUDA device ordinal                                                 */
  int32_t precision;  /* KRRT_F64: binary64 throughout, as the reference (robotsubscriber.h:14), bit-exact with the CPU checker.
                         KRRT_F32: binary32 throughout (8-byte position / velocity pairs, fp32 steering, scans and collision
                         tests): the planner proper (krrt_set_start, krrt_grow, krrt_tree_*, krrt_path, krrt_result_of, ...),
                         krrt_plan_batch* and the stateless operators — the SAME kernels compiled in float, bit-exact with the
                         checker's float build; costs and times within 1e-4 of binary64 on the same edge.  Arrays at this
                         interface stay binary64 (rounded on the way in, widened on the way out).  Every
                         planning entry point serves both precisions; the development diagnostics below report the
                         binary64 planner's phases only */
  int32_t capacity;   /* max tree nodes                                                      */
  int32_t S;          /* samples per edge for collision checks, both end points included     */
  double r;           /* input weight of J = int (1 + r|u|^2) dt; solver.set_weight(r) (…node.cpp:91-92) */
  double R;           /* env.collision_radius (…node.cpp:84)                                 */
  double lo[4];       /* sampler bounds x,y,vx,vy                                            */
  double hi[4];
  double gamma;       /* near radius rho(n) = min(rho_max, gamma (ln n / n)^(1/4))           */
  double rho_max;
  double p_direct;    /* sampler.set_direct_sample(true, p) (…node.cpp:225-226)              */
  double goal_tol;    /* KRRT_GOAL_TOLERANCE: node is a goal node iff |x - goal| <= goal_tol */
  uint64_t seed;      /* Philox4x32-10 key                                                   */
  uint32_t query_id;  /* Philox counter word 1                                               */
  int32_t batch;      /* speculative batch: samples evaluated against a frozen tree per round, then
                         committed in order (results do not depend on it). 0 -> default      */
  int32_t near_cap;   /* per-sample candidate capacity. 0 -> default                         */
  int32_t goal_mode;  /* KRRT_GOAL_CONNECT (default) / KRRT_GOAL_TOLERANCE                    */
} krrt_config;

typedef struct krrt_result {
  int32_t solved;
  int32_t goal_index;
  int32_t n_nodes;
  int32_t path_len;   /* true length of the root->goal chain (may exceed KRRT_MAX_WAYPOINTS) */
  double cost;
  double time;
  double waypoints[KRRT_MAX_WAYPOINTS][5]; /* x,y,vx,vy,t_arrival */
} krrt_result;

typedef struct krrt_stats {
  int64_t iters, inserted, fail_reject, fail_dup, fail_nofree;
  int64_t rewired, propagated, sum_near, steer_calls, collide_calls;
} krrt_stats;

typedef struct krrt_ctx krrt_ctx;

void krrt_default_config(krrt_config* c);
int krrt_create(const krrt_config* c, krrt_ctx** out);
void krrt_destroy(krrt_ctx* ctx);
/* change the planner parameters of a live context (solver.set_weight, env.collision_radius, sampler and radius
 * settings: …node.cpp:84-92,225-226).  device / precision / capacity / batch / near_cap must equal the
 * create-time values (else KRRT_ERR_INVALID: create a new context).  The tree is left as it is.            */
int krrt_set_params(krrt_ctx* ctx, const krrt_config* cfg);
const char* krrt_last_error(const krrt_ctx* ctx); /* ctx may be NULL: last krrt_create / krrt_plan_batch_static error of this thread */
int krrt_device_count(void);

int krrt_set_obstacles(krrt_ctx* ctx, const double* xyvxvy, int m, int dynamic);
int krrt_set_start(krrt_ctx* ctx, const double x0[4]);
/* n_iter exact sequential RRT* expansions.  goal: NULL or 4 doubles.  samples: NULL or n_iter x 4
 * host doubles; a row whose first entry is NaN means "draw from the internal Philox sampler". */
int krrt_grow(krrt_ctx* ctx, const double* goal, int n_iter, const double* samples, int* solved,
              int* goal_index);
/* The caller's loop `for (i < iteration) solved = rrt.grow(xg)` (…node.cpp:176-185) as ONE device call: the tree after the
 * call is the tree n_iter single krrt_grow(ctx, goal, 1, row i) calls leave, and records[i] holds what call i would have
 * returned and what the getters the loop reads after it would have said (tree size, last sample).  include/rrtstar.hpp
 * serves the unchanged per-iteration loop from such a record.  1 <= n_iter <= 256, tree size after the call <= 1024,
 * else KRRT_ERR_UNSUPPORTED (use krrt_grow).  An iteration the tree had no room for repeats the previous record. */
typedef struct krrt_iter_record {
  int32_t n_nodes;     /* tree.tree.size() after the iteration                       (…node.cpp:179-180) */
  int32_t solved;      /* what rrt.grow(xg) returns                                  (…node.cpp:177)     */
  int32_t goal_index;  /* rrt.goalIndex() at that point, -1 when not solved                               */
  int32_t inserted;    /* the iteration added a node ("ok" / "failed" of …node.cpp:179)                   */
  double sample[4];    /* sampler.last_sample()                                      (…node.cpp:178)     */
  double goal_cost;    /* cost-to-come of the goal node when solved                                       */
} krrt_iter_record;
int krrt_grow_logged(krrt_ctx* ctx, const double* goal, int n_iter, const double* samples, krrt_iter_record* records,
                     int* solved, int* goal_index);
/* how the last krrt_grow call ended: the iterations it really ran and whether it stopped because the tree had reached
 * `capacity` (the call itself returns KRRT_OK: the tree is valid, only shorter than asked for) */
int krrt_grow_done(krrt_ctx* ctx, int64_t* iterations_done, int* full);
int krrt_tree_size(krrt_ctx* ctx, int* n);
int krrt_iteration(krrt_ctx* ctx, int64_t* iter);
int krrt_last_sample(krrt_ctx* ctx, double out[4]);
int krrt_get_stats(krrt_ctx* ctx, krrt_stats* out);
/* any of the output pointers may be NULL.  states: n x 4, parent: n, cost/time: n              */
int krrt_tree_download(krrt_ctx* ctx, double* states, int32_t* parent, double* cost, double* time);
/* rows [first, first + count) of the same arrays; the range may end with the goal node of KRRT_GOAL_CONNECT (index ==
 * krrt_tree_size while krrt_goal_index says so): tree(rrt.goalIndex()) of robosoccer_visual.hpp:45                */
int krrt_tree_read(krrt_ctx* ctx, int first, int count, double* states, int32_t* parent, double* cost, double* time);
/* rrt.goalIndex() (robosoccer_visual.hpp:19): the goal node found by the last krrt_grow, -1 when there is none    */
int krrt_goal_index(krrt_ctx* ctx, int* goal_index);
/* tree.trajectories (robosoccer_visual.hpp:18): for every node i the stored edge parent[i] -> i as S rows
 * (x,y,vx,vy,t), h = k/(S-1); node 0 gives S copies of the root.  rows: cap_nodes x S x 5 doubles; with cap_nodes >
 * tree size the connected goal node's edge follows the tree's                                                  */
int krrt_tree_edges(krrt_ctx* ctx, int S, double* rows, int cap_nodes);
/* root -> idx chain; states cap x 4, times cap; *len = true chain length                        */
int krrt_path(krrt_ctx* ctx, int idx, double* states, double* times, int cap, int* len);
int krrt_result_of(krrt_ctx* ctx, int solved, int goal_index, krrt_result* out);
/* fine re-solve of the root -> idx path at step dt: rows (x,y,vx,vy,t) ; *n_rows = rows needed   */
int krrt_trajectory(krrt_ctx* ctx, int idx, double dt, double* rows, int cap_rows, int* n_rows);

/* stateless batched operators on host arrays (n pairs / edges) */
int krrt_steer_cost(krrt_ctx* ctx, const double* xi, const double* xf, int n, double* t, double* c);
int krrt_solve(krrt_ctx* ctx, const double xi[4], const double xf[4], double dt, double t0, double* rows,
               int cap_rows, int* n_rows);
/* krrt_solve plus the control inputs (ax, ay) at the same rows: inputs n x 2.  Serves the 2-D integrator model of the
 * compute_trajectory service (trajectory_generator_node.cpp:6-53 answers "angular_trajectory" only) */
int krrt_solve_inputs(krrt_ctx* ctx, const double xi[4], const double xf[4], double dt, double t0, double* rows,
                      double* inputs, int cap_rows, int* n_rows);
/* Connector slot on its own: the optimal edge of each of n pairs as S rows (x,y,vx,vy,t0+t); rows: n x S x 5 */
int krrt_edge_samples(krrt_ctx* ctx, const double* xi, const double* xf, const double* t0, int n, int S,
                      double* rows);
int krrt_edge_collide(krrt_ctx* ctx, const double* xi, const double* xf, const double* tau, const double* t0,
                      int n, unsigned char* hit);
/* radius query on the current tree: ascending indices, *count = true count, nearest = lowest index on ties */
int krrt_near(krrt_ctx* ctx, const double x[4], double radius2, int32_t* idx, int cap, int* count,
              int32_t* nearest, double* nearest_d2);

/* q independent queries, one warp each, trees in shared memory.  starts/goals q x 4, obstacles q x m x 4.
 * Philox query_id = query_base + i.  Uses ctx's config except capacity (= n_iter + 2).         */
int krrt_plan_batch(krrt_ctx* ctx, const double* starts, const double* goals, const double* obstacles, int q,
                    int m, int dynamic, int n_iter, uint32_t query_base, krrt_result* out);
/* same with device-resident inputs/outputs (pointers from the caller's CUDA allocator, e.g. for
 * timing with inputs already in HBM, or to feed an NCCL gather without a host round trip).  Asynchronous on
 * cuda_stream (the ctx stream when NULL) once the per-context buffers have reached their size (the first call of a
 * given q / n_iter allocates).  ONE batched plan may be in flight per context: work counter, counters and the redo
 * list are per-context scratch — use one context per concurrent stream.                                          */
int krrt_plan_batch_device(krrt_ctx* ctx, const double* d_starts, const double* d_goals,
                           const double* d_obstacles, int q, int m, int dynamic, int n_iter,
                           uint32_t query_base, krrt_result* d_out, void* cuda_stream);
/* same, and the kernels count finished queries per chunk: d_chunk_done[i / chunk_queries] (zeroed by the caller) is
 * incremented once query i's record is visible device-wide, so a consumer on ANOTHER stream can wait for a chunk to reach
 * its size (cuStreamWaitValue32 GEQ) and ship it while the rest of the batch is still being planned — what
 * libkrrt_mgpu's sweep does to overlap the NCCL gather with the planning (include/krrt_mgpu.h)                    */
int krrt_plan_batch_device_chunks(krrt_ctx* ctx, const double* d_starts, const double* d_goals,
                                  const double* d_obstacles, int q, int m, int dynamic, int n_iter,
                                  uint32_t query_base, krrt_result* d_out, int chunk_queries,
                                  uint32_t* d_chunk_done, void* cuda_stream);
/* krrt_plan_batch_device for callers that are not allowed to allocate device memory.  (No counterpart in the reference:
 * its planner is CPU code and never meets the question.)
 *
 * A krrt_ctx gets its device buffers from cudaMalloc — at krrt_create, and again whenever a batch is larger than any
 * before it.  That rules the context-based calls out for (a) a planner inside a robot's fixed-rate control cycle, where a
 * cudaMalloc / cudaFree (driver-wide lock, possibly a wait for the whole device) is a stall of unknown length, and (b) a
 * process whose device memory is owned by another component's pool, where a raw cudaMalloc fails or takes memory the pool
 * relies on.  This call needs no context and allocates nothing, on the host or on the device, on any successful path: the
 * planner's scratch (radius table, work counters, statistics, redo list) is a 264 KB array compiled into libkrrt.so, which
 * every GPU that loads the library holds until the process exits.  Hence the limits:
 *     q <= KRRT_STATIC_MAX_QUERIES and n_iter <= KRRT_STATIC_MAX_ITER per call,
 * beyond which the call returns KRRT_ERR_INVALID before it touches the GPU, and krrt_last_error(NULL) names both limits.
 *
 * cfg: the fields a batched plan reads (device, S, r, R, lo, hi, gamma, rho_max, p_direct, goal_tol, seed, goal_mode,
 * capacity); precision must be KRRT_F64, device < 16.  All pointers are DEVICE memory of cfg->device from the caller's own
 * allocator, laid out as for krrt_plan_batch_device; the records in d_out are the ones that call writes, byte for byte.
 * The work is queued on cuda_stream (NULL = the default stream; there is no context stream) and the call returns without
 * waiting: synchronise the stream before reading d_out.
 *
 * - The first call on each device also looks up where that device keeps the array and loads the kernels; make it during
 *   start-up, outside the time-critical loop and before any stream capture begins.
 * - There is ONE array per device and every call uses it.  Each call clears the counters it needs at zero and rewrites the
 *   table it reads, so calls one after the other on a stream are independent; two calls whose GPU work overlaps (different
 *   streams or threads without an event between them) race on it and return garbage.  Order them.
 * - No krrt_batch_stats / krrt_launch_count: there is no context to ask (each call launches 2 kernels).                 */
#define KRRT_STATIC_MAX_QUERIES 65536
#define KRRT_STATIC_MAX_ITER 1024
int krrt_plan_batch_static(const krrt_config* cfg, const double* d_starts, const double* d_goals,
                           const double* d_obstacles, int q, int m, int dynamic, int n_iter, uint32_t query_base,
                           krrt_result* d_out, void* cuda_stream);
/* Yaw (heading) profiles for q independent queries: (w, dw) = current heading / rate, yaw = target heading, limits
 * v_max / a_max (Controller::setLimit), n sample rows over [0, final_time] as getAngleTrajectory(ctrl, offset, 0, t, n,
 * &inputs) returns them.  states: q x max(n,1) x 2 (w wrapped to [-pi, pi], dw), times / inputs: q x max(n,1),
 * status: 0 ok, 1 = the reference's "case not match!" exception, 2 = its "control sequence error!" (rows are zero),
 * n_seg: entries of the control sequence incl. the initial state.  Bit-identical with the reference library.   */
int krrt_yaw_profiles(krrt_ctx* ctx, const double* w, const double* dw, const double* yaw, int q, double v_max,
                      double a_max, int n, double* states, double* times, double* inputs, double* final_time,
                      int32_t* status, int32_t* n_seg);
/* same with device-resident buffers (asynchronous on cuda_stream, or the ctx stream when NULL)                 */
int krrt_yaw_profiles_device(krrt_ctx* ctx, const double* d_w, const double* d_dw, const double* d_yaw, int q,
                             double v_max, double a_max, int n, double* d_states, double* d_times, double* d_inputs,
                             double* d_final_time, int32_t* d_status, int32_t* d_n_seg, void* cuda_stream);
/* Obstacle ingest for f frames at once: OminiVisionInfo robotinfo[] in nubot units (float32 cm, cm/s:
 * nubot_common/msgs/RobotInfo.msg:2,9-12) -> the robot's own state and the obstacle rows (x,y,vx,vy) in metres that
 * krrt_set_obstacles / krrt_plan_batch take.  Replaces RobotSubscriber::callback (robotsubscriber.cpp:44-77: own team,
 * x / 100.), ::rival_callback (robotsubscriber.cpp:29-42: rivals sign-flipped, -x * 1e-2) and ::getObstacles
 * (robotsubscriber.h:50-55: friends without the robot itself, then rivals).
 * team f x n_team x 5 (pos.x, pos.y, vtrans.x, vtrans.y, heading.theta), agent_id f x n_team, rival f x n_rival x 4,
 * self_id f.  Out: self_state f x 4 (zeros when no entry carries self_id), heading f, self_index f (-1 = absent,
 * else the LAST matching entry), obstacles f x (n_team + n_rival) x 4 (unused rows zero), n_obs f.               */
int krrt_ingest_vision(krrt_ctx* ctx, const float* team, const int32_t* agent_id, int n_team, const float* rival,
                       int n_rival, const int32_t* self_id, int f, double* self_state, double* heading,
                       int32_t* self_index, double* obstacles, int32_t* n_obs);
/* same with device-resident buffers (16-byte aligned; asynchronous on cuda_stream, or the ctx stream when NULL):
 * d_obstacles can be handed to krrt_plan_batch_device as is                                                       */
int krrt_ingest_vision_device(krrt_ctx* ctx, const float* d_team, const int32_t* d_agent_id, int n_team,
                              const float* d_rival, int n_rival, const int32_t* d_self_id, int f,
                              double* d_self_state, double* d_heading, int32_t* d_self_index, double* d_obstacles,
                              int32_t* d_n_obs, void* cuda_stream);
/* oracle-comparable counters summed over the queries of the last batched plan                    */
int krrt_batch_stats(krrt_ctx* ctx, krrt_stats* out);
/* integration / benchmarking helpers: page-locked host memory (cudaHostAlloc: copies at full PCIe rate, no staging),
 * and a cold-cache step for timing loops: overwrites a buffer twice the L2's size on the ctx stream and waits      */
int krrt_host_alloc(void** p, uint64_t bytes);
int krrt_host_free(void* p);
int krrt_flush_l2(krrt_ctx* ctx);
int krrt_synchronize(krrt_ctx* ctx);
/* kernels launched by this ctx since creation (bench.py's gpu_launches)                          */
int64_t krrt_launch_count(const krrt_ctx* ctx);
/* CUDA-event stopwatch on the ctx stream spanning any sequence of calls (end-to-end timing incl. copies) */
int krrt_timer_begin(krrt_ctx* ctx);
int krrt_timer_end(krrt_ctx* ctx, double* ms);
/* profile mode: CUDA events around every neighbour-scan launch of krrt_grow (adds a little launch gap)  */
int krrt_set_profile(krrt_ctx* ctx, int on);
/* krrt_set_profile(ctx, 2): an event after every kernel of every round of krrt_grow (first 4096 rounds).
 * ms / launches: [1] = k_round_scan, [3] = k_round_decide, [4] = k_round_validate, [5] = k_round_apply (others 0);
 * [7] = gaps between rounds (host read-back of the progress counter + enqueue)                           */
int krrt_kernel_times(krrt_ctx* ctx, double ms[8], int64_t launches[8]);
/* since krrt_set_start: out[0] = rounds whose validated prefix was committed in parallel, out[1] = rounds (or
 * single samples) committed by the sequential warp, out[2] = expansions committed in parallel; out[3..7] = what
 * ended the parallel prefixes: end of batch, in-batch neighbour, sample needing the sequential path, sample with
 * no collision-free parent, read/write conflict                                                             */
int krrt_wave_stats(krrt_ctx* ctx, int64_t out[8]);
/* of out[1] above: samples that needed the in-order path right behind a parallel prefix and were committed by the
 * sequential warp inside the same sub-round (one extra grid barrier instead of a sub-round of their own)        */
int krrt_tail_commits(krrt_ctx* ctx, int64_t* out);
/* development builds (-DKRRT_DEBUG_CLK): SM-clock sums of the steps inside decide / apply of CTA 0; zeros otherwise */
int krrt_debug_clocks(krrt_ctx* ctx, int64_t out[16]);
/* device time of the last krrt_grow / krrt_plan_batch in ms (CUDA events on the ctx stream), and the share
 * spent in the neighbour-scan kernel (for the roofline line).                                   */
int krrt_last_timing(krrt_ctx* ctx, double* total_ms, double* scan_ms, int64_t* scan_launches,
                     double* scan_node_samples);
/* Neighbour scan of the last krrt_grow call: (node, sample) pairs really evaluated (scan_node_samples above is the
 * brute-force equivalent), rebuilds of the spatial index since krrt_set_start, nodes it covers now.  */
int krrt_scan_stats(krrt_ctx* ctx, double* visited_pairs, int64_t* rebuilds, int32_t* indexed_nodes);

#ifdef __cplusplus
}
#endif
#endif
