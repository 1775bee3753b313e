// krrt_f32_api.h — what krrt_capi.cu calls in the binary32 build of the batched planner and of the stateless
// Cost / Environment operators (krrt_f32.cu; KRRT_F32 contexts of include/krrt.h).  Plain types only: this header is
// seen by krrt_f32.cu BEFORE it redefines `double`, so every `double` below is binary64 in both translation units.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "../../include/krrt.h"

namespace krrt_f32_api {

struct BatchArgs {              // one batched plan: device buffers of the C ABI (binary64) + the context's scratch
  const double* starts;         // q x 4
  const double* goals;          // q x 4
  const double* obstacles;      // q x m x 4
  void* out;                    // q x krrt_result
  int* next_query;              // work counter (zeroed by the caller)
  const int* list;              // redo pass: queries list[0 .. *list_count)
  const int* list_count;
  long long* stats;             // [10]
  const double* rho2;           // [cap + 1] squared near radii, binary32 values held in binary64
  int* redo_list;
  int* redo_count;
  unsigned* chunk_done;
  int chunk_queries;
  int q, m, dynamic, n_iter, cap, S, goal_mode;
  uint32_t query_base;
  double r, R, goal_tol;        // rounded to binary32 by the launcher (R2 = R * R and tol^2 in binary32)
  double lo[4], hi[4], p_direct;
  uint64_t seed;
};

size_t batch2_warp_smem_bytes(int cap, int m);
size_t batch_warp_smem_bytes(int cap, int m);
int batch2_max_warps();
// k_plan_batch2 / k_plan_batch of namespace krrt_f32; return a cudaError_t as int
int launch_batch2(const BatchArgs& a, int grid, int warps, size_t smem, cudaStream_t st);
int launch_batch(const BatchArgs& a, int grid, int warps, size_t smem, cudaStream_t st);
// squared near radius of a tree of n nodes, evaluated in binary32 exactly as the CPU checker's float build does
double near_radius2(double gamma, double rho_max, int n);
// stateless operators: device arrays in binary64 (as the C ABI hands them over), arithmetic in binary32
int steer_batch(const double* d_xi, const double* d_xf, int n, double r, double* d_t, double* d_c, cudaStream_t st);
int edge_collide_batch(const double* d_xi, const double* d_xf, const double* d_tau, const double* d_t0, int n,
                       const double* d_obs /* m x 4 */, int m, int dynamic, double R, int S, unsigned char* d_hit,
                       cudaStream_t st);

// Connector slot (solver.solve(xi, xf, dt), the stored edge samples): rows of the edges states[e] -> states[e + 1]
int traj_count(const double* d_states, int n_edges, double r, double dt, double* d_tau, int* d_rows, cudaStream_t st);
int traj_fill(const double* d_states, const double* d_t0, const double* d_tau, const int* d_offs, int n_edges, double dt,
              double* d_out, int cap_rows, double* d_inputs /* nullable */, cudaStream_t st);
int edge_samples(const double* d_xi, const double* d_xf, const double* d_t0, int n, int S, double r, double* d_rows,
                 cudaStream_t st);

// ---- the single-tree planner in binary32 (k_grow and the kernels around it compiled in float; krrt_f32.cu).  One Tree per
// KRRT_F32 context: krrt_capi.cu forwards the planner's entry points here.  Arrays are the C ABI's (binary64): rounded on the
// way in, widened on the way out.  Every function returns a KRRT_* status; tree_error() gives the text.
struct Tree;
int tree_create(const krrt_config* cfg, Tree** out, char* err, size_t err_len);
void tree_destroy(Tree* t);
const char* tree_error(const Tree* t);
int tree_set_params(Tree* t, const krrt_config* cfg);
int tree_set_obstacles(Tree* t, const double* xyvxvy, int m, int dynamic);
int tree_set_start(Tree* t, const double x0[4]);
int tree_grow(Tree* t, const double* goal, int n_iter, const double* samples, int* solved, int* goal_index);
// krrt_grow_logged: records != nullptr (1 <= n_iter <= 256, tree size after the call <= 1024)
int tree_grow_logged(Tree* t, const double* goal, int n_iter, const double* samples, krrt_iter_record* records, int* solved,
                     int* goal_index);
int tree_grow_done(Tree* t, int64_t* iterations_done, int* full, double* ms);
// krrt_near (radius2 and the returned squared distance in the float metric, widened)
int tree_near(Tree* t, const double x[4], double radius2, int32_t* idx, int cap, int* count, int32_t* nearest,
              double* nearest_d2);
int tree_info(Tree* t, int* n, int64_t* iteration, double last_sample[4], krrt_stats* stats, int* goal_index);
int tree_read(Tree* t, int first, int count, double* states, int32_t* parent, double* cost, double* time);
int tree_path(Tree* t, int idx, double* states, double* times, int cap, int* len);
int64_t tree_launches(const Tree* t);

}  // namespace krrt_f32_api
