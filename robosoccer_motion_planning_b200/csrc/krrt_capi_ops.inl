// krrt_capi_ops.inl — tree read-back, stateless operators and the batched planner entry points of include/krrt.h.
// Included at the end of krrt_capi.cu (same translation unit, same -fmad=false arithmetic contract).

static int current_size(krrt_ctx* ctx, int* n, int* n_with_goal = nullptr) {
  Ctl h;
  int rc = fetch_ctl(ctx, &h);
  if (rc != KRRT_OK) return rc;
  *n = h.n;
  // the goal node of KRRT_GOAL_CONNECT sits right behind the tree (slot n): the index-taking getters accept it
  if (n_with_goal) *n_with_goal = h.n + (h.goal_index == h.n ? 1 : 0);
  return KRRT_OK;
}

extern "C" {

int krrt_tree_download(krrt_ctx* ctx, double* states, int32_t* parent, double* cost, double* time) {
  if (!ctx) return KRRT_ERR_INVALID;
  if (ctx->cfg.precision == KRRT_F32) {
    int n32 = 0;
    int rc32 = krrt_f32_api::tree_info(ctx->f32, &n32, nullptr, nullptr, nullptr, nullptr);
    if (rc32 == KRRT_OK) rc32 = krrt_f32_api::tree_read(ctx->f32, 0, n32, states, parent, cost, time);
    if (rc32 != KRRT_OK) ctx->err = krrt_f32_api::tree_error(ctx->f32);
    return rc32;
  }
  REQUIRE(ctx->started, KRRT_ERR_STATE, "no tree: call krrt_set_start");
  CK(cudaSetDevice(ctx->cfg.device));
  int n = 0;
  int rc = current_size(ctx, &n);
  if (rc != KRRT_OK) return rc;
  cudaStream_t st = ctx->stream;
  if (states) {
    // interleave (pos, vel) -> n x 4 rows with two strided copies
    CK(cudaMemcpy2DAsync(states, 32, ctx->T.pos, 16, 16, n, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpy2DAsync(states + 2, 32, ctx->T.vel, 16, 16, n, cudaMemcpyDeviceToHost, st));
  }
  if (parent) CK(cudaMemcpyAsync(parent, ctx->T.parent, sizeof(int) * n, cudaMemcpyDeviceToHost, st));
  if (cost) CK(cudaMemcpyAsync(cost, ctx->T.ctc, sizeof(double) * n, cudaMemcpyDeviceToHost, st));
  if (time) CK(cudaMemcpyAsync(time, ctx->T.tim, sizeof(double) * n, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return KRRT_OK;
}

int krrt_tree_read(krrt_ctx* ctx, int first, int count, double* states, int32_t* parent, double* cost, double* time) {
  if (!ctx) return KRRT_ERR_INVALID;
  F32_FORWARD(krrt_f32_api::tree_read(ctx->f32, first, count, states, parent, cost, time));
  REQUIRE(ctx->started, KRRT_ERR_STATE, "no tree: call krrt_set_start");
  CK(cudaSetDevice(ctx->cfg.device));
  int n = 0, ng = 0;
  int rc = current_size(ctx, &n, &ng);
  if (rc != KRRT_OK) return rc;
  REQUIRE(first >= 0 && count >= 0 && first + count <= ng, KRRT_ERR_INVALID, "node range out of bounds");
  if (count == 0) return KRRT_OK;
  cudaStream_t st = ctx->stream;
  if (states) {
    CK(cudaMemcpy2DAsync(states, 32, ctx->T.pos + first, 16, 16, count, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpy2DAsync(states + 2, 32, ctx->T.vel + first, 16, 16, count, cudaMemcpyDeviceToHost, st));
  }
  if (parent) CK(cudaMemcpyAsync(parent, ctx->T.parent + first, sizeof(int) * count, cudaMemcpyDeviceToHost, st));
  if (cost) CK(cudaMemcpyAsync(cost, ctx->T.ctc + first, sizeof(double) * count, cudaMemcpyDeviceToHost, st));
  if (time) CK(cudaMemcpyAsync(time, ctx->T.tim + first, sizeof(double) * count, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return KRRT_OK;
}

int krrt_goal_index(krrt_ctx* ctx, int* goal_index) {
  if (!ctx || !goal_index) return KRRT_ERR_INVALID;
  F32_FORWARD(krrt_f32_api::tree_info(ctx->f32, nullptr, nullptr, nullptr, nullptr, goal_index));
  REQUIRE(ctx->started, KRRT_ERR_STATE, "no tree: call krrt_set_start");
  CK(cudaSetDevice(ctx->cfg.device));
  Ctl h;
  int rc = fetch_ctl(ctx, &h);
  if (rc != KRRT_OK) return rc;
  *goal_index = h.goal_index;
  return KRRT_OK;
}

int krrt_path(krrt_ctx* ctx, int idx, double* states, double* times, int cap, int* len) {
  if (!ctx || !len) return KRRT_ERR_INVALID;
  F32_FORWARD(krrt_f32_api::tree_path(ctx->f32, idx, states, times, cap, len));
  REQUIRE(ctx->started, KRRT_ERR_STATE, "no tree: call krrt_set_start");
  REQUIRE(cap >= 0, KRRT_ERR_INVALID, "cap < 0");
  CK(cudaSetDevice(ctx->cfg.device));
  int n = 0, ng = 0;
  int rc = current_size(ctx, &n, &ng);
  if (rc != KRRT_OK) return rc;
  REQUIRE(idx >= 0 && idx < ng, KRRT_ERR_INVALID, "node index out of range");
  n = ng;
  cudaStream_t st = ctx->stream;
  const size_t c = (size_t)std::max(cap, 1);
  if ((rc = ensure_scratch(ctx, 0, sizeof(double) * 4 * c)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 1, sizeof(double) * c)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 2, sizeof(int) * 4)) != KRRT_OK) return rc;
  double* d_states = (double*)ctx->d_scratch[0];
  double* d_times = (double*)ctx->d_scratch[1];
  int* d_len = (int*)ctx->d_scratch[2];
  k_path_extract<<<1, 32, 0, st>>>(ctx->T, n, idx, d_states, d_times, cap, d_len);
  ctx->launches++;
  CK(cudaGetLastError());
  int h_len = 0;
  CK(cudaMemcpyAsync(&h_len, d_len, sizeof(int), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  const int got = std::min(h_len, cap);
  if (states && got > 0) CK(cudaMemcpyAsync(states, d_states, sizeof(double) * 4 * got, cudaMemcpyDeviceToHost, st));
  if (times && got > 0) CK(cudaMemcpyAsync(times, d_times, sizeof(double) * got, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  *len = h_len;
  return KRRT_OK;
}

int krrt_result_of(krrt_ctx* ctx, int solved, int goal_index, krrt_result* out) {
  if (!ctx || !out) return KRRT_ERR_INVALID;
  if (ctx->cfg.precision == KRRT_F32) {  // the same record from the binary32 tree (path + the goal node's cost and arrival time)
    std::memset(out, 0, sizeof(*out));
    int n32 = 0;
    int rc32 = krrt_tree_size(ctx, &n32);
    if (rc32 != KRRT_OK) return rc32;
    out->solved = solved;
    out->goal_index = goal_index;
    out->n_nodes = n32;
    if (!solved) return KRRT_OK;
    double st4[KRRT_MAX_WAYPOINTS][4], tm[KRRT_MAX_WAYPOINTS];
    int len = 0;
    if ((rc32 = krrt_path(ctx, goal_index, &st4[0][0], tm, KRRT_MAX_WAYPOINTS, &len)) != KRRT_OK) return rc32;
    out->path_len = len;
    for (int k = 0; k < std::min(len, (int)KRRT_MAX_WAYPOINTS); ++k) {
      for (int d = 0; d < 4; ++d) out->waypoints[k][d] = st4[k][d];
      out->waypoints[k][4] = tm[k];
    }
    return krrt_tree_read(ctx, goal_index, 1, nullptr, nullptr, &out->cost, &out->time);
  }
  REQUIRE(ctx->started, KRRT_ERR_STATE, "no tree: call krrt_set_start");
  std::memset(out, 0, sizeof(*out));
  int n = 0;
  CK(cudaSetDevice(ctx->cfg.device));
  int rc = current_size(ctx, &n);
  if (rc != KRRT_OK) return rc;
  out->solved = solved;
  out->goal_index = goal_index;
  out->n_nodes = n;
  if (!solved) return KRRT_OK;
  double st4[KRRT_MAX_WAYPOINTS][4];
  double tm[KRRT_MAX_WAYPOINTS];
  int len = 0;
  rc = krrt_path(ctx, goal_index, &st4[0][0], tm, KRRT_MAX_WAYPOINTS, &len);
  if (rc != KRRT_OK) return rc;
  out->path_len = len;
  for (int k = 0; k < std::min(len, (int)KRRT_MAX_WAYPOINTS); ++k) {
    for (int d = 0; d < 4; ++d) out->waypoints[k][d] = st4[k][d];
    out->waypoints[k][4] = tm[k];
  }
  CK(cudaMemcpyAsync(&out->cost, ctx->T.ctc + goal_index, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaMemcpyAsync(&out->time, ctx->T.tim + goal_index, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return KRRT_OK;
}

// rows of every edge of a device-resident chain of `len` states
static int trajectory_of_chain(krrt_ctx* ctx, const double* d_states, const double* d_t0, int len, double dt,
                               double* rows, int cap_rows, int* n_rows, double* inputs = nullptr) {
  cudaStream_t st = ctx->stream;
  const int ne = len - 1;
  if (ne <= 0) {
    *n_rows = 0;
    return KRRT_OK;
  }
  int rc;
  if ((rc = ensure_scratch(ctx, 3, sizeof(double) * ne)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 4, sizeof(int) * (2 * ne + 2))) != KRRT_OK) return rc;
  double* d_tau = (double*)ctx->d_scratch[3];
  int* d_rows = (int*)ctx->d_scratch[4];
  int* d_offs = d_rows + ne;
  const bool f32 = ctx->cfg.precision == KRRT_F32;  // the Connector kernels of krrt_f32.cu (rows stay binary64 arrays)
  if (f32) CK((cudaError_t)krrt_f32_api::traj_count(d_states, ne, ctx->cfg.r, dt, d_tau, d_rows, st));
  else k_traj_count<<<(ne + 127) / 128, 128, 0, st>>>(d_states, ne, ctx->cfg.r, dt, d_tau, d_rows);
  k_traj_offsets<<<1, 32, 0, st>>>(d_rows, ne, d_offs);
  ctx->launches += 2;
  int total = 0;
  CK(cudaMemcpyAsync(&total, d_offs + ne, sizeof(int), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  *n_rows = total;
  const int got = std::min(total, cap_rows);
  if (!rows || got <= 0) return KRRT_OK;
  if ((rc = ensure_scratch(ctx, 5, sizeof(double) * 7 * (size_t)got)) != KRRT_OK) return rc;
  double* d_out = (double*)ctx->d_scratch[5];
  double* d_inp = inputs ? d_out + 5 * (size_t)got : nullptr;
  if (f32) CK((cudaError_t)krrt_f32_api::traj_fill(d_states, d_t0, d_tau, d_offs, ne, dt, d_out, got, d_inp, st));
  else k_traj_fill<<<ne, 128, 0, st>>>(d_states, d_t0, d_tau, d_offs, dt, d_out, got, d_inp);
  ctx->launches++;
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(rows, d_out, sizeof(double) * 5 * (size_t)got, cudaMemcpyDeviceToHost, st));
  if (inputs) CK(cudaMemcpyAsync(inputs, d_inp, sizeof(double) * 2 * (size_t)got, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return KRRT_OK;
}

int krrt_trajectory(krrt_ctx* ctx, int idx, double dt, double* rows, int cap_rows, int* n_rows) {
  if (!ctx || !n_rows) return KRRT_ERR_INVALID;
  if (ctx->cfg.precision == KRRT_F32) {  // the chain root -> idx of the binary32 tree, re-solved by the float Connector kernels
    REQUIRE(dt > 0.0 && cap_rows >= 0, KRRT_ERR_INVALID, "dt must be > 0");
    CK(cudaSetDevice(ctx->cfg.device));
    int n32 = 0;
    int rc32 = krrt_f32_api::tree_info(ctx->f32, &n32, nullptr, nullptr, nullptr, nullptr);
    std::vector<double> st(4 * (size_t)(n32 + 1)), tm((size_t)n32 + 1);
    int len = 0;
    if (rc32 == KRRT_OK) rc32 = krrt_f32_api::tree_path(ctx->f32, idx, st.data(), tm.data(), n32 + 1, &len);
    if (rc32 != KRRT_OK) {
      ctx->err = krrt_f32_api::tree_error(ctx->f32);
      return rc32;
    }
    if (len <= 0) {
      *n_rows = 0;
      return KRRT_OK;
    }
    if ((rc32 = ensure_scratch(ctx, 0, sizeof(double) * 4 * (size_t)len)) != KRRT_OK) return rc32;
    if ((rc32 = ensure_scratch(ctx, 1, sizeof(double) * (size_t)len)) != KRRT_OK) return rc32;
    CK(cudaMemcpyAsync(ctx->d_scratch[0], st.data(), sizeof(double) * 4 * (size_t)len, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(ctx->d_scratch[1], tm.data(), sizeof(double) * (size_t)len, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return trajectory_of_chain(ctx, (const double*)ctx->d_scratch[0], (const double*)ctx->d_scratch[1], len, dt, rows, cap_rows,
                               n_rows);
  }
  REQUIRE(ctx->started, KRRT_ERR_STATE, "no tree: call krrt_set_start");
  REQUIRE(dt > 0.0 && cap_rows >= 0, KRRT_ERR_INVALID, "dt must be > 0");
  CK(cudaSetDevice(ctx->cfg.device));
  int n = 0, ng = 0;
  int rc = current_size(ctx, &n, &ng);
  if (rc != KRRT_OK) return rc;
  REQUIRE(idx >= 0 && idx < ng, KRRT_ERR_INVALID, "node index out of range");
  n = ng;
  cudaStream_t st = ctx->stream;
  if ((rc = ensure_scratch(ctx, 0, sizeof(double) * 4 * (size_t)n)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 1, sizeof(double) * (size_t)n)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 2, sizeof(int) * 4)) != KRRT_OK) return rc;
  double* d_states = (double*)ctx->d_scratch[0];
  double* d_times = (double*)ctx->d_scratch[1];
  int* d_len = (int*)ctx->d_scratch[2];
  k_path_extract<<<1, 32, 0, st>>>(ctx->T, n, idx, d_states, d_times, n, d_len);
  ctx->launches++;
  int len = 0;
  CK(cudaMemcpyAsync(&len, d_len, sizeof(int), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return trajectory_of_chain(ctx, d_states, d_times, len, dt, rows, cap_rows, n_rows);
}

int krrt_solve(krrt_ctx* ctx, const double xi[4], const double xf[4], double dt, double t0, double* rows,
               int cap_rows, int* n_rows) {
  if (!ctx || !xi || !xf || !n_rows) return KRRT_ERR_INVALID;
  REQUIRE(dt > 0.0 && cap_rows >= 0, KRRT_ERR_INVALID, "dt must be > 0");
  CK(cudaSetDevice(ctx->cfg.device));
  int rc;
  if ((rc = ensure_scratch(ctx, 0, sizeof(double) * 8)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 1, sizeof(double) * 2)) != KRRT_OK) return rc;
  double h[8] = {xi[0], xi[1], xi[2], xi[3], xf[0], xf[1], xf[2], xf[3]};
  CK(cudaMemcpyAsync(ctx->d_scratch[0], h, sizeof(h), cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(ctx->d_scratch[1], &t0, sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return trajectory_of_chain(ctx, (const double*)ctx->d_scratch[0], (const double*)ctx->d_scratch[1], 2, dt, rows,
                             cap_rows, n_rows);
}

int krrt_solve_inputs(krrt_ctx* ctx, const double xi[4], const double xf[4], double dt, double t0, double* rows,
                      double* inputs, int cap_rows, int* n_rows) {
  if (!ctx || !xi || !xf || !n_rows) return KRRT_ERR_INVALID;
  REQUIRE(dt > 0.0 && cap_rows >= 0, KRRT_ERR_INVALID, "dt must be > 0");
  REQUIRE(cap_rows == 0 || (rows && inputs), KRRT_ERR_INVALID, "null buffer");
  CK(cudaSetDevice(ctx->cfg.device));
  int rc;
  if ((rc = ensure_scratch(ctx, 0, sizeof(double) * 8)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 1, sizeof(double) * 2)) != KRRT_OK) return rc;
  double h[8] = {xi[0], xi[1], xi[2], xi[3], xf[0], xf[1], xf[2], xf[3]};
  CK(cudaMemcpyAsync(ctx->d_scratch[0], h, sizeof(h), cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(ctx->d_scratch[1], &t0, sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return trajectory_of_chain(ctx, (const double*)ctx->d_scratch[0], (const double*)ctx->d_scratch[1], 2, dt, rows,
                             cap_rows, n_rows, inputs);
}

int krrt_steer_cost(krrt_ctx* ctx, const double* xi, const double* xf, int n, double* t, double* c) {
  if (!ctx) return KRRT_ERR_INVALID;
  REQUIRE(n >= 0 && (n == 0 || (xi && xf && t && c)), KRRT_ERR_INVALID, "null buffer");
  if (n == 0) return KRRT_OK;
  CK(cudaSetDevice(ctx->cfg.device));
  cudaStream_t st = ctx->stream;
  int rc;
  const size_t sb = sizeof(double) * 4 * (size_t)n, ob = sizeof(double) * (size_t)n;
  if ((rc = ensure_scratch(ctx, 0, sb)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 1, sb)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 3, ob)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 5, ob)) != KRRT_OK) return rc;
  CK(cudaMemcpyAsync(ctx->d_scratch[0], xi, sb, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(ctx->d_scratch[1], xf, sb, cudaMemcpyHostToDevice, st));
  if (ctx->cfg.precision == KRRT_F32)
    CK((cudaError_t)krrt_f32_api::steer_batch((const double*)ctx->d_scratch[0], (const double*)ctx->d_scratch[1], n, ctx->cfg.r,
                                              (double*)ctx->d_scratch[3], (double*)ctx->d_scratch[5], st));
  else
    k_steer_batch<<<(n + 127) / 128, 128, 0, st>>>((const double*)ctx->d_scratch[0], (const double*)ctx->d_scratch[1], n,
                                                   ctx->cfg.r, (double*)ctx->d_scratch[3], (double*)ctx->d_scratch[5]);
  ctx->launches++;
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(t, ctx->d_scratch[3], ob, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(c, ctx->d_scratch[5], ob, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return KRRT_OK;
}

int krrt_edge_collide(krrt_ctx* ctx, const double* xi, const double* xf, const double* tau, const double* t0, int n,
                      unsigned char* hit) {
  if (!ctx) return KRRT_ERR_INVALID;
  REQUIRE(n >= 0 && (n == 0 || (xi && xf && tau && t0 && hit)), KRRT_ERR_INVALID, "null buffer");
  if (n == 0) return KRRT_OK;
  CK(cudaSetDevice(ctx->cfg.device));
  cudaStream_t st = ctx->stream;
  int rc;
  const size_t sb = sizeof(double) * 4 * (size_t)n, ob = sizeof(double) * (size_t)n;
  if ((rc = ensure_scratch(ctx, 0, sb)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 1, sb)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 3, ob)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 5, ob)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 6, (size_t)n)) != KRRT_OK) return rc;
  CK(cudaMemcpyAsync(ctx->d_scratch[0], xi, sb, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(ctx->d_scratch[1], xf, sb, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(ctx->d_scratch[3], tau, ob, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(ctx->d_scratch[5], t0, ob, cudaMemcpyHostToDevice, st));
  const long long threads = (long long)n * 32;
  if (ctx->cfg.precision == KRRT_F32)
    CK((cudaError_t)krrt_f32_api::edge_collide_batch(
        (const double*)ctx->d_scratch[0], (const double*)ctx->d_scratch[1], (const double*)ctx->d_scratch[3],
        (const double*)ctx->d_scratch[5], n, reinterpret_cast<const double*>(ctx->d_obs), ctx->P.m, ctx->P.dynamic, ctx->cfg.R,
        ctx->P.S, (unsigned char*)ctx->d_scratch[6], st));
  else
    k_edge_collide_batch<<<(unsigned)((threads + 127) / 128), 128, 0, st>>>(
        (const double*)ctx->d_scratch[0], (const double*)ctx->d_scratch[1], (const double*)ctx->d_scratch[3],
        (const double*)ctx->d_scratch[5], n, ctx->d_obs, ctx->P.m, ctx->P.dynamic, ctx->P.R2, ctx->P.S,
        (unsigned char*)ctx->d_scratch[6]);
  ctx->launches++;
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(hit, ctx->d_scratch[6], (size_t)n, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return KRRT_OK;
}

int krrt_near(krrt_ctx* ctx, const double x[4], double radius2, int32_t* idx, int cap, int* count, int32_t* nearest,
              double* nearest_d2) {
  if (!ctx || !x) return KRRT_ERR_INVALID;
  F32_FORWARD(krrt_f32_api::tree_near(ctx->f32, x, radius2, idx, cap, count, nearest, nearest_d2));
  REQUIRE(ctx->started, KRRT_ERR_STATE, "no tree: call krrt_set_start");
  REQUIRE(cap >= 0 && (cap == 0 || idx), KRRT_ERR_INVALID, "null buffer");
  CK(cudaSetDevice(ctx->cfg.device));
  int n = 0;
  int rc = current_size(ctx, &n);
  if (rc != KRRT_OK) return rc;
  cudaStream_t st = ctx->stream;
  if ((rc = ensure_scratch(ctx, 4, sizeof(int) * (size_t)std::max(cap, 1))) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 2, 32)) != KRRT_OK) return rc;
  int* d_idx = (int*)ctx->d_scratch[4];
  int* d_cnt = (int*)ctx->d_scratch[2];
  int* d_nearest = d_cnt + 1;
  double* d_nd2 = (double*)(d_cnt + 2);
  State4 s = {x[0], x[1], x[2], x[3]};
  k_near_query<<<1, 1024, 0, st>>>(ctx->T, n, s, radius2, d_idx, cap, d_cnt, d_nearest, d_nd2);
  ctx->launches++;
  CK(cudaGetLastError());
  struct { int cnt, nearest; double nd2; } h;
  CK(cudaMemcpyAsync(&h, d_cnt, sizeof(h), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  const int got = std::min(h.cnt, cap);
  if (got > 0) CK(cudaMemcpyAsync(idx, d_idx, sizeof(int) * got, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  if (count) *count = h.cnt;
  if (nearest) *nearest = h.nearest;
  if (nearest_d2) *nearest_d2 = h.nd2;
  return KRRT_OK;
}

// ---- batched independent queries ---------------------------------------------------------------------
// One batched plan may be in flight per context (work counter, counters and redo list are per-context scratch); the
// near-radius table and the redo list are (re)built on the caller's stream, so the call stays asynchronous once the
// buffers have reached their size (they only grow: the first call of a given q / n_iter allocates).
static int launch_plan_batch(krrt_ctx* ctx, const double* d_starts, const double* d_goals, const double* d_obs, int q,
                             int m, int dynamic, int n_iter, uint32_t query_base, krrt_result* d_out,
                             cudaStream_t st, bool* redo_unserved = nullptr, unsigned* d_chunk_done = nullptr,
                             int chunk_queries = 0) {
  const krrt_config& c = ctx->cfg;
  const int cap = std::min(c.capacity, std::max(n_iter + 1, 2));  // root + one node per iteration at most (+ the goal node's slot)
  REQUIRE(cap >= 2, KRRT_ERR_INVALID, "n_iter / capacity too small");
  const size_t smem_max = ctx->prop.sharedMemPerBlockOptin;            // 227 KB on sm_100
  const size_t smem_sm = ctx->prop.sharedMemPerMultiprocessor;         // 228 KB
  // what fits: k_plan_batch2 (near lists in registers, 16-bit links) and its redo pass k_plan_batch (everything in shared
  // memory, 121 B per node against 74).  Decided BEFORE anything is launched.
  const bool f32 = c.precision == KRRT_F32;  // the binary32 build of the same two kernels (krrt_f32.cu)
  const size_t pw2 = f32 ? krrt_f32_api::batch2_warp_smem_bytes(cap, m) : batch2_warp_smem_bytes(cap, m);
  const int w2 = (int)std::min<size_t>(f32 ? krrt_f32_api::batch2_max_warps() : B2_MAX_WARPS, (smem_max - 1024) / pw2);
  const bool use_v2 = !ctx->batch_v1 && cap <= 65533 && w2 >= 1;
  const size_t pw = f32 ? krrt_f32_api::batch_warp_smem_bytes(cap, m) : batch_warp_smem_bytes(cap, m);
  const bool v1_fits = pw <= smem_max;
  REQUIRE(use_v2 || v1_fits, KRRT_ERR_UNSUPPORTED,
          "n_iter too large for the shared-memory batched planner; use krrt_grow per query");
  // trees too big for the redo pass: k_plan_batch2 alone, and the caller has to look at the redo count afterwards
  REQUIRE(v1_fits || redo_unserved, KRRT_ERR_UNSUPPORTED,
          "n_iter too large for the redo pass of the batched planner on a caller-owned stream; use krrt_plan_batch");
  if (redo_unserved) *redo_unserved = !v1_fits;
  if (ctx->rho2_batch_cap != cap) {
    if (ctx->rho2_batch_alloc < cap + 1) {
      if (ctx->d_rho2_batch) CK(cudaFree(ctx->d_rho2_batch));
      ctx->d_rho2_batch = nullptr;
      ctx->rho2_batch_alloc = 0;
      ctx->rho2_batch_cap = -1;
      CK(cudaMalloc((void**)&ctx->d_rho2_batch, sizeof(double) * (cap + 1)));
      ctx->rho2_batch_alloc = cap + 1;
    }
    ctx->h_rho2_batch.resize(cap + 1);  // (owned by the context: the copy below is asynchronous)
    for (int n = 0; n <= cap; ++n)
      ctx->h_rho2_batch[n] = f32 ? krrt_f32_api::near_radius2(c.gamma, c.rho_max, n) : near_radius2(c, n);
    CK(cudaMemcpyAsync(ctx->d_rho2_batch, ctx->h_rho2_batch.data(), sizeof(double) * (cap + 1), cudaMemcpyHostToDevice, st));
    ctx->rho2_batch_cap = cap;
  }
  BatchJob J;
  J.starts = d_starts; J.goals = d_goals; J.obstacles = d_obs;
  J.out = reinterpret_cast<ResultRec*>(d_out);
  J.next_query = ctx->d_counter;
  J.list = nullptr; J.list_count = nullptr;
  J.stats = ctx->d_batch_stats;
  J.rho2 = ctx->d_rho2_batch;
  J.q = q; J.m = m; J.dynamic = dynamic ? 1 : 0; J.n_iter = n_iter; J.cap = cap;
  J.query_base = query_base;
  J.r = c.r; J.R2 = c.R * c.R; J.goal_tol2 = c.goal_tol * c.goal_tol; J.S = c.S;
  J.goal_mode = c.goal_mode;
  J.chunk_done = d_chunk_done;
  J.chunk_queries = chunk_queries > 0 ? chunk_queries : 1;
  for (int k = 0; k < 4; ++k) { J.sp.lo[k] = c.lo[k]; J.sp.hi[k] = c.hi[k]; }
  J.sp.p_direct = c.p_direct;
  J.sp.seed = c.seed;
  CK(cudaMemsetAsync(ctx->d_counter, 0, sizeof(int), st));
  CK(cudaMemsetAsync(ctx->d_batch_stats, 0, sizeof(long long) * 10, st));
  krrt_f32_api::BatchArgs A;
  if (f32) {
    A.starts = d_starts; A.goals = d_goals; A.obstacles = d_obs; A.out = d_out;
    A.next_query = ctx->d_counter; A.list = nullptr; A.list_count = nullptr;
    A.stats = ctx->d_batch_stats; A.rho2 = ctx->d_rho2_batch;
    A.redo_list = nullptr; A.redo_count = nullptr;
    A.chunk_done = d_chunk_done; A.chunk_queries = J.chunk_queries;
    A.q = q; A.m = m; A.dynamic = J.dynamic; A.n_iter = n_iter; A.cap = cap; A.S = c.S; A.goal_mode = c.goal_mode;
    A.query_base = query_base;
    A.r = c.r; A.R = c.R; A.goal_tol = c.goal_tol;
    for (int k = 0; k < 4; ++k) { A.lo[k] = c.lo[k]; A.hi[k] = c.hi[k]; }
    A.p_direct = c.p_direct; A.seed = c.seed;
  }

  // ---- k_plan_batch2: near lists in registers, 16-bit links, one CTA of up to 16 warps per SM
  if (use_v2) {
    if (ctx->redo_cap < (size_t)q + 2) {
      if (ctx->d_redo) CK(cudaFree(ctx->d_redo));
      ctx->d_redo = nullptr;
      ctx->redo_cap = 0;
      CK(cudaMalloc((void**)&ctx->d_redo, sizeof(int) * ((size_t)q + 2)));
      ctx->redo_cap = (size_t)q + 2;
    }
    CK(cudaMemsetAsync(ctx->d_redo, 0, sizeof(int) * 2, st));
    Batch2Job J2;
    J2.b = J;
    J2.redo_count = ctx->d_redo + 1;
    J2.redo_list = ctx->d_redo + 2;
    // a batch that fits the resident warps in one go is spread over ALL SMs (512 queries: 128 CTAs of 4 warps, not 32 CTAs of
    // 16 — each warp gets four times the issue slots: 1.13 -> 0.94 ms, profiles/r2_wave_fit_s16.txt); a larger batch keeps the
    // most warps that fit, because warps pull queries from a counter and more slots even out the unequal queries
    int w_run = w2;
    if (ctx->b2_wave_fit) {
      const long long sms = ctx->prop.multiProcessorCount;
      if ((long long)q <= sms * w2) w_run = (int)std::max<long long>(1, ((long long)q + sms - 1) / sms);
    }
    const size_t smem2 = (size_t)w_run * pw2;
    int grid2 = ctx->prop.multiProcessorCount;
    grid2 = (int)std::max<long long>(1, std::min<long long>(grid2, ((long long)q + w_run - 1) / w_run));
    if (f32) {
      A.redo_count = J2.redo_count;
      A.redo_list = J2.redo_list;
      CK((cudaError_t)krrt_f32_api::launch_batch2(A, grid2, w_run, smem2, st));
    } else {
      CK(cudaFuncSetAttribute(k_plan_batch2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
      k_plan_batch2<<<grid2, w_run * 32, smem2, st>>>(J2);
    }
    ctx->launches++;
    CK(cudaGetLastError());
    // queries whose near set outgrew the registers (none on the soccer-field configurations) are planned by k_plan_batch
    J.next_query = ctx->d_redo;
    J.list = ctx->d_redo + 2;
    J.list_count = ctx->d_redo + 1;
    A.next_query = J.next_query; A.list = J.list; A.list_count = J.list_count;
    if (!v1_fits) return KRRT_OK;
  }

  // ---- k_plan_batch: every list in shared memory (the whole batch, or k_plan_batch2's redo list)
  int best_w = 1, best_total = 0, best_ctas = 1;
  for (int w = 1; w <= 8; ++w) {
    size_t per_cta = (size_t)w * pw;
    if (per_cta > smem_max) break;
    int ctas = (int)(smem_sm / (per_cta + 1024));
    ctas = std::min(ctas, 32);
    if (ctas < 1) continue;
    if (w * ctas >= best_total) {
      best_total = w * ctas;
      best_w = w;
      best_ctas = ctas;
    }
  }
  const size_t smem = (size_t)best_w * pw;
  long long warps_needed = q;   // (as the redo pass: every warp exits after one atomic when the list is empty)
  int grid = ctx->prop.multiProcessorCount * best_ctas;
  grid = (int)std::max<long long>(1, std::min<long long>(grid, (warps_needed + best_w - 1) / best_w));
  if (f32) {
    CK((cudaError_t)krrt_f32_api::launch_batch(A, grid, best_w, smem, st));
  } else {
    CK(cudaFuncSetAttribute(k_plan_batch, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_plan_batch<<<grid, best_w * 32, smem, st>>>(J);
  }
  ctx->launches++;
  CK(cudaGetLastError());
  return KRRT_OK;
}

int krrt_plan_batch_device(krrt_ctx* ctx, const double* d_starts, const double* d_goals, const double* d_obstacles,
                           int q, int m, int dynamic, int n_iter, uint32_t query_base, krrt_result* d_out,
                           void* cuda_stream) {
  if (!ctx) return KRRT_ERR_INVALID;
  REQUIRE(q >= 0 && m >= 0 && m <= KRRT_MAX_OBS && n_iter >= 0, KRRT_ERR_INVALID, "bad q / m / n_iter");
  if (q == 0) return KRRT_OK;
  REQUIRE(d_starts && d_goals && d_out && (m == 0 || d_obstacles), KRRT_ERR_INVALID, "null buffer");
  CK(cudaSetDevice(ctx->cfg.device));
  cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : ctx->stream;
  return launch_plan_batch(ctx, d_starts, d_goals, d_obstacles, q, m, dynamic, n_iter, query_base, d_out, st);
}

int krrt_plan_batch_device_chunks(krrt_ctx* ctx, const double* d_starts, const double* d_goals,
                                  const double* d_obstacles, int q, int m, int dynamic, int n_iter, uint32_t query_base,
                                  krrt_result* d_out, int chunk_queries, uint32_t* d_chunk_done, void* cuda_stream) {
  if (!ctx) return KRRT_ERR_INVALID;
  REQUIRE(q >= 0 && m >= 0 && m <= KRRT_MAX_OBS && n_iter >= 0, KRRT_ERR_INVALID, "bad q / m / n_iter");
  REQUIRE(chunk_queries >= 1 && d_chunk_done, KRRT_ERR_INVALID, "chunk_queries >= 1 and a counter array are needed");
  if (q == 0) return KRRT_OK;
  REQUIRE(d_starts && d_goals && d_out && (m == 0 || d_obstacles), KRRT_ERR_INVALID, "null buffer");
  CK(cudaSetDevice(ctx->cfg.device));
  cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : ctx->stream;
  return launch_plan_batch(ctx, d_starts, d_goals, d_obstacles, q, m, dynamic, n_iter, query_base, d_out, st, nullptr,
                           d_chunk_done, chunk_queries);
}

int krrt_plan_batch(krrt_ctx* ctx, const double* starts, const double* goals, const double* obstacles, int q, int m,
                    int dynamic, int n_iter, uint32_t query_base, krrt_result* out) {
  if (!ctx) return KRRT_ERR_INVALID;
  REQUIRE(q >= 0 && m >= 0 && m <= KRRT_MAX_OBS && n_iter >= 0, KRRT_ERR_INVALID, "bad q / m / n_iter");
  if (q == 0) return KRRT_OK;
  REQUIRE(starts && goals && out && (m == 0 || obstacles), KRRT_ERR_INVALID, "null buffer");
  CK(cudaSetDevice(ctx->cfg.device));
  cudaStream_t st = ctx->stream;
  int rc;
  const size_t sb = sizeof(double) * 4 * (size_t)q, ob = sizeof(double) * 4 * (size_t)q * std::max(m, 1);
  const size_t rb = sizeof(krrt_result) * (size_t)q;
  if ((rc = ensure_scratch(ctx, 0, sb)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 1, sb)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 6, ob)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 7, rb)) != KRRT_OK) return rc;
  CK(cudaEventRecord(ctx->ev0, st));
  CK(cudaMemcpyAsync(ctx->d_scratch[0], starts, sb, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(ctx->d_scratch[1], goals, sb, cudaMemcpyHostToDevice, st));
  if (m > 0) CK(cudaMemcpyAsync(ctx->d_scratch[6], obstacles, sizeof(double) * 4 * (size_t)q * m, cudaMemcpyHostToDevice, st));
  bool redo_unserved = false;
  rc = launch_plan_batch(ctx, (const double*)ctx->d_scratch[0], (const double*)ctx->d_scratch[1],
                         (const double*)ctx->d_scratch[6], q, m, dynamic, n_iter, query_base,
                         (krrt_result*)ctx->d_scratch[7], st, &redo_unserved);
  if (rc != KRRT_OK) return rc;
  CK(cudaMemcpyAsync(out, ctx->d_scratch[7], rb, cudaMemcpyDeviceToHost, st));
  CK(cudaEventRecord(ctx->ev1, st));
  CK(cudaStreamSynchronize(st));
  if (redo_unserved) {  // trees too big for the redo pass: fine as long as no query needed it
    int redo = 0;
    CK(cudaMemcpy(&redo, ctx->d_redo + 1, sizeof(int), cudaMemcpyDeviceToHost));
    REQUIRE(redo == 0, KRRT_ERR_NEAR_OVERFLOW,
            "a near set outgrew k_plan_batch2's 64 candidates and n_iter is too large for its redo pass; use krrt_grow per query");
  }
  float ms = 0.f;
  CK(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
  ctx->last_total_ms = ms;
  return KRRT_OK;
}

int krrt_batch_stats(krrt_ctx* ctx, krrt_stats* out) {
  if (!ctx || !out) return KRRT_ERR_INVALID;
  CK(cudaSetDevice(ctx->cfg.device));
  CK(cudaMemcpyAsync(out, ctx->d_batch_stats, sizeof(krrt_stats), cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return KRRT_OK;
}

int krrt_set_params(krrt_ctx* ctx, const krrt_config* nc) {
  if (!ctx || !nc) return KRRT_ERR_INVALID;
  krrt_config c = *nc;
  if (c.batch == 0) c.batch = ctx->cfg.batch;
  if (c.near_cap == 0) c.near_cap = ctx->cfg.near_cap;
  REQUIRE(c.device == ctx->cfg.device && c.precision == ctx->cfg.precision && c.capacity == ctx->cfg.capacity &&
              c.batch == ctx->cfg.batch && c.near_cap == ctx->cfg.near_cap,
          KRRT_ERR_INVALID, "device / precision / capacity / batch / near_cap are create-time fields");
  REQUIRE(c.S >= 1 && c.r > 0.0 && c.R >= 0.0, KRRT_ERR_INVALID, "bad S / r / R");
  REQUIRE(c.goal_mode == KRRT_GOAL_CONNECT || c.goal_mode == KRRT_GOAL_TOLERANCE, KRRT_ERR_INVALID, "bad goal_mode");
  CK(cudaSetDevice(c.device));
  if (c.gamma != ctx->cfg.gamma || c.rho_max != ctx->cfg.rho_max) {
    const size_t cap = (size_t)c.capacity;
    std::vector<double> rho2(cap + 1);
    for (size_t n = 0; n <= cap; ++n) rho2[n] = near_radius2(c, (int)n);
    CK(cudaMemcpyAsync(ctx->d_rho2, rho2.data(), sizeof(double) * (cap + 1), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->rho2_batch_cap = -1;  // the batched planner's table is rebuilt on its next call
  }
  ctx->cfg = c;
  Params& P = ctx->P;
  P.r = c.r;
  P.R2 = c.R * c.R;
  P.S = c.S;
  for (int k = 0; k < 4; ++k) {
    P.sp.lo[k] = c.lo[k];
    P.sp.hi[k] = c.hi[k];
  }
  P.sp.p_direct = c.p_direct;
  P.sp.seed = c.seed;
  P.query_id = c.query_id;
  P.goal_tol2 = c.goal_tol * c.goal_tol;
  P.goal_mode = c.goal_mode;
  if (ctx->f32) {
    const int rc = krrt_f32_api::tree_set_params(ctx->f32, &ctx->cfg);
    if (rc != KRRT_OK) ctx->err = krrt_f32_api::tree_error(ctx->f32);
    return rc;
  }
  return KRRT_OK;
}

int krrt_edge_samples(krrt_ctx* ctx, const double* xi, const double* xf, const double* t0, int n, int S, double* rows) {
  if (!ctx) return KRRT_ERR_INVALID;
  REQUIRE(n >= 0 && S >= 1 && (n == 0 || (xi && xf && t0 && rows)), KRRT_ERR_INVALID, "null buffer / S < 1");
  if (n == 0) return KRRT_OK;
  CK(cudaSetDevice(ctx->cfg.device));
  cudaStream_t st = ctx->stream;
  int rc;
  const size_t sb = sizeof(double) * 4 * (size_t)n, ob = sizeof(double) * (size_t)n;
  const size_t rb = sizeof(double) * 5 * (size_t)n * S;
  if ((rc = ensure_scratch(ctx, 0, sb)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 1, sb)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 3, ob)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 7, rb)) != KRRT_OK) return rc;
  CK(cudaMemcpyAsync(ctx->d_scratch[0], xi, sb, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(ctx->d_scratch[1], xf, sb, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(ctx->d_scratch[3], t0, ob, cudaMemcpyHostToDevice, st));
  const long long total = (long long)n * S;
  if (ctx->cfg.precision == KRRT_F32)
    CK((cudaError_t)krrt_f32_api::edge_samples((const double*)ctx->d_scratch[0], (const double*)ctx->d_scratch[1],
                                               (const double*)ctx->d_scratch[3], n, S, ctx->cfg.r, (double*)ctx->d_scratch[7], st));
  else
    k_edge_samples<<<(unsigned)((total + 127) / 128), 128, 0, st>>>((const double*)ctx->d_scratch[0],
                                                                    (const double*)ctx->d_scratch[1],
                                                                    (const double*)ctx->d_scratch[3], n, S, ctx->cfg.r,
                                                                    (double*)ctx->d_scratch[7]);
  ctx->launches++;
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(rows, ctx->d_scratch[7], rb, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return KRRT_OK;
}

int krrt_tree_edges(krrt_ctx* ctx, int S, double* rows, int cap_nodes) {
  if (!ctx || !rows) return KRRT_ERR_INVALID;
  if (ctx->cfg.precision == KRRT_F32) {
    // binary32 tree: the stored edge parent -> node is the optimal edge between the two states departing at the parent's
    // arrival time, i.e. what the float Connector kernel gives for (states[parent], states[node], time[parent])
    REQUIRE(S >= 1 && cap_nodes >= 0, KRRT_ERR_INVALID, "S < 1 / cap_nodes < 0");
    int n32 = 0, gi32 = -1;
    int rc32 = krrt_f32_api::tree_info(ctx->f32, &n32, nullptr, nullptr, nullptr, &gi32);
    if (rc32 != KRRT_OK) {
      ctx->err = krrt_f32_api::tree_error(ctx->f32);
      return rc32;
    }
    const int n = std::min(n32 + (gi32 == n32 ? 1 : 0), cap_nodes);
    if (n == 0) return KRRT_OK;
    std::vector<double> st(4 * (size_t)n), tm((size_t)n), xi(4 * (size_t)n), t0((size_t)n);
    std::vector<int32_t> par((size_t)n);
    if ((rc32 = krrt_tree_read(ctx, 0, n, st.data(), par.data(), nullptr, tm.data())) != KRRT_OK) return rc32;
    for (int i = 1; i < n; ++i) {
      for (int d = 0; d < 4; ++d) xi[4 * (size_t)i + d] = st[4 * (size_t)par[i] + d];
      t0[i] = tm[par[i]];
    }
    for (int k = 0; k < S; ++k) {
      double* row = rows + 5 * (size_t)k;
      row[0] = st[0]; row[1] = st[1]; row[2] = st[2]; row[3] = st[3]; row[4] = 0.0;
    }
    if (n == 1) return KRRT_OK;
    return krrt_edge_samples(ctx, xi.data() + 4, st.data() + 4, t0.data() + 1, n - 1, S, rows + 5 * (size_t)S);
  }
  REQUIRE(ctx->started, KRRT_ERR_STATE, "no tree: call krrt_set_start");
  REQUIRE(S >= 1 && cap_nodes >= 0, KRRT_ERR_INVALID, "S < 1 / cap_nodes < 0");
  CK(cudaSetDevice(ctx->cfg.device));
  int n = 0, ng = 0;
  int rc = current_size(ctx, &n, &ng);
  if (rc != KRRT_OK) return rc;
  n = std::min(ng, cap_nodes);  // (cap_nodes > tree size: the goal node's edge follows the tree's)
  if (n == 0) return KRRT_OK;
  cudaStream_t st = ctx->stream;
  const size_t rb = sizeof(double) * 5 * (size_t)n * S;
  if ((rc = ensure_scratch(ctx, 7, rb)) != KRRT_OK) return rc;
  const long long total = (long long)n * S;
  k_tree_edges<<<(unsigned)((total + 127) / 128), 128, 0, st>>>(ctx->T, n, S, (double*)ctx->d_scratch[7]);
  ctx->launches++;
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(rows, ctx->d_scratch[7], rb, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return KRRT_OK;
}

// ---- yaw profiles: Trajectory1D::AngleController for q queries at once
static int launch_yaw(krrt_ctx* ctx, const double* d_w, const double* d_dw, const double* d_yaw, int q, double v_max,
                      double a_max, int n, double* d_states, double* d_times, double* d_inputs, double* d_ft,
                      int32_t* d_status, int32_t* d_nseg, cudaStream_t st) {
  YawJob J;
  J.w = d_w; J.dw = d_dw; J.yaw = d_yaw;
  J.q = q; J.n = n; J.v_max = v_max; J.a_max = a_max;
  J.states = d_states; J.times = d_times; J.inputs = d_inputs; J.final_time = d_ft;
  J.status = d_status; J.n_seg = d_nseg;
  const int ctas = (int)std::min<long long>(((long long)q + YAW_WARPS - 1) / YAW_WARPS,
                                            (long long)ctx->prop.multiProcessorCount * 16);
  k_yaw_profiles<<<std::max(1, ctas), YAW_WARPS * 32, 0, st>>>(J);
  ctx->launches++;
  CK(cudaGetLastError());
  return KRRT_OK;
}

int krrt_yaw_profiles_device(krrt_ctx* ctx, const double* d_w, const double* d_dw, const double* d_yaw, int q,
                             double v_max, double a_max, int n, double* d_states, double* d_times, double* d_inputs,
                             double* d_final_time, int32_t* d_status, int32_t* d_n_seg, void* cuda_stream) {
  if (!ctx) return KRRT_ERR_INVALID;
  REQUIRE(q >= 0 && n >= 0 && v_max > 0.0 && a_max > 0.0, KRRT_ERR_INVALID, "q, n >= 0 and v_max, a_max > 0");
  if (q == 0) return KRRT_OK;
  REQUIRE(d_w && d_dw && d_yaw && d_states && d_times && d_inputs && d_final_time && d_status && d_n_seg,
          KRRT_ERR_INVALID, "null buffer");
  CK(cudaSetDevice(ctx->cfg.device));
  return launch_yaw(ctx, d_w, d_dw, d_yaw, q, v_max, a_max, n, d_states, d_times, d_inputs, d_final_time, d_status,
                    d_n_seg, cuda_stream ? (cudaStream_t)cuda_stream : ctx->stream);
}

int krrt_yaw_profiles(krrt_ctx* ctx, const double* w, const double* dw, const double* yaw, int q, double v_max,
                      double a_max, int n, double* states, double* times, double* inputs, double* final_time,
                      int32_t* status, int32_t* n_seg) {
  if (!ctx) return KRRT_ERR_INVALID;
  REQUIRE(q >= 0 && n >= 0 && v_max > 0.0 && a_max > 0.0, KRRT_ERR_INVALID, "q, n >= 0 and v_max, a_max > 0");
  if (q == 0) return KRRT_OK;
  REQUIRE(w && dw && yaw && states && times && inputs && final_time && status && n_seg, KRRT_ERR_INVALID, "null buffer");
  CK(cudaSetDevice(ctx->cfg.device));
  cudaStream_t st = ctx->stream;
  const size_t rows = (size_t)q * (size_t)(n > 0 ? n : 1);
  const size_t qb = sizeof(double) * (size_t)q;
  int rc;
  if ((rc = ensure_scratch(ctx, 0, 3 * qb)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 1, sizeof(double) * 4 * rows)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 2, qb + 2 * sizeof(int32_t) * (size_t)q)) != KRRT_OK) return rc;
  double* d_in = (double*)ctx->d_scratch[0];
  double* d_rows = (double*)ctx->d_scratch[1];
  double* d_ft = (double*)ctx->d_scratch[2];
  int32_t* d_st = (int32_t*)(d_ft + q);
  CK(cudaMemcpyAsync(d_in, w, qb, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(d_in + q, dw, qb, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(d_in + 2 * (size_t)q, yaw, qb, cudaMemcpyHostToDevice, st));
  rc = launch_yaw(ctx, d_in, d_in + q, d_in + 2 * (size_t)q, q, v_max, a_max, n, d_rows, d_rows + 2 * rows,
                  d_rows + 3 * rows, d_ft, d_st, d_st + q, st);
  if (rc != KRRT_OK) return rc;
  CK(cudaMemcpyAsync(states, d_rows, sizeof(double) * 2 * rows, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(times, d_rows + 2 * rows, sizeof(double) * rows, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(inputs, d_rows + 3 * rows, sizeof(double) * rows, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(final_time, d_ft, qb, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(status, d_st, sizeof(int32_t) * (size_t)q, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(n_seg, d_st + q, sizeof(int32_t) * (size_t)q, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return KRRT_OK;
}

// ---- obstacle ingest: RobotSubscriber callbacks for f frames at once
static int launch_ingest(krrt_ctx* ctx, const IngestJob& J, cudaStream_t st) {
  const long long total = (long long)J.f * (J.n_team + J.n_rival + 1);
  const long long ctas = std::min<long long>((total + 255) / 256, (long long)ctx->prop.multiProcessorCount * 8);
  k_ingest_vision<<<(unsigned)std::max<long long>(1, ctas), 256, 0, st>>>(J);
  ctx->launches++;
  CK(cudaGetLastError());
  return KRRT_OK;
}

int krrt_ingest_vision_device(krrt_ctx* ctx, const float* d_team, const int32_t* d_agent_id, int n_team,
                              const float* d_rival, int n_rival, const int32_t* d_self_id, int f,
                              double* d_self_state, double* d_heading, int32_t* d_self_index, double* d_obstacles,
                              int32_t* d_n_obs, void* cuda_stream) {
  if (!ctx) return KRRT_ERR_INVALID;
  REQUIRE(f >= 0 && n_team >= 0 && n_rival >= 0, KRRT_ERR_INVALID, "f, n_team, n_rival >= 0");
  if (f == 0) return KRRT_OK;
  REQUIRE((d_team && d_agent_id) || n_team == 0, KRRT_ERR_INVALID, "null team buffer");
  REQUIRE(d_rival || n_rival == 0, KRRT_ERR_INVALID, "null rival buffer");
  REQUIRE(d_self_id && d_self_state && d_heading && d_self_index && d_n_obs && (d_obstacles || n_team + n_rival == 0),
          KRRT_ERR_INVALID, "null buffer");
  CK(cudaSetDevice(ctx->cfg.device));
  IngestJob J = {d_team, d_agent_id, d_rival, d_self_id, d_self_state, d_heading, d_self_index, d_obstacles, d_n_obs,
                 f, n_team, n_rival};
  return launch_ingest(ctx, J, cuda_stream ? (cudaStream_t)cuda_stream : ctx->stream);
}

int krrt_ingest_vision(krrt_ctx* ctx, const float* team, const int32_t* agent_id, int n_team, const float* rival,
                       int n_rival, const int32_t* self_id, int f, double* self_state, double* heading,
                       int32_t* self_index, double* obstacles, int32_t* n_obs) {
  if (!ctx) return KRRT_ERR_INVALID;
  REQUIRE(f >= 0 && n_team >= 0 && n_rival >= 0, KRRT_ERR_INVALID, "f, n_team, n_rival >= 0");
  if (f == 0) return KRRT_OK;
  REQUIRE((team && agent_id) || n_team == 0, KRRT_ERR_INVALID, "null team buffer");
  REQUIRE(rival || n_rival == 0, KRRT_ERR_INVALID, "null rival buffer");
  REQUIRE(self_id && self_state && heading && self_index && n_obs && (obstacles || n_team + n_rival == 0),
          KRRT_ERR_INVALID, "null buffer");
  CK(cudaSetDevice(ctx->cfg.device));
  cudaStream_t st = ctx->stream;
  const size_t F = (size_t)f, slots = (size_t)n_team + n_rival;
  const size_t b_team = sizeof(float) * 5 * F * n_team, b_ids = sizeof(int32_t) * F * n_team;
  const size_t b_rival = sizeof(float) * 4 * F * n_rival, b_self = sizeof(int32_t) * F;
  const size_t b_obs = sizeof(double) * 4 * F * slots;
  int rc;
  if ((rc = ensure_scratch(ctx, 0, b_team + 16)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 1, b_rival + 16)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 2, b_ids + b_self + 16)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 3, b_obs + 16)) != KRRT_OK) return rc;
  if ((rc = ensure_scratch(ctx, 4, sizeof(double) * 5 * F + 2 * sizeof(int32_t) * F)) != KRRT_OK) return rc;
  float* d_team = (float*)ctx->d_scratch[0];
  float* d_rival = (float*)ctx->d_scratch[1];
  int32_t* d_ids = (int32_t*)ctx->d_scratch[2];
  int32_t* d_self = d_ids + F * n_team;
  double* d_obs = (double*)ctx->d_scratch[3];
  double* d_state = (double*)ctx->d_scratch[4];
  double* d_head = d_state + 4 * F;
  int32_t* d_sidx = (int32_t*)(d_head + F);
  int32_t* d_nobs = d_sidx + F;
  if (b_team) CK(cudaMemcpyAsync(d_team, team, b_team, cudaMemcpyHostToDevice, st));
  if (b_ids) CK(cudaMemcpyAsync(d_ids, agent_id, b_ids, cudaMemcpyHostToDevice, st));
  if (b_rival) CK(cudaMemcpyAsync(d_rival, rival, b_rival, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(d_self, self_id, b_self, cudaMemcpyHostToDevice, st));
  IngestJob J = {d_team, d_ids, d_rival, d_self, d_state, d_head, d_sidx, d_obs, d_nobs, f, n_team, n_rival};
  if ((rc = launch_ingest(ctx, J, st)) != KRRT_OK) return rc;
  CK(cudaMemcpyAsync(self_state, d_state, sizeof(double) * 4 * F, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(heading, d_head, sizeof(double) * F, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(self_index, d_sidx, sizeof(int32_t) * F, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(n_obs, d_nobs, sizeof(int32_t) * F, cudaMemcpyDeviceToHost, st));
  if (b_obs) CK(cudaMemcpyAsync(obstacles, d_obs, b_obs, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return KRRT_OK;
}

}  // extern "C"
