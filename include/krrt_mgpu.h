/* krrt_mgpu.h — C ABI of libkrrt_mgpu.so: the batched planning sweep over the GPUs of one box.
 *
 * What it replaces: solve_rrt() (robosoccer_motion_planning/src/robosoccer_motion_planning_node.cpp:167-196: setStart,
 * setObstacles, for(i<iteration) rrt.grow(xg), goalIndex/path) run for Q independent queries (robots, seeds, start/goal
 * pairs) — BASELINE.json configs[4], SURVEY.md 8(b) `krrt_sweep_multi_gpu`, 8(e).
 *
 * Queries are independent, so the path shards with NO data-path collective: GPU g plans the contiguous block
 * [g*Q/G, (g+1)*Q/G) with libkrrt's batched planner (Philox is keyed by the GLOBAL query id: results do not depend on
 * G), and the only exchange is the gather of the fixed-size result records (krrt_result, 1304 B) over NCCL
 * (NVLink 5 / NVSwitch).  The gather is overlapped with the planning: the planner counts finished queries per chunk
 * (krrt_plan_batch_device_chunks), a second stream waits for each chunk's counter (cuStreamWaitValue32) and all-gathers
 * the chunk while later chunks are still being planned; with NCCL >= 2.28 the buffers are symmetric windows and the
 * communicator asks for the zero-CTA policy, so the transfers run on the copy engines and take no SM from the planner.
 *
 * This library links libkrrt.so, libcudart and libnccl; libkrrt.so itself keeps linking libcudart only.  Plain C,
 * opaque handle, caller-owned HOST buffers, int status codes of krrt.h, no exceptions; a handle is not thread-safe.
 * There is no CPU fallback.
 *
 * Two ways to stand it up:
 *   A. one process drives n_gpus devices (krrt_mgpu_create): every call fans out to all of them from the calling thread;
 *   B. one process per GPU, e.g. under `python -m torch.distributed.run` or mpirun (krrt_mgpu_create_rank): rank 0 makes
 *      a 128-byte id (krrt_mgpu_unique_id), hands it to the other ranks by whatever means the launcher offers (a file,
 *      MPI_Bcast, a TCP store), and every rank calls every entry point below collectively, in the same order.
 */
#ifndef KRRT_MGPU_H
#define KRRT_MGPU_H

#include "krrt.h"

#ifdef __cplusplus
extern "C" {
#endif

#define KRRT_MGPU_ID_BYTES 128

typedef struct krrt_mgpu krrt_mgpu;

/* A. n_gpus devices of this process; devices: NULL (0 .. n_gpus-1) or n_gpus CUDA ordinals.  cfg->device is ignored. */
int krrt_mgpu_create(const krrt_config* cfg, int n_gpus, const int* devices, krrt_mgpu** out);
/* B. this process is rank `rank` of `world` and drives cfg->device.                                                 */
int krrt_mgpu_unique_id(void* id /* KRRT_MGPU_ID_BYTES */);
int krrt_mgpu_create_rank(const krrt_config* cfg, int rank, int world, const void* id, krrt_mgpu** out);
void krrt_mgpu_destroy(krrt_mgpu* mg);
const char* krrt_mgpu_last_error(const krrt_mgpu* mg); /* mg may be NULL: last create error */
/* world = GPUs of the whole job, local = GPUs this handle drives, first = global rank of its first GPU, nccl = 1 when
 * the gather goes over NCCL (world > 1), nccl_version = ncclGetVersion (0 when world == 1)                        */
int krrt_mgpu_info(const krrt_mgpu* mg, int* world, int* local, int* first, int* nccl, int* nccl_version);

/* pure host logic (callable without a GPU): the block of rank r, and where chunk c of rank r lies in the gathered
 * array: the gather buffer is laid out [chunk][rank][records of that chunk], chunks of `chunk` queries of every rank's
 * PADDED block (pad = ceil(q / world)); *offset = first record, *count = real records of (r, c), *global = index of
 * the first one in the q-long result array                                                                          */
void krrt_mgpu_shard(int q, int rank, int world, int* lo, int* hi);
int krrt_mgpu_chunk_layout(int q, int world, int chunk, int rank, int c, int64_t* offset, int* count, int* global);

/* The sweep in one call: Q queries from host arrays (starts / goals q x 4, obstacles q x m x 4; every rank passes the
 * FULL arrays in mode B), results for all Q queries in global order in `out` (on every rank).  Philox query id =
 * query_base + i.  Uses cfg of the handle except capacity (= n_iter + 1 nodes per tree).                           */
int krrt_sweep_multi_gpu(krrt_mgpu* mg, const double* starts, const double* goals, const double* obstacles, int q,
                         int m, int dynamic, int n_iter, uint32_t query_base, krrt_result* out);
/* The same in three steps, for callers that keep the scenes resident in HBM between sweeps (and for timing):
 * load copies every GPU's block of the inputs to it; run plans all blocks and gathers the records (device time in ms
 * between the first kernel and the last gather, max over the GPUs of this handle; max over all ranks with
 * krrt_mgpu_max); results copies the gathered records to the host in global order.                                 */
int krrt_sweep_load(krrt_mgpu* mg, const double* starts, const double* goals, const double* obstacles, int q, int m);
int krrt_sweep_run(krrt_mgpu* mg, int dynamic, int n_iter, uint32_t query_base, double* device_ms);
int krrt_sweep_results(krrt_mgpu* mg, krrt_result* out);
/* device times of the last calls on this handle's first GPU, CUDA events on its planning stream: run_ms = the last
 * krrt_sweep_run; e2e_ms = from the first host-to-device copy of the last krrt_sweep_load to the last device-to-host
 * copy of the last krrt_sweep_results (what krrt_sweep_multi_gpu spans, host gaps included)                        */
int krrt_sweep_last_timing(krrt_mgpu* mg, double* run_ms, double* e2e_ms);
/* oracle-comparable counters of the last run, summed over ALL queries of the sweep (all ranks)                      */
int krrt_sweep_stats(krrt_mgpu* mg, krrt_stats* out);
/* kernels launched by this handle's contexts since creation                                                         */
int64_t krrt_mgpu_launch_count(const krrt_mgpu* mg);
/* 1 when the last run shipped its chunks while later ones were still being planned (stream wait-value available,
 * more than one chunk), *chunks = chunks per block                                                                  */
int krrt_sweep_overlap(const krrt_mgpu* mg, int* overlapped, int* chunks);

/* collective helpers for timing loops: barrier (device-side, all ranks), max / sum of one double over all ranks     */
int krrt_mgpu_barrier(krrt_mgpu* mg);
int krrt_mgpu_max(krrt_mgpu* mg, double* x);

#ifdef __cplusplus
}
#endif
#endif
