/* ingest_oracle.h — CPU restatement of the reference's obstacle ingest (TEST INFRASTRUCTURE: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu legs may load it; the product never does).
 * Follows RobotSubscriber::callback (robotsubscriber.cpp:44-77), ::rival_callback (robotsubscriber.cpp:29-42) and
 * ::getObstacles (robotsubscriber.h:50-55) of /root/reference/src/robosoccer_motion_planning; field types from
 * gazebo_visual/src/nubot_common/msgs/RobotInfo.msg:2,9-12 (int32 AgentID, float32 Point2d pos / vtrans, float32 heading).
 * PARITY PINNED: the reference source for this piece compiles here — robotsubscriber.cpp where it lies, against the ROS
 * stand-ins of tests/native/ros_stub (oracle/Makefile target `ref` -> oracle/_ref/libingest_ref.so, wrapper
 * oracle/ref_wrap/ingest_ref_wrap.cpp) — and this restatement equals it bit for bit on frames with missing and repeated
 * own entries, no friends, no rivals (tests/test_ingest.py: test_restatement_equals_reference_subscriber); the CUDA kernel
 * is held against the same library on the GPU box (test_gpu_ingest_equals_reference_subscriber). */
#ifndef INGEST_ORACLE_H
#define INGEST_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
/* team: f x n_team x 5 float32 (pos.x, pos.y, vtrans.x, vtrans.y, heading.theta) [cm, cm/s, rad]; agent_id: f x n_team
 * rival: f x n_rival x 4 float32 (pos.x, pos.y, vtrans.x, vtrans.y); self_id: f
 * self_state: f x 4 [m, m/s] (zeros when no entry carries self_id), heading: f, self_index: f (-1 = absent, else the
 * LAST matching entry, as the callback's loop overwrites), obstacles: f x (n_team + n_rival) x 4 (friends in message
 * order without self, then rivals; unused rows zero), n_obs: f */
void io_ingest_vision(const float* team, const int32_t* agent_id, int n_team, const float* rival, int n_rival,
                      const int32_t* self_id, int f, double* self_state, double* heading, int32_t* self_index,
                      double* obstacles, int32_t* n_obs);
#ifdef __cplusplus
}
#endif
#endif
