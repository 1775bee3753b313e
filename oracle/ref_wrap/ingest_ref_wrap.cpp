// ingest_ref_wrap.cpp — TEST INFRASTRUCTURE: a C entry point around the REFERENCE's own obstacle ingest, so that the
// restatement (oracle/ingest_oracle.c) and the CUDA kernel (k_ingest_vision) can be held against the real thing.
// Compiled together with /root/reference/src/robosoccer_motion_planning/src/robotsubscriber.cpp WHERE IT LIES (nothing is
// copied; oracle/Makefile target `ref` -> oracle/_ref/libingest_ref.so), with tests/native/ros_stub/ standing in for
// roscpp / tf / the message headers (the stand-ins carry no arithmetic: fields and callbacks only).
// Per frame: a fresh RobotSubscriber(node, self_id, NUBOT) is handed the rival message (rival_callback,
// robotsubscriber.cpp:29-42) and the team message (callback, :44-77); getState(), getHeading().first and getObstacles()
// (robotsubscriber.h:48-56) are what comes out.  Same argument layout as io_ingest_vision (oracle/ingest_oracle.h).
// Two conventions are this wrapper's, not the reference's: a frame without the robot's own entry reports a zero state /
// heading and self_index -1 (the reference leaves `state` as it was — uninitialised in a fresh object), and self_index is
// found here (the reference does not keep it).
#include <cstdint>
#include <cstring>
#include <memory>

#include "robotsubscriber.h"

extern "C" int ingest_ref_vision(const float* team, const int32_t* agent_id, int n_team, const float* rival, int n_rival,
                                 const int32_t* self_id, int f, double* self_state, double* heading, int32_t* self_index,
                                 double* obstacles, int32_t* n_obs) {
  ros::NodeHandle node;
  const int rows = n_team + n_rival;
  for (int q = 0; q < f; ++q) {
    auto mt = std::make_shared<nubot_common::OminiVisionInfo>();
    auto mr = std::make_shared<nubot_common::OminiVisionInfo>();
    int self = -1;
    for (int i = 0; i < n_team; ++i) {
      const float* p = team + ((size_t)q * n_team + i) * 5;
      nubot_common::RobotInfo r;
      r.AgentID = agent_id[(size_t)q * n_team + i];
      r.pos.x = p[0]; r.pos.y = p[1]; r.vtrans.x = p[2]; r.vtrans.y = p[3]; r.heading.theta = p[4];
      mt->robotinfo.push_back(r);
      if (r.AgentID == self_id[q]) self = i;
    }
    for (int i = 0; i < n_rival; ++i) {
      const float* p = rival + ((size_t)q * n_rival + i) * 4;
      nubot_common::RobotInfo r;
      r.pos.x = p[0]; r.pos.y = p[1]; r.vtrans.x = p[2]; r.vtrans.y = p[3];
      mr->robotinfo.push_back(r);
    }
    RobotSubscriber sub(node, (size_t)self_id[q], RobotSubscriber::NUBOT);
    sub.rival_callback(mr);
    sub.callback(mt);
    double* st = self_state + (size_t)q * 4;
    self_index[q] = self;
    if (self >= 0) {
      const auto& s = sub.getState();
      for (int k = 0; k < 4; ++k) st[k] = s[k];
      heading[q] = sub.getHeading().first;
    } else {
      std::memset(st, 0, 4 * sizeof(double));
      heading[q] = 0.0;
    }
    const auto obs = sub.getObstacles();
    if ((int)obs.size() > rows) return 1;
    n_obs[q] = (int32_t)obs.size();
    double* o = obstacles + (size_t)q * rows * 4;
    std::memset(o, 0, (size_t)rows * 4 * sizeof(double));
    for (size_t j = 0; j < obs.size(); ++j)
      for (int k = 0; k < 4; ++k) o[j * 4 + k] = obs[j][k];
  }
  return 0;
}
