/* ingest_oracle.c — see ingest_oracle.h (TEST INFRASTRUCTURE ONLY; parity pinned by oracle/_ref/libingest_ref.so; CPU restatement of
 * robotsubscriber.cpp:29-77). */
#include "ingest_oracle.h"

#include <string.h>

void io_ingest_vision(const float* team, const int32_t* agent_id, int n_team, const float* rival, int n_rival,
                      const int32_t* self_id, int f, double* self_state, double* heading, int32_t* self_index,
                      double* obstacles, int32_t* n_obs) {
  const int slots = n_team + n_rival;
  for (int q = 0; q < f; ++q) {
    double* ob = obstacles + (size_t)q * slots * 4;
    double* st = self_state + (size_t)q * 4;
    memset(ob, 0, sizeof(double) * 4 * (size_t)slots);
    st[0] = st[1] = st[2] = st[3] = 0.0;
    heading[q] = 0.0;
    self_index[q] = -1;
    int k = 0;
    /* callback(): own team, cm -> m by DIVISION by 100. (robotsubscriber.cpp:51-54); the entry with AgentID == id is
       the robot's own state (:55-69), every other one an obstacle (:70-77) */
    for (int i = 0; i < n_team; ++i) {
      const float* r = team + ((size_t)q * n_team + i) * 5;
      const double px = (double)r[0] / 100., py = (double)r[1] / 100.;
      const double vx = (double)r[2] / 100., vy = (double)r[3] / 100.;
      if (agent_id[(size_t)q * n_team + i] == self_id[q]) {
        st[0] = px; st[1] = py; st[2] = vx; st[3] = vy;
        heading[q] = (double)r[4];
        self_index[q] = i;
      } else {
        ob[4 * k + 0] = px; ob[4 * k + 1] = py; ob[4 * k + 2] = vx; ob[4 * k + 3] = vy;
        ++k;
      }
    }
    /* rival_callback(): flipped ("nubot's rule") and scaled by MULTIPLICATION with 1e-2 (robotsubscriber.cpp:36-39);
       getObstacles() appends them after the friends (robotsubscriber.h:50-55) */
    for (int i = 0; i < n_rival; ++i) {
      const float* r = rival + ((size_t)q * n_rival + i) * 4;
      ob[4 * k + 0] = -(double)r[0] * 1e-2;
      ob[4 * k + 1] = -(double)r[1] * 1e-2;
      ob[4 * k + 2] = -(double)r[2] * 1e-2;
      ob[4 * k + 3] = -(double)r[3] * 1e-2;
      ++k;
    }
    n_obs[q] = k;
  }
}
