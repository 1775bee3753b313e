// TEST INFRASTRUCTURE (see ros/ros.h in this directory): tf::TransformBroadcaster (counts what is sent)
#ifndef KRRT_ROS_STUB_TF_BROADCASTER_H
#define KRRT_ROS_STUB_TF_BROADCASTER_H
#include <tf/transform_datatypes.h>
namespace tf {
struct TransformBroadcaster {
  void sendTransform(const StampedTransform&) { ++sent; }
  unsigned long sent = 0;
};
}
#endif
