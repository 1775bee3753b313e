// TEST INFRASTRUCTURE (see ros/ros.h in this directory): geometry_msgs/{Point,Quaternion,Pose,PoseStamped}
#ifndef KRRT_ROS_STUB_POSESTAMPED_H
#define KRRT_ROS_STUB_POSESTAMPED_H
#include <std_msgs/Header.h>
namespace geometry_msgs {
struct Point { double x = 0, y = 0, z = 0; };
struct Quaternion { double x = 0, y = 0, z = 0, w = 1; };
struct Pose { Point position; Quaternion orientation; };
struct PoseStamped {
  std_msgs::Header header;
  Pose pose;
  typedef std::shared_ptr<const PoseStamped> ConstPtr;
};
typedef std::shared_ptr<const PoseStamped> PoseStampedConstPtr;
}
#endif
