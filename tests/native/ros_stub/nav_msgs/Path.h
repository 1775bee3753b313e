// TEST INFRASTRUCTURE (see ros/ros.h in this directory): nav_msgs/Path
#ifndef KRRT_ROS_STUB_PATH_H
#define KRRT_ROS_STUB_PATH_H
#include <geometry_msgs/PoseStamped.h>
namespace nav_msgs {
struct Path {
  std_msgs::Header header;
  std::vector<geometry_msgs::PoseStamped> poses;
};
}
#endif
