// TEST INFRASTRUCTURE (see ros/ros.h in this directory): std_msgs/Header
#ifndef KRRT_ROS_STUB_HEADER_H
#define KRRT_ROS_STUB_HEADER_H
#include <ros/ros.h>
namespace std_msgs {
struct Header {
  uint32_t seq = 0;
  ros::Time stamp;
  std::string frame_id;
};
}
#endif
