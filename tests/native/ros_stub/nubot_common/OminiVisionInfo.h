// TEST INFRASTRUCTURE (see ros/ros.h in this directory): nubot_common/OminiVisionInfo with the fields of
// /root/reference/src/gazebo_visual/src/nubot_common/msgs/{OminiVisionInfo,RobotInfo,Point2d,Angle}.msg the subscriber reads
#ifndef KRRT_ROS_STUB_OMINIVISIONINFO_H
#define KRRT_ROS_STUB_OMINIVISIONINFO_H
#include <std_msgs/Header.h>
namespace nubot_common {
struct Point2d { float x = 0, y = 0; };
struct Angle { float theta = 0; };
struct RobotInfo {
  std_msgs::Header header;
  int32_t AgentID = 0;
  Point2d pos;
  Angle heading;
  float vrot = 0;
  Point2d vtrans;
  bool isvalid = true;
};
struct OminiVisionInfo {
  std_msgs::Header header;
  std::vector<RobotInfo> robotinfo;
  typedef std::shared_ptr<const OminiVisionInfo> ConstPtr;
};
}
#endif
