// TEST INFRASTRUCTURE (see ros/ros.h in this directory): tf2/LinearMath/Quaternion.h — the reference includes it and uses tf:: only
#ifndef KRRT_ROS_STUB_TF2_QUATERNION_H
#define KRRT_ROS_STUB_TF2_QUATERNION_H
#endif
