// TEST INFRASTRUCTURE (see ros/ros.h in this directory): the slice of tf the reference's subscribers / publisher call
#ifndef KRRT_ROS_STUB_TF_DATATYPES_H
#define KRRT_ROS_STUB_TF_DATATYPES_H
#include <cmath>
#include <geometry_msgs/PoseStamped.h>
namespace tf {
struct Vector3 {
  double x = 0, y = 0, z = 0;
  Vector3() {}
  Vector3(double a, double b, double c) : x(a), y(b), z(c) {}
};
struct Quaternion {
  double x = 0, y = 0, z = 0, w = 1;
  void setRPY(double roll, double pitch, double yaw) {
    const double cr = std::cos(roll / 2), sr = std::sin(roll / 2), cp = std::cos(pitch / 2), sp = std::sin(pitch / 2),
                 cy = std::cos(yaw / 2), sy = std::sin(yaw / 2);
    x = sr * cp * cy - cr * sp * sy;
    y = cr * sp * cy + sr * cp * sy;
    z = cr * cp * sy - sr * sp * cy;
    w = cr * cp * cy + sr * sp * sy;
  }
};
inline void quaternionMsgToTF(const geometry_msgs::Quaternion& m, Quaternion& q) { q.x = m.x, q.y = m.y, q.z = m.z, q.w = m.w; }
inline void quaternionTFToMsg(const Quaternion& q, geometry_msgs::Quaternion& m) { m.x = q.x, m.y = q.y, m.z = q.z, m.w = q.w; }
inline double getYaw(const Quaternion& q) { return std::atan2(2 * (q.w * q.z + q.x * q.y), 1 - 2 * (q.y * q.y + q.z * q.z)); }
inline double getYaw(const geometry_msgs::Quaternion& m) { Quaternion q; quaternionMsgToTF(m, q); return getYaw(q); }
struct Transform {
  Vector3 origin;
  Quaternion rotation;
  void setOrigin(const Vector3& v) { origin = v; }
  void setRotation(const Quaternion& q) { rotation = q; }
};
struct StampedTransform : Transform {
  ros::Time stamp;
  std::string frame, child;
  StampedTransform(const Transform& t, const ros::Time& s, const std::string& f, const std::string& c)
      : Transform(t), stamp(s), frame(f), child(c) {}
};
}
#endif
