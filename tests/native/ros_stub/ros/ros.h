// TEST INFRASTRUCTURE — a minimal in-process stand-in for roscpp, just enough for the reference's OWN sources
// (/root/reference/src/robosoccer_motion_planning/src/*.cpp, include/*.h*) to compile and run UNMODIFIED against
// include/*.hpp of this repo (tests/test_reference_node.py).  Not part of the product; no ROS behaviour beyond what those
// files call: init / ok / NodeHandle::subscribe+advertise / Subscriber / Publisher / Time / Duration / Rate /
// AsyncSpinner / param::get+has / ROS_INFO+ROS_WARN.  Topics are an in-process bus (ros_stub_runtime.cpp feeds it).
#ifndef KRRT_ROS_STUB_ROS_H
#define KRRT_ROS_STUB_ROS_H

#include <array>
#include <cstdint>
#include <cstdio>
#include <functional>
#include <memory>
#include <stdexcept>
#include <string>
#include <typeindex>
#include <vector>

namespace ros {

struct Duration {
  double s = 0;
  Duration() {}
  explicit Duration(double x) : s(x) {}
  double toSec() const { return s; }
};
struct Time {
  double s = 0;
  static Time now();
  double toSec() const { return s; }
  Time operator+(const Duration& d) const { Time t; t.s = s + d.s; return t; }
  Duration operator-(const Time& o) const { return Duration(s - o.s); }
  bool operator<(const Time& o) const { return s < o.s; }
  bool operator>(const Time& o) const { return s > o.s; }
};

void init(int& argc, char** argv, const std::string& name);
bool ok();

struct Rate {
  explicit Rate(double hz) : period(1.0 / hz) {}
  void sleep();
  double period;
};
struct AsyncSpinner {
  explicit AsyncSpinner(int) {}
  void start();  // the stub delivers the scenario's messages here, synchronously
};

namespace stub {
using Handler = std::function<void(const std::shared_ptr<const void>&)>;
void add_handler(const std::string& topic, std::type_index type, Handler h);
void deliver(const std::string& topic, std::type_index type, const std::shared_ptr<const void>& msg);
void record(const std::string& topic, std::type_index type, const void* msg);  // Publisher::publish lands here
int publishers(const std::string& topic);
bool verbose();
template <typename... A>
inline void log(const char* level, const char* fmt, A... a) {
  if (level[0] == 'I' && !verbose()) return;  // ROS_INFO is per-iteration chatter; warnings always print
  std::fprintf(stderr, "[%s] ", level);
  if constexpr (sizeof...(A) == 0) std::fputs(fmt, stderr);
  else std::fprintf(stderr, fmt, a...);
  std::fputc('\n', stderr);
}
bool param_raw(const std::string& name, std::string& out);  // environment variable ROSPARAM_<name>
}  // namespace stub

class Subscriber {
 public:
  Subscriber() {}
  explicit Subscriber(std::string t) : topic_(std::move(t)) {}
  uint32_t getNumPublishers() const { return (uint32_t)stub::publishers(topic_); }
 private:
  std::string topic_;
};
class Publisher {
 public:
  Publisher() {}
  Publisher(std::string t, std::type_index ty) : topic_(std::move(t)), type_(ty) {}
  template <typename M>
  void publish(const M& m) const { stub::record(topic_, std::type_index(typeid(M)), &m); }
 private:
  std::string topic_;
  std::type_index type_ = std::type_index(typeid(void));
};
class NodeHandle {
 public:
  template <typename M, typename T>
  Subscriber subscribe(const std::string& topic, uint32_t, void (T::*fp)(const std::shared_ptr<const M>&), T* obj) {
    stub::add_handler(topic, std::type_index(typeid(M)), [fp, obj](const std::shared_ptr<const void>& p) {
      (obj->*fp)(std::static_pointer_cast<const M>(p));
    });
    return Subscriber(topic);
  }
  template <typename M>
  Publisher advertise(const std::string& topic, uint32_t) { return Publisher(topic, std::type_index(typeid(M))); }
};

namespace param {
inline bool has(const std::string& name) { std::string v; return stub::param_raw(name, v); }
inline bool get(const std::string& name, std::string& out) { return stub::param_raw(name, out); }
inline bool get(const std::string& name, int& out) {
  std::string v; if (!stub::param_raw(name, v)) return false; out = std::stoi(v); return true;
}
inline bool get(const std::string& name, double& out) {
  std::string v; if (!stub::param_raw(name, v)) return false; out = std::stod(v); return true;
}
inline bool get(const std::string& name, bool& out) {
  std::string v; if (!stub::param_raw(name, v)) return false; out = (v == "1" || v == "true"); return true;
}
}  // namespace param
}  // namespace ros

#pragma GCC diagnostic ignored "-Wformat-security"
#define ROS_INFO(...) ::ros::stub::log("INFO", __VA_ARGS__)
#define ROS_WARN(...) ::ros::stub::log("WARN", __VA_ARGS__)
#define ROS_ERROR(...) ::ros::stub::log("ERROR", __VA_ARGS__)

#endif
