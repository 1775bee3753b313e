// TEST INFRASTRUCTURE — runtime of the roscpp stand-in (ros/ros.h in this directory) plus the SCENARIO that plays the
// simulator's part: when the node starts its spinner, one OminiVisionInfo message per team is delivered to the
// subscribers (cm and cm/s, rivals mirrored, as nubot_gazebo publishes them: SURVEY.md Appendix A), describing the C1/C2
// scene of robosoccer_motion_planning_b200/scenes.py: own robot 1 at (-8.5, 0), four team mates and five rivals at the other
// spawn poses (robot_up.sh:11-14).  ros::ok() lets the main loop run KRRT_STUB_CYCLES times (default 1).  Every
// nav_msgs::Path the node publishes is counted; the last one per topic is written to $KRRT_STUB_PATH_CSV.<topic>.csv
// with 17 significant digits so that a test can compare it bit for bit with the CPU oracle.
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <map>
#include <thread>

#include <geometry_msgs/PoseStamped.h>
#include <nav_msgs/Path.h>
#include <nubot_common/OminiVisionInfo.h>
#include <ros/ros.h>

namespace {
struct Bus {
  std::multimap<std::string, std::pair<std::type_index, ros::stub::Handler>> handlers;
  std::map<std::string, int> publishers;
  std::map<std::string, unsigned long> published;
  bool spinning = false;
  int ok_calls = 0, cycles = 1;
  std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
};
Bus& bus() {
  static Bus b;
  return b;
}
const char* env(const char* k) { return std::getenv(k); }

// scene: 9 other robots (x, y, vx, vy) in metres; velocities only when ROSPARAM_environment=dynamic (C2-like)
const double SPAWN[9][2] = {{-2, 1}, {-2, -1}, {-5, 2}, {-5, -2}, {8.5, 0}, {2, 1}, {2, -1}, {5, 2}, {5, -2}};
const double VEL[9][2] = {{0.5, -0.25}, {-0.5, 0.75}, {0.25, 0.5}, {-0.75, -0.5}, {-1, 0.25}, {0.5, 0.5}, {-0.25, -0.75}, {0.75, 0.25}, {-0.5, -0.25}};

void feed_scene() {
  const bool dynamic = env("ROSPARAM_environment") && !std::strcmp(env("ROSPARAM_environment"), "dynamic");
  auto own = std::make_shared<nubot_common::OminiVisionInfo>();
  auto rival = std::make_shared<nubot_common::OminiVisionInfo>();
  nubot_common::RobotInfo me;
  me.AgentID = 1;
  me.pos.x = -850.f, me.pos.y = 0.f;
  own->robotinfo.push_back(me);
  for (int i = 0; i < 9; ++i) {
    nubot_common::RobotInfo r;
    const double sgn = i < 4 ? 1.0 : -1.0;  // rivals report mirrored coordinates (robotsubscriber.cpp:35-39)
    r.AgentID = i < 4 ? i + 2 : i - 3;
    r.pos.x = (float)(sgn * SPAWN[i][0] * 100.0), r.pos.y = (float)(sgn * SPAWN[i][1] * 100.0);
    if (dynamic) r.vtrans.x = (float)(sgn * VEL[i][0] * 100.0), r.vtrans.y = (float)(sgn * VEL[i][1] * 100.0);
    (i < 4 ? own : rival)->robotinfo.push_back(r);
  }
  own->header.stamp = rival->header.stamp = ros::Time::now();
  ros::stub::deliver("/nubot1/omnivision/OmniVisionInfo", std::type_index(typeid(nubot_common::OminiVisionInfo)), own);
  ros::stub::deliver("/rival1/omnivision/OmniVisionInfo", std::type_index(typeid(nubot_common::OminiVisionInfo)), rival);
}

// the operator's "2D Nav Goal" (RViz publishes it on move_base_simple/goal): KRRT_STUB_GOAL="x,y", delivered when the
// main loop starts, i.e. after the node has seeded the subscriber with randomGoal() (…node.cpp:221-222)
void feed_goal() {
  const char* g = env("KRRT_STUB_GOAL");
  if (!g) return;
  auto msg = std::make_shared<geometry_msgs::PoseStamped>();
  if (std::sscanf(g, "%lf,%lf", &msg->pose.position.x, &msg->pose.position.y) != 2) return;
  msg->header.stamp = ros::Time::now();
  ros::stub::deliver("move_base_simple/goal", std::type_index(typeid(geometry_msgs::PoseStamped)), msg);
}

void dump_path(const std::string& topic, const nav_msgs::Path& p) {
  const char* base = env("KRRT_STUB_PATH_CSV");
  if (!base) return;
  std::string name = topic;
  for (auto& c : name)
    if (c == '/') c = '_';
  std::FILE* f = std::fopen((std::string(base) + "." + name + ".csv").c_str(), "w");
  if (!f) return;
  std::fputs("seq,stamp,x,y,qz,qw\n", f);
  for (const auto& q : p.poses)
    std::fprintf(f, "%u,%.17g,%.17g,%.17g,%.17g,%.17g\n", q.header.seq, q.header.stamp.toSec(), q.pose.position.x,
                 q.pose.position.y, q.pose.orientation.z, q.pose.orientation.w);
  std::fclose(f);
}
}  // namespace

namespace ros {
Time Time::now() {
  Time t;
  t.s = 1000.0 + std::chrono::duration<double>(std::chrono::steady_clock::now() - bus().t0).count();
  return t;
}
void init(int&, char**, const std::string&) {
  if (env("KRRT_STUB_CYCLES")) bus().cycles = std::atoi(env("KRRT_STUB_CYCLES"));
}
// called once per pass of "wait for topic" (…node.cpp:204) and once per planning cycle (:251)
bool ok() {
  Bus& b = bus();
  if (!b.spinning) return true;
  if (b.ok_calls == 1) feed_goal();
  if (b.ok_calls++ < b.cycles + 1) return true;  // +1: the wait loop's own call after the spinner started
  std::printf("ros_stub: %d cycle(s) done;", b.cycles);
  for (const auto& kv : b.published) std::printf(" %s x%lu", kv.first.c_str(), kv.second);
  std::printf("\n");
  return false;
}
void Rate::sleep() { std::this_thread::sleep_for(std::chrono::duration<double>(period < 0.01 ? period : 0.01)); }
void AsyncSpinner::start() {
  bus().spinning = true;
  bus().publishers["/nubot1/omnivision/OmniVisionInfo"] = 1;
  bus().publishers["/rival1/omnivision/OmniVisionInfo"] = 1;
  feed_scene();
}

namespace stub {
void add_handler(const std::string& topic, std::type_index type, Handler h) {
  bus().handlers.insert({topic, {type, std::move(h)}});
}
void deliver(const std::string& topic, std::type_index type, const std::shared_ptr<const void>& msg) {
  auto range = bus().handlers.equal_range(topic);
  for (auto it = range.first; it != range.second; ++it)
    if (it->second.first == type) it->second.second(msg);
}
void record(const std::string& topic, std::type_index type, const void* msg) {
  ++bus().published[topic];
  if (type == std::type_index(typeid(nav_msgs::Path))) dump_path(topic, *static_cast<const nav_msgs::Path*>(msg));
}
int publishers(const std::string& topic) {
  auto it = bus().publishers.find(topic);
  return it == bus().publishers.end() ? 0 : it->second;
}
bool verbose() {
  static const bool v = env("KRRT_STUB_VERBOSE") != nullptr;
  return v;
}
bool param_raw(const std::string& name, std::string& out) {
  const char* v = env(("ROSPARAM_" + name).c_str());
  if (!v) return false;
  out = v;
  return true;
}
}  // namespace stub
}  // namespace ros
