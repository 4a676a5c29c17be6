// TEST-ONLY stand-in for roscpp: just enough declarations for `g++ -fsyntax-only -DSEPB_WITH_ROS host/ros_node.cpp`
// in an image without ROS (tests/test_host_services.py::test_ros_node_compiles).  Never linked, never shipped.
#pragma once
#include <cstdio>
#include <functional>
#include <memory>
#include <string>
#include <vector>

namespace boost {
template <class T>
using function = std::function<T>;
template <class T>
using shared_ptr = std::shared_ptr<T>;
}  // namespace boost

#define ROS_ERROR(...) std::fprintf(stderr, __VA_ARGS__)
#define ROS_INFO(...) std::fprintf(stderr, __VA_ARGS__)

namespace ros {
void init(int& argc, char** argv, const std::string& name);
void spin();
struct Time {
  double sec = 0;
  bool operator>(const Time& o) const { return sec > o.sec; }
};
struct Duration {
  explicit Duration(double s) : sec(s) {}
  double sec;
  bool sleep() const;
};
struct ServiceServer {};
struct ServiceClient {
  template <class Srv>
  bool call(Srv& srv);
};
struct NodeHandle {
  template <class T>
  bool getParam(const std::string& key, T& value) const;
  template <class Req, class Res>
  ServiceServer advertiseService(const std::string& name, const boost::function<bool(Req&, Res&)>& cb);
  template <class Srv>
  ServiceClient serviceClient(const std::string& name);
};
namespace topic {
template <class M>
boost::shared_ptr<M const> waitForMessage(const std::string& topic, NodeHandle& nh);
}
}  // namespace ros
