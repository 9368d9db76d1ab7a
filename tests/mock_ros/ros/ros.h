// MINIMAL STAND-IN for <ros/ros.h> -- test infrastructure only (tests/test_ros_shim.py): just enough of the roscpp surface the reference's
// WholeBodyController touches (wbc.cpp:22-59, 122-148, 579-683) to compile include/anymal_wbc_b200/whole_body_controller_ros.hpp and to drive its
// control loop in-process.  Published messages are kept per topic; subscriptions are kept as callbacks that a test delivers messages to.
#pragma once
#include <any>
#include <functional>
#include <iostream>
#include <map>
#include <sstream>
#include <string>
#include <vector>
namespace mock {
struct Bus {
    std::map<std::string, std::any> last; std::map<std::string, int> count; std::map<std::string, std::any> subs; std::map<std::string, std::any> params;
    int okBudget = 0;                                            // ros::ok() returns true this many times
    static Bus& get() { static Bus b; return b; }
};
template <class M> void deliver(const std::string& topic, const M& m) { std::any_cast<std::function<void(M)>&>(Bus::get().subs.at(topic))(m); }
template <class M> const M& lastPublished(const std::string& topic) { return std::any_cast<const M&>(Bus::get().last.at(topic)); }
template <class T> void setParam(const std::string& k, const T& v) { Bus::get().params[k] = v; }
}
namespace ros {
struct Duration { double s; explicit Duration(double v = 0) : s(v) {} };
struct Time {
    double t; explicit Time(double v = 0) : t(v) {}
    static Time now() { static double c = 0; c += 1e-3; return Time(c); }
    bool operator==(const Time& o) const { return t == o.t; } bool operator<=(const Time& o) const { return t <= o.t; }
    Time operator+(const Duration& d) const { return Time(t + d.s); }
};
struct Rate { explicit Rate(double) {} void sleep() {} };
struct Publisher {
    std::string topic;
    template <class M> void publish(const M& m) const { mock::Bus::get().last[topic] = m; mock::Bus::get().count[topic]++; }
};
struct Subscriber {};
struct NodeHandle {
    template <class M> Publisher advertise(const std::string& topic, int) { Publisher p; p.topic = topic; return p; }
    template <class M, class C> Subscriber subscribe(const std::string& topic, int, void (C::*fn)(M), C* obj) {
        mock::Bus::get().subs[topic] = std::function<void(M)>([obj, fn](M m) { (obj->*fn)(m); }); return Subscriber(); }
    template <class T> bool getParam(const std::string& k, T& v) { auto it = mock::Bus::get().params.find(k); if (it == mock::Bus::get().params.end()) return false; v = std::any_cast<T>(it->second); return true; }
};
inline bool ok() { return mock::Bus::get().okBudget-- > 0; }
inline void spin() {}
inline void init(int, char**, const std::string&) {}
}
#define ROS_INFO(...) do { } while (0)
#define ROS_WARN(...) do { std::cerr << "[WARN] " << #__VA_ARGS__ << std::endl; } while (0)
#define ROS_ERROR(...) do { std::cerr << "[ERROR] " << #__VA_ARGS__ << std::endl; } while (0)
#define ROS_INFO_STREAM(x) do { std::ostringstream s_; s_ << x; } while (0)
#define ROS_ERROR_STREAM(x) do { std::cerr << "[ERROR] " << x << std::endl; } while (0)
