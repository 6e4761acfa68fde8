// Minimal stand-in for <ros/ros.h>: only what integration/ros1/*.cpp touches, so that the adapter sources can be
// SYNTAX-checked in an image without ROS (tests/test_export.py). Nothing here runs.
#pragma once
#include <stdexcept>
#include <string>
namespace ros {
struct Time { static void init() {} static Time now() { return Time(); } double toSec() const { return 0.0; } };
struct Rate { explicit Rate(double) {} void sleep() {} };
struct Subscriber {};
struct Publisher { template <class M> void publish(const M&) {} };
struct NodeHandle {
    template <class M, class T> Subscriber subscribe(const std::string&, int, void (T::*)(const M&), T*) { return Subscriber(); }
    template <class M> Publisher advertise(const std::string&, int) { return Publisher(); }
};
inline void init(int&, char**, const std::string&) {}
inline bool ok() { return false; }
inline void spinOnce() {}
}  // namespace ros
