#pragma once
#include <string>
#include <ros/ros.h>
namespace tf {
struct Vector3 { double x, y, z; Vector3(double a = 0, double b = 0, double c = 0) : x(a), y(b), z(c) {} };
struct Quaternion { double r = 0, p = 0, y = 0; void setRPY(double a, double b, double c) { r = a; p = b; y = c; } };
struct Transform { Vector3 o; Quaternion q; void setOrigin(const Vector3& v) { o = v; } void setRotation(const Quaternion& v) { q = v; } };
struct StampedTransform { Transform t; ros::Time stamp; std::string parent, child; StampedTransform(const Transform& a, const ros::Time& s, const std::string& p, const std::string& c) : t(a), stamp(s), parent(p), child(c) {} };
struct TransformBroadcaster { int sent = 0; StampedTransform* last = nullptr; void sendTransform(const StampedTransform& s) { sent++; delete last; last = new StampedTransform(s); } };
}
