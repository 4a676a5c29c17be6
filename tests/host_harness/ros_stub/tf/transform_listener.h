// TEST-ONLY stand-in for tf (see ros/ros.h in this directory).
#pragma once
#include <stdexcept>
#include <string>

#include <ros/ros.h>

namespace tf {
struct Vector3 {
  double v[3];
  double operator[](int i) const { return v[i]; }
};
struct Matrix3x3 {
  Vector3 r[3];
  const Vector3& operator[](int i) const { return r[i]; }
};
struct StampedTransform {
  const Matrix3x3& getBasis() const;
  const Vector3& getOrigin() const;
};
struct TransformException : std::runtime_error {
  using std::runtime_error::runtime_error;
};
struct TransformListener {
  bool waitForTransform(const std::string& target, const std::string& source, const ros::Time& t, const ros::Duration& d);
  void lookupTransform(const std::string& target, const std::string& source, const ros::Time& t, StampedTransform& out);
};
}  // namespace tf
