// rt::core::Vec3 — host-side double 3-vector with the reference's surface
// (src/core/vec3.h:8-105).  Arithmetic order is part of the contract: the SAH
// builder and Camera::Initialize must reproduce the reference's doubles bit for
// bit, so Dot sums left to right, Cross is (a1*b2 - a2*b1, ...), and division
// by a scalar multiplies by the rounded reciprocal (vec3.h:40-42,91-93).
#pragma once
#include <cmath>
#include <ostream>

namespace rt::core {

class Vec3 {
 public:
  constexpr Vec3() : c_{0.0, 0.0, 0.0} {}
  constexpr Vec3(double x, double y, double z) : c_{x, y, z} {}

  constexpr double x() const { return c_[0]; }
  constexpr double y() const { return c_[1]; }
  constexpr double z() const { return c_[2]; }
  constexpr double operator[](int i) const { return c_[i]; }
  double& operator[](int i) { return c_[i]; }

  constexpr Vec3 operator-() const { return {-c_[0], -c_[1], -c_[2]}; }
  Vec3& operator+=(const Vec3& o) {
    for (int i = 0; i < 3; ++i) c_[i] += o.c_[i];
    return *this;
  }
  Vec3& operator*=(double s) {
    for (int i = 0; i < 3; ++i) c_[i] *= s;
    return *this;
  }
  Vec3& operator/=(double s) { return *this *= (1.0 / s); }

  constexpr double length_squared() const { return c_[0] * c_[0] + c_[1] * c_[1] + c_[2] * c_[2]; }
  double length() const { return std::sqrt(length_squared()); }
  bool NearZero() const {
    constexpr double tiny = 1e-8;
    return std::fabs(c_[0]) < tiny && std::fabs(c_[1]) < tiny && std::fabs(c_[2]) < tiny;
  }

 private:
  double c_[3];
};

using Point3 = Vec3;

inline std::ostream& operator<<(std::ostream& os, const Vec3& v) {
  return os << v[0] << ' ' << v[1] << ' ' << v[2];
}
inline Vec3 operator+(const Vec3& a, const Vec3& b) { return {a[0] + b[0], a[1] + b[1], a[2] + b[2]}; }
inline Vec3 operator-(const Vec3& a, const Vec3& b) { return {a[0] - b[0], a[1] - b[1], a[2] - b[2]}; }
inline Vec3 operator*(const Vec3& a, const Vec3& b) { return {a[0] * b[0], a[1] * b[1], a[2] * b[2]}; }
inline Vec3 operator*(double s, const Vec3& v) { return {s * v[0], s * v[1], s * v[2]}; }
inline Vec3 operator*(const Vec3& v, double s) { return s * v; }
inline Vec3 operator/(const Vec3& v, double s) { return (1.0 / s) * v; }
inline double Dot(const Vec3& a, const Vec3& b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
inline Vec3 Cross(const Vec3& a, const Vec3& b) {
  return {a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]};
}

}  // namespace rt::core
