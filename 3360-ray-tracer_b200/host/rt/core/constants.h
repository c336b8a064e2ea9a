// Constants of src/core/constants.h:9-33.
#pragma once
#include <limits>
namespace rt::core {
constexpr double kPi = 3.14159265358979323846;
constexpr double kTwoPi = 2.0 * kPi;
constexpr double kHalfPi = 0.5 * kPi;
constexpr double kInvPi = 1.0 / kPi;
constexpr double kInvTwoPi = 1.0 / (2.0 * kPi);
constexpr double kInfinity = std::numeric_limits<double>::infinity();
constexpr double kEpsilon = 1e-8;
constexpr double DegreesToRadians(double deg) { return deg * (kPi / 180.0); }
constexpr double RadiansToDegrees(double rad) { return rad * (180.0 / kPi); }
constexpr double Clamp(double x, double lo, double hi) { return x < lo ? lo : (x > hi ? hi : x); }
}  // namespace rt::core
