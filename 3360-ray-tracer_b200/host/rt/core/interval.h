// rt::core::Interval (src/core/interval.h:9-45).  Default is the empty interval.
#pragma once
#include <algorithm>
#include "constants.h"
namespace rt::core {
class Interval {
 public:
  double min_, max_;
  Interval() : min_(+kInfinity), max_(-kInfinity) {}
  Interval(double lo, double hi) : min_(lo), max_(hi) {}
  Interval(const Interval& a, const Interval& b) : min_(std::min(a.min_, b.min_)), max_(std::max(a.max_, b.max_)) {}
  double Size() const { return max_ - min_; }
  bool Contains(double x) const { return min_ <= x && x <= max_; }
  bool Surrounds(double x) const { return min_ < x && x < max_; }
  double Clamp(double x) const { return x < min_ ? min_ : (x > max_ ? max_ : x); }
  // The reference's Expand shifts instead of widening (interval.h:23-26); kept as is, unused.
  Interval Expand(double delta) const { return Interval(min_ + delta / 2, max_ + delta / 2); }
};
}  // namespace rt::core
