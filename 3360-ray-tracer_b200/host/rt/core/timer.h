// rt::core::Timer (src/core/timer.h:7-21).
#pragma once
#include <chrono>
namespace rt::core {
class Timer {
 public:
  Timer() { reset(); }
  void reset() { t0_ = std::chrono::steady_clock::now(); }
  double elapsed() const { return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0_).count(); }
 private:
  std::chrono::steady_clock::time_point t0_;
};
}  // namespace rt::core
