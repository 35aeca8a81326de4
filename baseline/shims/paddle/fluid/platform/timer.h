// Shim for the reference arm: wall-clock stopwatch with the reference's method names.
#pragma once
#include <chrono>
#include <cstdint>
namespace paddle {
namespace platform {
class Timer {
 public:
  Timer() { Reset(); }
  void Reset() { elapsed_us_ = 0; paused_ = true; count_ = 0; }
  void Start() { start_ = clock::now(); paused_ = false; }
  void Pause() {
    if (paused_) return;
    elapsed_us_ += std::chrono::duration_cast<std::chrono::microseconds>(clock::now() - start_).count();
    ++count_;
    paused_ = true;
  }
  void Resume() { Start(); }
  int Count() { return count_; }
  double ElapsedUS() { return (double)elapsed_us_; }
  double ElapsedMS() { return elapsed_us_ / 1e3; }
  double ElapsedSec() { return elapsed_us_ / 1e6; }
 private:
  using clock = std::chrono::steady_clock;
  clock::time_point start_;
  int64_t elapsed_us_;
  int count_;
  bool paused_;
};
}  // namespace platform
}  // namespace paddle
