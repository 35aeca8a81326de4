// Host runtime utilities: error plumbing, blocking MPMC channel, thread pool.
// Reference roles: framework/channel.h:39-481 (ChannelObject), framework/threadpool.h
// (ThreadPool, parallel_run_range), platform/timer.h.
#pragma once
#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <deque>
#include <functional>
#include <future>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace pbh {

void set_error(const char* fmt, ...);

#define PBH_REQUIRE(cond, ...)        \
  do {                                \
    if (!(cond)) {                    \
      ::pbh::set_error(__VA_ARGS__);  \
      return -1;                      \
    }                                 \
  } while (0)

// Bounded blocking MPMC queue with batched reads/writes and close semantics.
template <typename T>
class Channel {
 public:
  explicit Channel(size_t capacity = 0) : cap_(capacity ? capacity : SIZE_MAX) {}
  void set_capacity(size_t c) { std::lock_guard<std::mutex> g(mu_); cap_ = c ? c : SIZE_MAX; }
  void open() { std::lock_guard<std::mutex> g(mu_); closed_ = false; }
  void close() {
    { std::lock_guard<std::mutex> g(mu_); closed_ = true; }
    not_empty_.notify_all();
    not_full_.notify_all();
  }
  bool closed() const { std::lock_guard<std::mutex> g(mu_); return closed_; }
  size_t size() const { std::lock_guard<std::mutex> g(mu_); return q_.size(); }
  void clear() { std::lock_guard<std::mutex> g(mu_); q_.clear(); }

  bool put(T&& v) {
    std::unique_lock<std::mutex> lk(mu_);
    not_full_.wait(lk, [&] { return closed_ || q_.size() < cap_; });
    if (closed_) return false;
    q_.emplace_back(std::move(v));
    lk.unlock();
    not_empty_.notify_one();
    return true;
  }
  // moves every element of `vs` in (blocks while full); returns number written
  size_t write_move(std::vector<T>& vs) {
    size_t n = 0;
    for (auto& v : vs) {
      if (!put(std::move(v))) break;
      ++n;
    }
    vs.clear();
    return n;
  }
  bool get(T& out) {
    std::unique_lock<std::mutex> lk(mu_);
    not_empty_.wait(lk, [&] { return closed_ || !q_.empty(); });
    if (q_.empty()) return false;
    out = std::move(q_.front());
    q_.pop_front();
    lk.unlock();
    not_full_.notify_one();
    return true;
  }
  // reads up to `max_n` elements, blocking only for the first one
  size_t read(std::vector<T>& out, size_t max_n) {
    std::unique_lock<std::mutex> lk(mu_);
    not_empty_.wait(lk, [&] { return closed_ || !q_.empty(); });
    size_t n = 0;
    while (n < max_n && !q_.empty()) {
      out.emplace_back(std::move(q_.front()));
      q_.pop_front();
      ++n;
    }
    lk.unlock();
    if (n) not_full_.notify_all();
    return n;
  }

 private:
  mutable std::mutex mu_;
  std::condition_variable not_empty_, not_full_;
  std::deque<T> q_;
  size_t cap_;
  bool closed_ = false;
};

class ThreadPool {
 public:
  explicit ThreadPool(int n) {
    if (n < 1) n = 1;
    for (int i = 0; i < n; ++i) workers_.emplace_back([this] { loop(); });
  }
  ~ThreadPool() {
    { std::lock_guard<std::mutex> g(mu_); stop_ = true; }
    cv_.notify_all();
    for (auto& t : workers_) t.join();
  }
  int size() const { return (int)workers_.size(); }
  std::future<void> run(std::function<void()> fn) {
    auto task = std::make_shared<std::packaged_task<void()>>(std::move(fn));
    std::future<void> f = task->get_future();
    { std::lock_guard<std::mutex> g(mu_); q_.emplace_back([task] { (*task)(); }); }
    cv_.notify_one();
    return f;
  }
  // fn(begin, end) over [0, n) split evenly (reference parallel_run_range)
  void parallel_range(size_t n, const std::function<void(size_t, size_t, int)>& fn) {
    int parts = size();
    if ((size_t)parts > n) parts = (int)(n ? n : 1);
    std::vector<std::future<void>> fs;
    size_t per = (n + parts - 1) / parts;
    for (int p = 0; p < parts; ++p) {
      size_t b = p * per, e = std::min(n, b + per);
      if (b >= e) break;
      fs.emplace_back(run([=, &fn] { fn(b, e, p); }));
    }
    for (auto& f : fs) f.get();
  }
  // (only for threads known to be alive: an exited, unjoined thread has tid 0 and the call would pin the
  //  CALLER — short-lived threads must pin themselves with pthread_self())
  static bool set_cpu_affinity(std::thread& t, int core);

 private:
  void loop() {
    for (;;) {
      std::function<void()> job;
      {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [&] { return stop_ || !q_.empty(); });
        if (stop_ && q_.empty()) return;
        job = std::move(q_.front());
        q_.pop_front();
      }
      job();
    }
  }
  std::vector<std::thread> workers_;
  std::deque<std::function<void()>> q_;
  std::mutex mu_;
  std::condition_variable cv_;
  bool stop_ = false;
};

struct Timer {
  using clk = std::chrono::steady_clock;
  clk::time_point t0;
  double total = 0;
  void start() { t0 = clk::now(); }
  void pause() { total += std::chrono::duration<double>(clk::now() - t0).count(); }
  double seconds() const { return total; }
};

// In-place LSD radix sort of 64-bit keys (8-bit digits; a digit that is the same in every key costs one
// counting pass and no move).  ~3x std::sort on random keys above a few thousand elements.
inline void radix_sort_u64(std::vector<uint64_t>& v) {
  const size_t n = v.size();
  if (n < 4096) { std::sort(v.begin(), v.end()); return; }
  std::vector<uint64_t> tmp(n);
  uint64_t* src = v.data();
  uint64_t* dst = tmp.data();
  for (int shift = 0; shift < 64; shift += 8) {
    size_t cnt[256] = {0};
    for (size_t i = 0; i < n; ++i) ++cnt[(src[i] >> shift) & 255];
    bool single = false;
    for (int d = 0; d < 256; ++d) if (cnt[d] == n) { single = true; break; }
    if (single) continue;
    size_t pos[256];
    size_t acc = 0;
    for (int d = 0; d < 256; ++d) { pos[d] = acc; acc += cnt[d]; }
    for (size_t i = 0; i < n; ++i) dst[pos[(src[i] >> shift) & 255]++] = src[i];
    std::swap(src, dst);
  }
  if (src != v.data()) memcpy(v.data(), src, n * sizeof(uint64_t));
}

inline uint64_t mix64(uint64_t k) {
  k ^= k >> 33; k *= 0xff51afd7ed558ccdull; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ull; k ^= k >> 33;
  return k;
}

}  // namespace pbh
