// Native self-test of the host runtime primitives (role of the reference's gtests threadpool_test.cc and
// the channel tests): blocking MPMC channel semantics and the thread pool's range split.
// Built and run by tests/test_host_runtime.py::test_native_util_selftest.
#include <cassert>
#include <cstdio>
#include <numeric>

#include "../util.h"

namespace pbh {
void set_error(const char*, ...) {}
}  // namespace pbh

#define CHECK(cond)                                                        \
  do {                                                                     \
    if (!(cond)) {                                                         \
      fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond);    \
      return 1;                                                            \
    }                                                                      \
  } while (0)

int main() {
  using namespace pbh;
  // ---- channel: every item produced by 4 writers reaches exactly one of 3 readers; close() drains
  {
    Channel<int> ch(64);
    std::atomic<long long> sum{0};
    std::atomic<int> got{0};
    std::vector<std::thread> readers, writers;
    for (int r = 0; r < 3; ++r)
      readers.emplace_back([&] {
        std::vector<int> block;
        for (;;) {
          block.clear();
          if (ch.read(block, 17) == 0) {
            if (ch.closed()) return;
            continue;
          }
          for (int v : block) { sum += v; ++got; }
        }
      });
    for (int w = 0; w < 4; ++w)
      writers.emplace_back([&, w] { for (int i = 0; i < 5000; ++i) ch.put(w * 5000 + i); });
    for (auto& t : writers) t.join();
    ch.close();
    for (auto& t : readers) t.join();
    CHECK(got.load() == 20000);
    CHECK(sum.load() == 19999LL * 20000 / 2);
    CHECK(!ch.put(1));                       // a closed channel refuses writes
    int v;
    CHECK(!ch.get(v));                       // and reads nothing once drained
    ch.open();
    CHECK(ch.put(7) && ch.get(v) && v == 7);
  }
  // ---- bounded capacity blocks the writer until a reader makes room
  {
    Channel<int> ch(2);
    std::atomic<int> written{0};
    std::thread w([&] { for (int i = 0; i < 5; ++i) { ch.put(int(i)); ++written; } });
    std::this_thread::sleep_for(std::chrono::milliseconds(50));
    CHECK(written.load() == 2);
    int v;
    for (int i = 0; i < 5; ++i) CHECK(ch.get(v) && v == i);
    w.join();
  }
  // ---- thread pool: parallel_range covers [0, n) exactly once, also when n < threads
  {
    ThreadPool pool(5);
    for (size_t n : {0ul, 1ul, 3ul, 5ul, 1000ul, 1003ul}) {
      std::vector<std::atomic<int>> hits(n);
      pool.parallel_range(n, [&](size_t b, size_t e, int) { for (size_t i = b; i < e; ++i) ++hits[i]; });
      for (size_t i = 0; i < n; ++i) CHECK(hits[i].load() == 1);
    }
    std::atomic<int> ran{0};
    std::vector<std::future<void>> fs;
    for (int i = 0; i < 100; ++i) fs.push_back(pool.run([&] { ++ran; }));
    for (auto& f : fs) f.get();
    CHECK(ran.load() == 100);
  }
  CHECK(mix64(1) != mix64(2) && mix64(0) == 0);
  printf("util selftest ok\n");
  return 0;
}
