// parallel_for.h - the host builder's only threading primitive: run fn(begin, end, tid) over
// [0, n) cut into `chunks` equal ranges that the threads pull from a shared counter.
#pragma once
#include <algorithm>
#include <atomic>
#include <thread>
#include <vector>

namespace rtp {

inline int host_threads(int n_threads) {
  if (n_threads <= 0) n_threads = (int)std::thread::hardware_concurrency();
  return n_threads > 0 ? n_threads : 1;
}

template <class F>
void parallel_for(size_t n, int n_threads, size_t grain, F fn) {
  n_threads = host_threads(n_threads);
  if (grain < 1) grain = 1;
  const size_t chunks = (n + grain - 1) / grain;
  if (n_threads > (int)chunks) n_threads = (int)chunks;
  if (n_threads <= 1) {
    if (n) fn((size_t)0, n, 0);
    return;
  }
  std::atomic<size_t> next{0};
  auto work = [&](int tid) {
    for (;;) {
      size_t c = next.fetch_add(1, std::memory_order_relaxed);
      if (c >= chunks) return;
      fn(c * grain, std::min(n, (c + 1) * grain), tid);
    }
  };
  std::vector<std::thread> th;
  th.reserve((size_t)n_threads - 1);
  for (int t = 1; t < n_threads; ++t) th.emplace_back(work, t);
  work(0);
  for (auto &t : th) t.join();
}

}  // namespace rtp
