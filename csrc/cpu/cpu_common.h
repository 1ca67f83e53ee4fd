// Shared by the host compute ops (csrc/cpu): ISA clones picked at load time, and a small fork-join over index ranges.
#pragma once

#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#if defined(__GNUC__) && defined(__x86_64__)
#define DFM_CLONES __attribute__((target_clones("arch=haswell", "default")))
#else
#define DFM_CLONES
#endif

namespace dfcpu {

// fn(lo, hi) over [0, n) on up to `threads` threads (the caller's thread takes the first range); ranges smaller than
// `min_per_thread` are not worth a thread start
template <typename Fn>
void parallel_ranges(int64_t n, int threads, int64_t min_per_thread, Fn fn) {
  int t = threads < 1 ? 1 : threads;
  if (n / t < min_per_thread) t = static_cast<int>(n / min_per_thread);
  if (t <= 1) {
    fn(int64_t(0), n);
    return;
  }
  std::vector<std::thread> pool;
  const int64_t step = (n + t - 1) / t;
  for (int i = 1; i < t; ++i) {
    const int64_t lo = i * step, hi = lo + step < n ? lo + step : n;
    if (lo < hi) pool.emplace_back([=] { fn(lo, hi); });
  }
  fn(int64_t(0), step < n ? step : n);
  for (auto& th : pool) th.join();
}


}  // namespace dfcpu
