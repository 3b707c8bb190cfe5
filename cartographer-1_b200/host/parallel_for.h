// Runs fn(i) for i in [0, n) on a few short-lived host threads.  Used for the
// per-search host preparation of a batch (search parameters, angle tables, rotational
// prefilter): independent, libm-heavy, a few microseconds to a millisecond each.
#ifndef CSM_HOST_PARALLEL_FOR_H_
#define CSM_HOST_PARALLEL_FOR_H_

#include <algorithm>
#include <atomic>
#include <cstddef>
#include <thread>
#include <vector>

namespace cartographer {
namespace common {

template <typename F>
void ParallelFor(const size_t n, const F& fn, const size_t min_items_per_thread = 8) {
  const size_t hw = std::max(1u, std::thread::hardware_concurrency());
  const size_t threads = std::min({hw, static_cast<size_t>(8), n / std::max<size_t>(1, min_items_per_thread)});
  if (threads <= 1) {
    for (size_t i = 0; i != n; ++i) fn(i);
    return;
  }
  std::atomic<size_t> next{0};
  auto work = [&]() {
    for (size_t i = next.fetch_add(1); i < n; i = next.fetch_add(1)) fn(i);
  };
  std::vector<std::thread> pool;
  pool.reserve(threads - 1);
  for (size_t t = 1; t < threads; ++t) pool.emplace_back(work);
  work();
  for (std::thread& t : pool) t.join();
}

}  // namespace common
}  // namespace cartographer

#endif  // CSM_HOST_PARALLEL_FOR_H_
