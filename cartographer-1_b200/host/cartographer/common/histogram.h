// common::Histogram as the constraint builders use it (reference: cartographer/common/histogram.{h,cc}):
// collects the scores of the constraints found and prints them as `buckets` equal-width buckets
// between the smallest and the largest value, with a 20-character bar and running totals
// (constraint_builder_2d.cc:238-240,289-293).
#ifndef CARTOGRAPHER_COMMON_HISTOGRAM_H_
#define CARTOGRAPHER_COMMON_HISTOGRAM_H_

#include <algorithm>
#include <cstdio>
#include <numeric>
#include <string>
#include <vector>

#include "cartographer_compat.h"

namespace cartographer {
namespace common {

class Histogram {
 public:
  void Add(float value) { values_.push_back(value); }
  size_t size() const { return values_.size(); }

  std::string ToString(int buckets) const {
    CSM_CHECK(buckets >= 1);
    if (values_.empty()) return "Count: 0";
    const auto range = std::minmax_element(values_.begin(), values_.end());
    const float min = *range.first, max = *range.second;
    const float mean = std::accumulate(values_.begin(), values_.end(), 0.f) / values_.size();
    char line[256];
    std::snprintf(line, sizeof(line), "Count: %zu  Min: %g  Max: %g  Mean: %g", values_.size(), min, max, mean);
    std::string result = line;
    if (min == max) return result;
    float lower = min;
    size_t total = 0;
    for (int i = 0; i != buckets; ++i) {
      const bool last = i + 1 == buckets;
      const float upper = last ? max : (max * (i + 1) / buckets + min * (buckets - i - 1) / buckets);
      size_t count = 0;
      for (const float v : values_)
        if (lower <= v && (last ? v <= upper : v < upper)) ++count;
      total += count;
      constexpr int kMaxBarChars = 20;
      const int bar = static_cast<int>((count * kMaxBarChars + values_.size() / 2) / values_.size());
      std::snprintf(line, sizeof(line), "\n[%f, %f%c\t", lower, upper, last ? ']' : ')');
      result += line;
      result += std::string(kMaxBarChars - bar, ' ') + std::string(bar, '#');
      std::snprintf(line, sizeof(line), "\tCount: %zu (%g%%)\tTotal: %zu (%g%%)", count,
                    count * 1e2f / values_.size(), total, total * 1e2f / values_.size());
      result += line;
      lower = upper;
    }
    return result;
  }

 private:
  std::vector<float> values_;
};

}  // namespace common
}  // namespace cartographer

#endif  // CARTOGRAPHER_COMMON_HISTOGRAM_H_
