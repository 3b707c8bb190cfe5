// common::FixedRatioSampler (reference: cartographer/common/fixed_ratio_sampler.{h,cc}):
// Pulse() returns true for a fixed ratio of calls, deterministically.
#ifndef CARTOGRAPHER_COMMON_FIXED_RATIO_SAMPLER_H_
#define CARTOGRAPHER_COMMON_FIXED_RATIO_SAMPLER_H_

#include <cstdint>

#include "cartographer_compat.h"

namespace cartographer {
namespace common {

class FixedRatioSampler {
 public:
  explicit FixedRatioSampler(double ratio) : ratio_(ratio) {
    CSM_CHECK(ratio >= 0.);
    CSM_CHECK(ratio <= 1.);
  }
  // fixed_ratio_sampler.cc:32-39: sample iff num_samples / num_pulses < ratio.
  bool Pulse() {
    ++num_pulses_;
    if (static_cast<double>(num_samples_) / num_pulses_ < ratio_) {
      ++num_samples_;
      return true;
    }
    return false;
  }

 private:
  const double ratio_;
  int64_t num_pulses_ = 0;
  int64_t num_samples_ = 0;
};

}  // namespace common
}  // namespace cartographer

#endif  // CARTOGRAPHER_COMMON_FIXED_RATIO_SAMPLER_H_
