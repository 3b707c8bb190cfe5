// RotationalScanMatcher behind the reference's interface
// (cartographer/mapping/internal/3d/scan_matching/rotational_scan_matcher.h:28-52).
// Host code: O(num_angles x histogram_size) fp32 work per search whose results gate
// which yaw slices reach the device (fast_correlative_scan_matcher_3d.cc:274-281).
#ifndef CARTOGRAPHER_MAPPING_INTERNAL_3D_SCAN_MATCHING_ROTATIONAL_SCAN_MATCHER_H_
#define CARTOGRAPHER_MAPPING_INTERNAL_3D_SCAN_MATCHING_ROTATIONAL_SCAN_MATCHER_H_

#include <vector>

#include "cartographer_compat_3d.h"

namespace cartographer {
namespace mapping {
namespace scan_matching {

class RotationalScanMatcher {
 public:
  // Rotates 'histogram' by 'angle'; a fractional bucket is interpolated linearly.
  static Eigen::VectorXf RotateHistogram(const Eigen::VectorXf& histogram, float angle);

  explicit RotationalScanMatcher(const Eigen::VectorXf* histogram);

  // Scores in [0, 1] of 'histogram' rotated by 'initial_angle' + each of 'angles'
  // against the submap histogram given at construction.
  std::vector<float> Match(const Eigen::VectorXf& histogram, float initial_angle,
                           const std::vector<float>& angles) const;

 private:
  const Eigen::VectorXf* histogram_;
};

}  // namespace scan_matching
}  // namespace mapping
}  // namespace cartographer

#endif  // CARTOGRAPHER_MAPPING_INTERNAL_3D_SCAN_MATCHING_ROTATIONAL_SCAN_MATCHER_H_
