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

// The yaw prefilter of the 3D matcher: a node's histogram of surface directions (angles folded into
// [0, pi), one weight per bucket; made by the sensor pipeline, rotational_scan_matcher.cc:164-176) is
// turned by every candidate yaw and compared with the submap's.  A candidate whose score stays
// under min_rotational_score never reaches the device.
class RotationalScanMatcher {
 public:
  // The submap's histogram; borrowed, must outlive the matcher.
  explicit RotationalScanMatcher(const Eigen::VectorXf* histogram);

  // One score in [0, 1] per entry of `angles`: the histogram turned by initial_angle + angles[i]
  // against the submap's, as the dot product of the two normalised vectors -- 1.f outright when the
  // product of the norms is under 1e-3 (rotational_scan_matcher.cc:121-133,178-189).  fp32 in the
  // reference's operation order.
  std::vector<float> Match(const Eigen::VectorXf& histogram, float initial_angle,
                           const std::vector<float>& angles) const;

  // `histogram` turned by `angle`: the turn is a (generally fractional) number of buckets, and each
  // output bucket is the linear blend of the two input buckets it falls between, the sequence
  // wrapping around at pi (rotational_scan_matcher.cc:135-162).
  static Eigen::VectorXf RotateHistogram(const Eigen::VectorXf& histogram, float angle);

 private:
  const Eigen::VectorXf* histogram_;
};

}  // namespace scan_matching
}  // namespace mapping
}  // namespace cartographer

#endif  // CARTOGRAPHER_MAPPING_INTERNAL_3D_SCAN_MATCHING_ROTATIONAL_SCAN_MATCHER_H_
