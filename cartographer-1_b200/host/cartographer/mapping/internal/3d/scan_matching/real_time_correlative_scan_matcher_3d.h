// B200 implementation behind the reference's RealTimeCorrelativeScanMatcher3D interface
// (cartographer/mapping/internal/3d/scan_matching/real_time_correlative_scan_matcher_3d.h:34-60).
//
// Match keeps the reference's signature.  The host generates the exhaustive search
// transforms, composes them with the initial pose and evaluates the motion penalty
// (libm); the device scores every candidate pose (csm_rt_match3d_batch).  The live
// HybridGrid is uploaded as a dense probability box on every Match call, as local SLAM
// changes it between calls; DeviceHybridGrid + MatchBatch serve callers that match
// several clouds against one grid state.
#ifndef CARTOGRAPHER_MAPPING_INTERNAL_3D_SCAN_MATCHING_REAL_TIME_CORRELATIVE_SCAN_MATCHER_3D_H_
#define CARTOGRAPHER_MAPPING_INTERNAL_3D_SCAN_MATCHING_REAL_TIME_CORRELATIVE_SCAN_MATCHER_3D_H_

#include <vector>

#include "cartographer/mapping/internal/2d/scan_matching/fast_correlative_scan_matcher_2d.h"
#include "cartographer_compat_3d.h"
#include "csm_b200.h"

namespace cartographer {
namespace mapping {
namespace scan_matching {

// A HybridGrid's probabilities resident on one device.
class DeviceHybridGrid {
 public:
  DeviceHybridGrid(DeviceContext* context, const HybridGrid& hybrid_grid);
  ~DeviceHybridGrid();
  DeviceHybridGrid(const DeviceHybridGrid&) = delete;
  DeviceHybridGrid& operator=(const DeviceHybridGrid&) = delete;
  const csm_grid3d* get() const { return grid_; }
  float resolution() const { return resolution_; }

 private:
  csm_grid3d* grid_ = nullptr;
  float resolution_;
  csm_ctx* build_context_ = nullptr;
};

class RealTimeCorrelativeScanMatcher3D {
 public:
  explicit RealTimeCorrelativeScanMatcher3D(const proto::RealTimeCorrelativeScanMatcherOptions& options);

  RealTimeCorrelativeScanMatcher3D(const RealTimeCorrelativeScanMatcher3D&) = delete;
  RealTimeCorrelativeScanMatcher3D& operator=(const RealTimeCorrelativeScanMatcher3D&) = delete;

  // Aligns 'point_cloud' within the 'hybrid_grid' given an 'initial_pose_estimate' then
  // updates 'pose_estimate' with the result and returns the score.
  float Match(const transform::Rigid3d& initial_pose_estimate, const sensor::PointCloud& point_cloud,
              const HybridGrid& hybrid_grid, transform::Rigid3d* pose_estimate) const;

  struct Query {
    transform::Rigid3d initial_pose_estimate;
    const sensor::PointCloud* point_cloud = nullptr;
    const DeviceHybridGrid* grid = nullptr;
  };
  struct Answer {
    float score = 0.f;
    transform::Rigid3d pose_estimate;
    int best_index = 0;       // generation index of the winning transform
    int num_candidates = 0;
  };
  std::vector<Answer> MatchBatch(DeviceContext* context, const std::vector<Query>& queries) const;

  // Visible for testing: the reference's private GenerateExhaustiveSearchTransforms.
  std::vector<transform::Rigid3f> GenerateExhaustiveSearchTransforms(
      float resolution, const sensor::PointCloud& point_cloud) const;

 private:
  const proto::RealTimeCorrelativeScanMatcherOptions options_;
};

}  // namespace scan_matching
}  // namespace mapping
}  // namespace cartographer

#endif  // CARTOGRAPHER_MAPPING_INTERNAL_3D_SCAN_MATCHING_REAL_TIME_CORRELATIVE_SCAN_MATCHER_3D_H_
