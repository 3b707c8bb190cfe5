// B200 implementation behind the reference's CeresScanMatcher2D interface
// (cartographer/mapping/internal/2d/scan_matching/ceres_scan_matcher_2d.h:38-60).
//
// Match() keeps the reference's signature: it refines `initial_pose_estimate` against a
// ProbabilityGrid by minimising the occupied-space, translation-delta and rotation-delta
// residuals (ceres_scan_matcher_2d.cc:64-110).  The minimisation is the trust-region
// Levenberg-Marquardt iteration ceres::Solve runs for these options, executed on the
// device (csm_ceres_match2d_batch, one warp per problem, fp64).  MatchBatch refines many
// poses in one launch -- what ConstraintBuilder2D needs right after a batch of
// branch-and-bound searches (constraint_builder_2d.cc:245-249).  ProbabilityGrid
// (occupied_space_cost_function_2d.cc) and TSDF2D (tsdf_match_cost_function_2d.cc) grids.
// No CPU fallback.
#ifndef CARTOGRAPHER_MAPPING_INTERNAL_2D_SCAN_MATCHING_CERES_SCAN_MATCHER_2D_H_
#define CARTOGRAPHER_MAPPING_INTERNAL_2D_SCAN_MATCHING_CERES_SCAN_MATCHER_2D_H_

#include <vector>

#include "cartographer/mapping/internal/2d/scan_matching/fast_correlative_scan_matcher_2d.h"
#include "cartographer_compat.h"
#include "csm_b200.h"

namespace cartographer {
namespace mapping {
namespace scan_matching {

// Align scans with an existing map using the device Levenberg-Marquardt solver.
class CeresScanMatcher2D {
 public:
  explicit CeresScanMatcher2D(const proto::CeresScanMatcherOptions2D& options);
  CeresScanMatcher2D(const proto::CeresScanMatcherOptions2D& options, DeviceContext* context);
  virtual ~CeresScanMatcher2D();

  CeresScanMatcher2D(const CeresScanMatcher2D&) = delete;
  CeresScanMatcher2D& operator=(const CeresScanMatcher2D&) = delete;

  // Aligns 'point_cloud' within the 'grid' given an 'initial_pose_estimate' and returns a
  // 'pose_estimate' and the solver 'summary' (reference :48-54).
  void Match(const transform::Vector2d& target_translation,
             const transform::Rigid2d& initial_pose_estimate,
             const sensor::PointCloud& point_cloud, const Grid2D& grid,
             transform::Rigid2d* pose_estimate, ceres::Solver::Summary* summary) const;

  struct Query {
    transform::Vector2d target_translation;
    transform::Rigid2d initial_pose_estimate;
    const sensor::PointCloud* point_cloud = nullptr;
    const Grid2D* grid = nullptr;
    // Optional: the grid's cells already resident on the matcher's device (the submap's
    // precomputation stack keeps them, FastCorrelativeScanMatcher2D::device_cells()); then the
    // batch uploads nothing for this grid.  `grid` still supplies limits and the value table.
    const uint16* device_cells = nullptr;
  };
  struct Answer {
    transform::Rigid2d pose_estimate;
    ceres::Solver::Summary summary;
  };
  std::vector<Answer> MatchBatch(const std::vector<Query>& queries) const;

 private:
  const proto::CeresScanMatcherOptions2D options_;
  DeviceContext* context_ = nullptr;  // nullptr: the calling thread's default context
};

}  // namespace scan_matching
}  // namespace mapping
}  // namespace cartographer

#endif  // CARTOGRAPHER_MAPPING_INTERNAL_2D_SCAN_MATCHING_CERES_SCAN_MATCHER_2D_H_
