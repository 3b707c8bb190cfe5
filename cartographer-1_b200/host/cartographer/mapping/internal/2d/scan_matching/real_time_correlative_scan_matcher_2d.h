// B200 implementation behind the reference's RealTimeCorrelativeScanMatcher2D
// interface (cartographer/mapping/internal/2d/scan_matching/
// real_time_correlative_scan_matcher_2d.h:55-89).
//
// Match() keeps the reference's signature and semantics: exhaustive search of the
// (rotation, x, y) window around the initial pose directly on the live grid, score =
// mean probability times the motion penalty, first maximum in generation order
// wins.  The libm-dependent pieces (SearchParameters from the pre-rotated cloud, the
// per-scan half-angle table, the fp64 penalty exp(-(hypot*w_t + |theta|*w_r)^2) per
// candidate) are computed here; rotation, discretisation and scoring run on the
// device through csm_rt_match2d_batch.  MatchBatch coalesces many scans (one per
// robot / trajectory) into one launch.  No CPU fallback.
#ifndef CARTOGRAPHER_MAPPING_INTERNAL_2D_SCAN_MATCHING_REAL_TIME_CORRELATIVE_SCAN_MATCHER_2D_H_
#define CARTOGRAPHER_MAPPING_INTERNAL_2D_SCAN_MATCHING_REAL_TIME_CORRELATIVE_SCAN_MATCHER_2D_H_

#include <vector>

#include "cartographer/mapping/internal/2d/scan_matching/correlative_scan_matcher_2d.h"
#include "cartographer/mapping/internal/2d/scan_matching/fast_correlative_scan_matcher_2d.h"
#include "cartographer_compat.h"
#include "csm_b200.h"

namespace cartographer {
namespace mapping {
namespace scan_matching {

class RealTimeCorrelativeScanMatcher2D {
 public:
  explicit RealTimeCorrelativeScanMatcher2D(
      const proto::RealTimeCorrelativeScanMatcherOptions& options);
  // Same, on an explicit device context.
  RealTimeCorrelativeScanMatcher2D(const proto::RealTimeCorrelativeScanMatcherOptions& options,
                                   DeviceContext* context);

  RealTimeCorrelativeScanMatcher2D(const RealTimeCorrelativeScanMatcher2D&) = delete;
  RealTimeCorrelativeScanMatcher2D& operator=(const RealTimeCorrelativeScanMatcher2D&) = delete;

  // Aligns 'point_cloud' within the 'grid' given an 'initial_pose_estimate' then
  // updates 'pose_estimate' with the result and returns the score (reference :60-66).
  double Match(const transform::Rigid2d& initial_pose_estimate,
               const sensor::PointCloud& point_cloud, const Grid2D& grid,
               transform::Rigid2d* pose_estimate) const;

  // Computes the score for each Candidate2D in a collection.  The cost is computed as the sum of
  // probabilities or normalized TSD values, different from the Ceres CostFunctions:
  // http://ceres-solver.org/modeling.html  Visible for testing (reference :68-79, .cc:151-176).
  // The per-candidate sums run on the device (csm_rt_score_candidates2d), the motion penalty is
  // applied here in double like the reference does.
  void ScoreCandidates(const Grid2D& grid, const std::vector<DiscreteScan2D>& discrete_scans,
                       const SearchParameters& search_parameters,
                       std::vector<Candidate2D>* candidates) const;

  // Generation order scan -> x -> y of the exhaustive window (reference :81-83, .cc:83-115).
  std::vector<Candidate2D> GenerateExhaustiveSearchCandidates(
      const SearchParameters& search_parameters) const;

  struct Query {
    transform::Rigid2d initial_pose_estimate;
    const sensor::PointCloud* point_cloud = nullptr;
    const Grid2D* grid = nullptr;
  };
  struct Answer {
    double score = 0.;
    transform::Rigid2d pose_estimate;
    Candidate2D best_candidate{0, 0, 0, SearchParameters(0, 0, 0., 0.)};
  };
  // All queries in one device batch (grids shared between queries are uploaded once).
  std::vector<Answer> MatchBatch(const std::vector<Query>& queries) const;

 private:
  const proto::RealTimeCorrelativeScanMatcherOptions options_;
  DeviceContext* context_ = nullptr;  // nullptr: the calling thread's default context
};

}  // namespace scan_matching
}  // namespace mapping
}  // namespace cartographer

#endif  // CARTOGRAPHER_MAPPING_INTERNAL_2D_SCAN_MATCHING_REAL_TIME_CORRELATIVE_SCAN_MATCHER_2D_H_
