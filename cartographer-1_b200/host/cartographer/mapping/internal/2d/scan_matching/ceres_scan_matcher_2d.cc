#include "cartographer/mapping/internal/2d/scan_matching/ceres_scan_matcher_2d.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace cartographer {
namespace mapping {
namespace scan_matching {

CeresScanMatcher2D::CeresScanMatcher2D(const proto::CeresScanMatcherOptions2D& options)
    : options_(options) {}

CeresScanMatcher2D::CeresScanMatcher2D(const proto::CeresScanMatcherOptions2D& options,
                                       DeviceContext* context)
    : options_(options), context_(context) {}

CeresScanMatcher2D::~CeresScanMatcher2D() {}

void CeresScanMatcher2D::Match(const transform::Vector2d& target_translation,
                               const transform::Rigid2d& initial_pose_estimate,
                               const sensor::PointCloud& point_cloud, const Grid2D& grid,
                               transform::Rigid2d* const pose_estimate,
                               ceres::Solver::Summary* const summary) const {
  CSM_CHECK(pose_estimate != nullptr);
  Query q;
  q.target_translation = target_translation;
  q.initial_pose_estimate = initial_pose_estimate;
  q.point_cloud = &point_cloud;
  q.grid = &grid;
  const Answer a = MatchBatch({q})[0];
  *pose_estimate = a.pose_estimate;
  if (summary != nullptr) *summary = a.summary;
}

std::vector<CeresScanMatcher2D::Answer> CeresScanMatcher2D::MatchBatch(
    const std::vector<Query>& queries) const {
  const size_t n = queries.size();
  std::vector<Answer> answers(n);
  if (n == 0) return answers;
  // reference :75, :94, :99
  CSM_CHECK(options_.occupied_space_weight() > 0.);
  CSM_CHECK(options_.translation_weight() > 0.);
  CSM_CHECK(options_.rotation_weight() > 0.);
  DeviceContext* ctx = context_ != nullptr ? context_ : DeviceContext::ThreadDefault();
  std::vector<std::vector<float>> xyz(n);
  std::vector<csm_ceres_job2d> jobs(n);
  for (size_t i = 0; i != n; ++i) {
    const Query& q = queries[i];
    CSM_CHECK(q.point_cloud != nullptr && q.grid != nullptr);
    CSM_CHECK(!q.point_cloud->empty());
    const MapLimits& limits = q.grid->limits();
    xyz[i].resize(3 * q.point_cloud->size());
    for (size_t p = 0; p != q.point_cloud->size(); ++p) {
      const sensor::Vector3f& v = (*q.point_cloud)[p].position;
      xyz[i][3 * p] = v.x();
      xyz[i][3 * p + 1] = v.y();
      xyz[i][3 * p + 2] = v.z();
    }
    csm_ceres_job2d& job = jobs[i];
    std::memset(&job, 0, sizeof(job));
    job.cells = q.grid->correspondence_cost_cells().data();
    job.cells_dev = q.device_cells;
    job.num_x = limits.cell_limits().num_x_cells;
    job.num_y = limits.cell_limits().num_y_cells;
    job.cost_table = q.grid->value_to_correspondence_cost_table().data();
    job.points_xyz = xyz[i].data();
    job.num_points = static_cast<int32_t>(q.point_cloud->size());
    job.max_num_iterations = options_.ceres_solver_options().max_num_iterations();
    job.use_nonmonotonic_steps = options_.ceres_solver_options().use_nonmonotonic_steps() ? 1 : 0;
    job.max_x = limits.max().x();
    job.max_y = limits.max().y();
    job.resolution = limits.resolution();
    job.initial_pose[0] = q.initial_pose_estimate.translation().x();
    job.initial_pose[1] = q.initial_pose_estimate.translation().y();
    job.initial_pose[2] = q.initial_pose_estimate.rotation().angle();
    job.target_translation[0] = q.target_translation.x();
    job.target_translation[1] = q.target_translation.y();
    job.occupied_space_weight = options_.occupied_space_weight();
    job.translation_weight = options_.translation_weight();
    job.rotation_weight = options_.rotation_weight();
    if (q.grid->GetGridType() == GridType::TSDF) {  // reference :84-91
      const TSDF2D& tsdf = static_cast<const TSDF2D&>(*q.grid);
      job.cost_table = tsdf.value_converter().value_to_tsd_table().data();
      job.weight_cells = tsdf.weight_cells().data();
      job.weight_table = tsdf.value_converter().value_to_weight_table().data();
      job.truncation_distance = tsdf.GetMaxCorrespondenceCost();
    }
  }
  std::vector<csm_ceres_result2d> results(n);
  if (csm_ceres_match2d_batch(ctx->get(), jobs.data(), static_cast<int>(n), results.data()) != CSM_OK) {
    std::fprintf(stderr, "csm_ceres_match2d_batch failed: %s\n", csm_last_error());
    std::abort();
  }
  static const char* const kWhy[] = {"gradient tolerance reached", "parameter tolerance reached",
                                     "function tolerance reached", "maximum number of iterations reached",
                                     "minimum trust region radius reached", "too many invalid steps"};
  for (size_t i = 0; i != n; ++i) {
    const csm_ceres_result2d& r = results[i];
    Answer& a = answers[i];
    a.pose_estimate = transform::Rigid2d(transform::Vector2d(r.pose[0], r.pose[1]), r.pose[2]);
    a.summary.initial_cost = r.initial_cost;
    a.summary.final_cost = r.final_cost;
    a.summary.num_iterations = r.iterations;
    a.summary.num_successful_steps = r.num_successful_steps;
    a.summary.num_unsuccessful_steps = r.num_unsuccessful_steps;
    a.summary.termination_type = r.termination <= 2 ? ceres::CONVERGENCE
                                 : r.termination <= 4 ? ceres::NO_CONVERGENCE
                                                      : ceres::FAILURE;
    a.summary.message = kWhy[r.termination >= 0 && r.termination <= 5 ? r.termination : 5];
  }
  return answers;
}

}  // namespace scan_matching
}  // namespace mapping
}  // namespace cartographer
