#include "cartographer/mapping/internal/3d/scan_matching/real_time_correlative_scan_matcher_3d.h"

#include <algorithm>
#include <cmath>

#include "cartographer/mapping/internal/3d/scan_matching/fast_correlative_scan_matcher_3d.h"
#include "parallel_for.h"

namespace cartographer {
namespace mapping {
namespace scan_matching {
namespace {

void CheckOk(int status, const char* what) {
  if (status != CSM_OK) {
    std::fprintf(stderr, "%s failed: %s\n", what, csm_last_error());
    std::abort();
  }
}

// transform.h:33-37 with FloatType = float
float GetAngle(const transform::Rigid3f& t) {
  const transform::Quaternionf& q = t.rotation();
  return 2.f * std::atan2(transform::Norm(transform::Vector3f(q.x(), q.y(), q.z())), std::abs(q.w()));
}

}  // namespace

DeviceHybridGrid::DeviceHybridGrid(DeviceContext* const context, const HybridGrid& hybrid_grid)
    : resolution_(hybrid_grid.resolution()), build_context_(context->get()) {
  std::vector<int32_t> xyz;
  std::vector<uint16> values;
  hybrid_grid.Export(&xyz, &values);
  CheckOk(csm_grid3d_build(build_context_, xyz.data(), values.data(), static_cast<int>(values.size()),
                           ValueToProbabilityTable().data(), &grid_),
          "csm_grid3d_build");
}

DeviceHybridGrid::~DeviceHybridGrid() { csm_grid3d_destroy(build_context_, grid_); }

RealTimeCorrelativeScanMatcher3D::RealTimeCorrelativeScanMatcher3D(
    const proto::RealTimeCorrelativeScanMatcherOptions& options)
    : options_(options) {}

float RealTimeCorrelativeScanMatcher3D::Match(const transform::Rigid3d& initial_pose_estimate,
                                              const sensor::PointCloud& point_cloud,
                                              const HybridGrid& hybrid_grid,
                                              transform::Rigid3d* const pose_estimate) const {
  CSM_CHECK(pose_estimate != nullptr);
  DeviceContext* ctx = DeviceContext::ThreadDefault();
  const DeviceHybridGrid grid(ctx, hybrid_grid);
  Query q;
  q.initial_pose_estimate = initial_pose_estimate;
  q.point_cloud = &point_cloud;
  q.grid = &grid;
  const Answer a = MatchBatch(ctx, {q})[0];
  *pose_estimate = a.pose_estimate;
  return a.score;
}

// reference :55-98
std::vector<transform::Rigid3f> RealTimeCorrelativeScanMatcher3D::GenerateExhaustiveSearchTransforms(
    const float resolution, const sensor::PointCloud& point_cloud) const {
  std::vector<transform::Rigid3f> result;
  const int linear_window_size = common::RoundToInt(options_.linear_search_window() / resolution);
  // Something on the order of resolution keeps the acos() below defined.
  float max_scan_range = 3.f * resolution;
  for (const sensor::RangefinderPoint& point : point_cloud)
    max_scan_range = std::max(transform::Norm(point.position), max_scan_range);
  const float kSafetyMargin = 1.f - 1e-3f;
  const float angular_step_size =
      kSafetyMargin *
      std::acos(1.f - (resolution * resolution) / (2.f * (max_scan_range * max_scan_range)));
  const int angular_window_size = common::RoundToInt(options_.angular_search_window() / angular_step_size);
  const size_t linear = 2 * linear_window_size + 1, angular = 2 * angular_window_size + 1;
  result.reserve(linear * linear * linear * angular * angular * angular);
  for (int z = -linear_window_size; z <= linear_window_size; ++z)
    for (int y = -linear_window_size; y <= linear_window_size; ++y)
      for (int x = -linear_window_size; x <= linear_window_size; ++x)
        for (int rz = -angular_window_size; rz <= angular_window_size; ++rz)
          for (int ry = -angular_window_size; ry <= angular_window_size; ++ry)
            for (int rx = -angular_window_size; rx <= angular_window_size; ++rx)
              result.emplace_back(
                  transform::Vector3f(x * resolution, y * resolution, z * resolution),
                  transform::AngleAxisVectorToRotationQuaternion(transform::Vector3f(
                      rx * angular_step_size, ry * angular_step_size, rz * angular_step_size)));
  return result;
}

std::vector<RealTimeCorrelativeScanMatcher3D::Answer> RealTimeCorrelativeScanMatcher3D::MatchBatch(
    DeviceContext* const context, const std::vector<Query>& queries) const {
  const size_t n = queries.size();
  std::vector<Answer> answers(n);
  if (n == 0) return answers;
  std::vector<std::vector<float>> poses(n), points(n);
  std::vector<std::vector<double>> penalty(n);
  std::vector<csm_rt_job3d> jobs(n);
  for (size_t i = 0; i != n; ++i) {
    const Query& q = queries[i];
    CSM_CHECK(q.point_cloud != nullptr && q.grid != nullptr);
    CSM_CHECK(!q.point_cloud->empty());  // the reference divides by the size and CHECKs score > 0
    const std::vector<transform::Rigid3f> transforms =
        GenerateExhaustiveSearchTransforms(q.grid->resolution(), *q.point_cloud);
    const transform::Rigid3f initial = transform::CastToFloat(q.initial_pose_estimate);
    const size_t k = transforms.size();
    poses[i].resize(7 * k);
    penalty[i].resize(k);
    common::ParallelFor(k, [&](const size_t c) {
      const transform::Rigid3f& t = transforms[c];
      const transform::Rigid3f candidate = initial * t;  // reference :45-46
      float* p = poses[i].data() + 7 * c;
      p[0] = candidate.translation().x();
      p[1] = candidate.translation().y();
      p[2] = candidate.translation().z();
      p[3] = candidate.rotation().w();
      p[4] = candidate.rotation().x();
      p[5] = candidate.rotation().y();
      p[6] = candidate.rotation().z();
      // reference :108-112: float norm and angle times double weights, squared, exp
      const float angle = GetAngle(t);
      const double cost = transform::Norm(t.translation()) * options_.translation_delta_cost_weight() +
                          angle * options_.rotation_delta_cost_weight();
      penalty[i][c] = std::exp(-(cost * cost));
    }, 4096);
    points[i].reserve(3 * q.point_cloud->size());
    for (const sensor::RangefinderPoint& point : *q.point_cloud)
      for (const float v : {point.position.x(), point.position.y(), point.position.z()}) points[i].push_back(v);
    csm_rt_job3d& job = jobs[i];
    job.grid = q.grid->get();
    job.points_xyz = points[i].data();
    job.num_points = static_cast<int32_t>(q.point_cloud->size());
    job.num_candidates = static_cast<int32_t>(k);
    job.candidate_poses = poses[i].data();
    job.penalty = penalty[i].data();
    job.resolution = q.grid->resolution();
    answers[i].num_candidates = static_cast<int>(k);
  }
  std::vector<csm_rt_result3d> results(n);
  CheckOk(csm_rt_match3d_batch(context->get(), jobs.data(), static_cast<int>(n), results.data()),
          "csm_rt_match3d_batch");
  for (size_t i = 0; i != n; ++i) {
    const float* p = poses[i].data() + 7 * results[i].best_index;
    answers[i].score = results[i].score;
    answers[i].best_index = results[i].best_index;
    answers[i].pose_estimate = transform::Rigid3d(transform::Vector3d(p[0], p[1], p[2]),
                                                  transform::Quaterniond(p[3], p[4], p[5], p[6]));
  }
  return answers;
}

}  // namespace scan_matching
}  // namespace mapping
}  // namespace cartographer
