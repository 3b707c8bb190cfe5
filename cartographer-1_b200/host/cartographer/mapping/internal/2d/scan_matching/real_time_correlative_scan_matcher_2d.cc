#include "cartographer/mapping/internal/2d/scan_matching/real_time_correlative_scan_matcher_2d.h"

#include "parallel_for.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace cartographer {
namespace mapping {
namespace scan_matching {
namespace {

// sensor::TransformPointCloud with Rigid3f::Rotation(AngleAxisf(angle, UnitZ))
// (reference: sensor/point_cloud.cc:56-64, transform/rigid_transform.h:129-131,
// 192-196): Eigen's quaternion * vector, fp32, no FMA (this file is compiled with
// -ffp-contract=off), then "+ translation" with a zero translation.  Only the x/y
// norms of the result are used here (SearchParameters); the device repeats the same
// rotation for the scoring itself.
sensor::PointCloud RotateYaw(const sensor::PointCloud& cloud, const float angle) {
  const float half = 0.5f * angle;
  const float w = std::cos(half), s = std::sin(half);
  sensor::PointCloud out;
  for (const sensor::RangefinderPoint& p : cloud) {
    const float vx = p.position.x(), vy = p.position.y(), vz = p.position.z();
    // q.vec = (0, 0, s): uv = 2 * (q.vec x v), c = q.vec x uv
    float uvx = 0.f * vz - s * vy, uvy = s * vx - 0.f * vz, uvz = 0.f * vy - 0.f * vx;
    uvx += uvx;
    uvy += uvy;
    uvz += uvz;
    const float cx = 0.f * uvz - s * uvy, cy = s * uvx - 0.f * uvz, cz = 0.f * uvy - 0.f * uvx;
    const float rx = (vx + w * uvx) + cx, ry = (vy + w * uvy) + cy, rz = (vz + w * uvz) + cz;
    out.push_back({sensor::Vector3f(rx + 0.f, ry + 0.f, rz + 0.f)});
  }
  return out;
}

void CheckOk(int status, const char* what) {
  if (status != CSM_OK) {
    std::fprintf(stderr, "%s failed: %s\n", what, csm_last_error());
    std::abort();
  }
}

}  // namespace

RealTimeCorrelativeScanMatcher2D::RealTimeCorrelativeScanMatcher2D(
    const proto::RealTimeCorrelativeScanMatcherOptions& options)
    : options_(options) {}

RealTimeCorrelativeScanMatcher2D::RealTimeCorrelativeScanMatcher2D(
    const proto::RealTimeCorrelativeScanMatcherOptions& options, DeviceContext* context)
    : options_(options), context_(context) {}

std::vector<Candidate2D> RealTimeCorrelativeScanMatcher2D::GenerateExhaustiveSearchCandidates(
    const SearchParameters& search_parameters) const {
  std::vector<Candidate2D> candidates;
  for (int scan_index = 0; scan_index != search_parameters.num_scans; ++scan_index) {
    const SearchParameters::LinearBounds& b = search_parameters.linear_bounds[scan_index];
    for (int x = b.min_x; x <= b.max_x; ++x)
      for (int y = b.min_y; y <= b.max_y; ++y)
        candidates.emplace_back(scan_index, x, y, search_parameters);
  }
  return candidates;
}

void RealTimeCorrelativeScanMatcher2D::ScoreCandidates(const Grid2D& grid,
                                                       const std::vector<DiscreteScan2D>& discrete_scans,
                                                       const SearchParameters& search_parameters,
                                                       std::vector<Candidate2D>* const candidates) const {
  CSM_CHECK(candidates != nullptr);
  if (candidates->empty()) return;
  CSM_CHECK(!discrete_scans.empty());
  const size_t n = discrete_scans[0].size();
  CSM_CHECK(n > 0);
  std::vector<int32_t> xy(2 * n * discrete_scans.size());
  for (size_t s = 0; s != discrete_scans.size(); ++s) {
    CSM_CHECK(discrete_scans[s].size() == n);  // rotated copies of one cloud
    for (size_t p = 0; p != n; ++p) {
      xy[2 * (s * n + p)] = discrete_scans[s][p].x();
      xy[2 * (s * n + p) + 1] = discrete_scans[s][p].y();
    }
  }
  std::vector<int32_t> list(3 * candidates->size());
  for (size_t c = 0; c != candidates->size(); ++c) {
    list[3 * c] = (*candidates)[c].scan_index;
    list[3 * c + 1] = (*candidates)[c].x_index_offset;
    list[3 * c + 2] = (*candidates)[c].y_index_offset;
  }
  const MapLimits& limits = grid.limits();
  csm_rt_job2d job;
  std::memset(&job, 0, sizeof(job));
  job.cells = grid.correspondence_cost_cells().data();
  job.num_x = limits.cell_limits().num_x_cells;
  job.num_y = limits.cell_limits().num_y_cells;
  job.cost_table = grid.value_to_correspondence_cost_table().data();
  job.num_points = static_cast<int32_t>(n);
  job.max_x = limits.max().x();
  job.max_y = limits.max().y();
  job.resolution = limits.resolution();
  job.out_of_bounds_probability = kMinProbability;
  if (grid.GetGridType() == GridType::TSDF) {
    const TSDF2D& tsdf = static_cast<const TSDF2D&>(grid);
    job.cost_table = tsdf.value_converter().value_to_tsd_table().data();
    job.weight_cells = tsdf.weight_cells().data();
    job.weight_table = tsdf.value_converter().value_to_weight_table().data();
    job.max_correspondence_cost = tsdf.GetMaxCorrespondenceCost();
    job.out_of_bounds_tsd = tsdf.value_converter().getMinTSD();
  }
  std::vector<float> scores(candidates->size());
  DeviceContext* ctx = context_ != nullptr ? context_ : DeviceContext::ThreadDefault();
  CheckOk(csm_rt_score_candidates2d(ctx->get(), &job, xy.data(), static_cast<int>(discrete_scans.size()),
                                    list.data(), static_cast<int>(list.size() / 3), scores.data()),
          "csm_rt_score_candidates2d");
  (void)search_parameters;  // the reference does not read it either
  for (size_t c = 0; c != candidates->size(); ++c) {
    Candidate2D& candidate = (*candidates)[c];
    candidate.score = scores[c];
    // reference :170-174: float *= double, evaluated in double and narrowed
    const double t = std::hypot(candidate.x, candidate.y) * options_.translation_delta_cost_weight() +
                     std::abs(candidate.orientation) * options_.rotation_delta_cost_weight();
    candidate.score = static_cast<float>(candidate.score * std::exp(-(t * t)));
  }
}

double RealTimeCorrelativeScanMatcher2D::Match(const transform::Rigid2d& initial_pose_estimate,
                                               const sensor::PointCloud& point_cloud,
                                               const Grid2D& grid,
                                               transform::Rigid2d* pose_estimate) const {
  CSM_CHECK(pose_estimate != nullptr);
  Query q;
  q.initial_pose_estimate = initial_pose_estimate;
  q.point_cloud = &point_cloud;
  q.grid = &grid;
  const Answer a = MatchBatch({q})[0];
  *pose_estimate = a.pose_estimate;
  return a.score;
}

std::vector<RealTimeCorrelativeScanMatcher2D::Answer> RealTimeCorrelativeScanMatcher2D::MatchBatch(
    const std::vector<Query>& queries) const {
  const size_t n = queries.size();
  std::vector<Answer> answers(n);
  if (n == 0) return answers;
  DeviceContext* ctx = context_ != nullptr ? context_ : DeviceContext::ThreadDefault();
  std::vector<SearchParameters> params(n, SearchParameters(0, 0, 0., 0.));
  std::vector<std::vector<float>> xyz(n), quats(n);
  std::vector<std::vector<double>> penalty(n);
  std::vector<csm_rt_job2d> jobs(n);
  // Host side of every scan (search parameters, angle and penalty tables: libm): independent, so
  // large batches spread it over a few threads (at batch 1024 it was the whole call's time).
  common::ParallelFor(n, [&](const size_t i) {
    const Query& q = queries[i];
    CSM_CHECK(q.point_cloud != nullptr && q.grid != nullptr);
    CSM_CHECK(!q.point_cloud->empty());
    const MapLimits& limits = q.grid->limits();
    const float yaw = static_cast<float>(q.initial_pose_estimate.rotation().angle());
    // The search parameters come from the PRE-ROTATED cloud (reference :124-130).
    params[i] = SearchParameters(options_.linear_search_window(), options_.angular_search_window(),
                                 RotateYaw(*q.point_cloud, yaw), limits.resolution());
    const SearchParameters& sp = params[i];
    xyz[i].resize(3 * q.point_cloud->size());
    for (size_t p = 0; p != q.point_cloud->size(); ++p) {
      const sensor::Vector3f& v = (*q.point_cloud)[p].position;
      xyz[i][3 * p] = v.x();
      xyz[i][3 * p + 1] = v.y();
      xyz[i][3 * p + 2] = v.z();
    }
    quats[i] = sp.ScanQuaternionTable();
    // Motion penalty per candidate, generation order scan -> x -> y (reference :100-112,
    // :170-174), fp64: exp(-(hypot(x, y) * w_t + |orientation| * w_r)^2) with
    // x = -y_offset * resolution, y = -x_offset * resolution and orientation =
    // (scan - num_angular_perturbations) * step.  The value depends only on
    // (|x_offset|, |y_offset|, |scan - nap|), so it is evaluated once per distinct
    // triple with exactly the reference's operations and copied to the mirror images.
    {
      const int lin = sp.linear_bounds[0].max_x;
      const int nap = sp.num_angular_perturbations;
      std::vector<double> reach(static_cast<size_t>(lin + 1) * (lin + 1));
      for (int ax = 0; ax <= lin; ++ax)
        for (int ay = 0; ay <= lin; ++ay) {
          const Candidate2D c(nap, ax, ay, sp);
          reach[ax * (lin + 1) + ay] = std::hypot(c.x, c.y) * options_.translation_delta_cost_weight();
        }
      // (|scan - nap|, |x|, |y|) -> factor; the device indexes it per candidate (csm_rt_job2d::penalty_table).
      std::vector<double>& table = penalty[i];
      table.resize(static_cast<size_t>(nap + 1) * reach.size());
      for (int a = 0; a <= nap; ++a) {
        const Candidate2D c(nap + a, 0, 0, sp);
        const double turn = std::abs(c.orientation) * options_.rotation_delta_cost_weight();
        for (size_t k = 0; k != reach.size(); ++k) {
          const double t = reach[k] + turn;
          table[a * reach.size() + k] = std::exp(-(t * t));
        }
      }
    }
    csm_rt_job2d& job = jobs[i];
    std::memset(&job, 0, sizeof(job));
    job.cells = q.grid->correspondence_cost_cells().data();
    job.num_x = limits.cell_limits().num_x_cells;
    job.num_y = limits.cell_limits().num_y_cells;
    job.cost_table = q.grid->value_to_correspondence_cost_table().data();
    job.points_xyz = xyz[i].data();
    job.num_points = static_cast<int32_t>(q.point_cloud->size());
    job.num_scans = sp.num_scans;
    job.scan_quat = quats[i].data();
    const float half_yaw = 0.5f * yaw;
    job.init_quat_w = std::cos(half_yaw);
    job.init_quat_s = std::sin(half_yaw);
    job.init_tx = static_cast<float>(q.initial_pose_estimate.translation().x());
    job.init_ty = static_cast<float>(q.initial_pose_estimate.translation().y());
    job.max_x = limits.max().x();
    job.max_y = limits.max().y();
    job.resolution = limits.resolution();
    job.linear_bound = sp.linear_bounds[0].max_x;
    job.penalty = nullptr;
    job.penalty_table = penalty[i].data();
    job.num_angular_perturbations = sp.num_angular_perturbations;
    job.out_of_bounds_probability = kMinProbability;  // probability_grid.cc:78-82
    if (q.grid->GetGridType() == GridType::TSDF) {  // reference :38-59, :163-168
      const TSDF2D& tsdf = static_cast<const TSDF2D&>(*q.grid);
      job.cost_table = tsdf.value_converter().value_to_tsd_table().data();
      job.weight_cells = tsdf.weight_cells().data();
      job.weight_table = tsdf.value_converter().value_to_weight_table().data();
      job.max_correspondence_cost = tsdf.GetMaxCorrespondenceCost();
      job.out_of_bounds_tsd = tsdf.value_converter().getMinTSD();  // tsdf_2d.cc:96-97
    }
  });
  std::vector<csm_result2d> results(n);
  CheckOk(csm_rt_match2d_batch(ctx->get(), jobs.data(), static_cast<int>(n), results.data()),
          "csm_rt_match2d_batch");
  for (size_t i = 0; i != n; ++i) {
    const csm_result2d& r = results[i];
    Answer& a = answers[i];
    a.best_candidate = Candidate2D(r.scan_index, r.x_index_offset, r.y_index_offset, params[i]);
    a.best_candidate.score = r.score;
    a.score = r.score;  // float widened to double, as the reference returns it
    const transform::Rigid2d& init = queries[i].initial_pose_estimate;
    a.pose_estimate = transform::Rigid2d(
        transform::Vector2d(init.translation().x() + a.best_candidate.x,
                            init.translation().y() + a.best_candidate.y),
        init.rotation() * transform::Rotation2Dd(a.best_candidate.orientation));
  }
  return answers;
}

}  // namespace scan_matching
}  // namespace mapping
}  // namespace cartographer
