#include "cartographer/mapping/internal/2d/scan_matching/correlative_scan_matcher_2d.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>

#include "cartographer/mapping/internal/2d/scan_matching/fast_correlative_scan_matcher_2d.h"

namespace cartographer {
namespace mapping {
namespace scan_matching {

// Reference: correlative_scan_matcher_2d.cc:27-55.  fp32 range (Eigen
// head<2>().norm()), fp32 square of the range (common::Pow2<float>), the rest fp64.
SearchParameters::SearchParameters(const double linear_search_window,
                                   const double angular_search_window,
                                   const sensor::PointCloud& point_cloud, const double resolution)
    : resolution(resolution) {
  float longest = 3.f * resolution;
  for (const sensor::RangefinderPoint& point : point_cloud) {
    const float px = point.position.x(), py = point.position.y();
    longest = std::max(std::sqrt(px * px + py * py), longest);
  }
  const float longest_squared = longest * (longest * 1.f);
  const double margin = 1. - 1e-3;
  angular_perturbation_step_size =
      margin * std::acos(1. - (resolution * (resolution * 1.)) / (2. * longest_squared));
  num_angular_perturbations =
      static_cast<int>(std::ceil(angular_search_window / angular_perturbation_step_size));
  num_scans = 2 * num_angular_perturbations + 1;
  const int reach = static_cast<int>(std::ceil(linear_search_window / resolution));
  linear_bounds.assign(num_scans, LinearBounds{-reach, reach, -reach, reach});
}

SearchParameters::SearchParameters(const int num_linear_perturbations,
                                   const int num_angular_perturbations,
                                   const double angular_perturbation_step_size,
                                   const double resolution)
    : num_angular_perturbations(num_angular_perturbations),
      angular_perturbation_step_size(angular_perturbation_step_size),
      resolution(resolution),
      num_scans(2 * num_angular_perturbations + 1) {
  linear_bounds.assign(num_scans, LinearBounds{-num_linear_perturbations, num_linear_perturbations,
                                               -num_linear_perturbations, num_linear_perturbations});
}

std::vector<float> SearchParameters::ScanQuaternionTable() const {
  std::vector<float> table(2 * static_cast<size_t>(num_scans));
  double delta_theta = -num_angular_perturbations * angular_perturbation_step_size;
  for (int scan = 0; scan < num_scans; ++scan, delta_theta += angular_perturbation_step_size) {
    // Eigen: Quaternionf(AngleAxisf(float(delta_theta), UnitZ)): w = cos(a/2), z = sin(a/2).
    const float half = 0.5f * static_cast<float>(delta_theta);
    table[2 * scan] = std::cos(half);
    table[2 * scan + 1] = std::sin(half);
  }
  return table;
}

namespace {
void CheckDevice(const int status, const char* what) {
  if (status != CSM_OK) {
    std::fprintf(stderr, "%s failed: %s\n", what, csm_last_error());
    std::abort();
  }
}
}  // namespace

std::vector<sensor::PointCloud> GenerateRotatedScans(const sensor::PointCloud& point_cloud,
                                                     const SearchParameters& search_parameters) {
  const size_t n = point_cloud.size();
  const int num_scans = search_parameters.num_scans;
  std::vector<sensor::PointCloud> rotated_scans(num_scans);
  if (n == 0 || num_scans == 0) return rotated_scans;
  std::vector<float> xyz(3 * n), out(3 * n * num_scans);
  for (size_t p = 0; p != n; ++p) {
    xyz[3 * p] = point_cloud[p].position.x();
    xyz[3 * p + 1] = point_cloud[p].position.y();
    xyz[3 * p + 2] = point_cloud[p].position.z();
  }
  const std::vector<float> quats = search_parameters.ScanQuaternionTable();
  CheckDevice(csm_rotate_scans2d(DeviceContext::ThreadDefault()->get(), xyz.data(), static_cast<int>(n),
                                 quats.data(), num_scans, out.data()),
              "csm_rotate_scans2d");
  for (int s = 0; s != num_scans; ++s)
    for (size_t p = 0; p != n; ++p) {
      const float* v = out.data() + 3 * (static_cast<size_t>(s) * n + p);
      rotated_scans[s].push_back({sensor::Vector3f(v[0], v[1], v[2])});
    }
  return rotated_scans;
}

std::vector<DiscreteScan2D> DiscretizeScans(const MapLimits& map_limits,
                                            const std::vector<sensor::PointCloud>& scans,
                                            const Eigen::Translation2f& initial_translation) {
  std::vector<DiscreteScan2D> discrete_scans(scans.size());
  size_t total = 0;
  for (const sensor::PointCloud& scan : scans) total += scan.size();
  if (total == 0) return discrete_scans;
  std::vector<float> xyz;
  xyz.reserve(3 * total);
  for (const sensor::PointCloud& scan : scans)
    for (const sensor::RangefinderPoint& point : scan) {
      xyz.push_back(point.position.x());
      xyz.push_back(point.position.y());
      xyz.push_back(point.position.z());
    }
  std::vector<int32_t> cells(2 * total);
  CheckDevice(csm_discretize_points2d(DeviceContext::ThreadDefault()->get(), xyz.data(),
                                      static_cast<long long>(total), initial_translation.x(),
                                      initial_translation.y(), map_limits.max().x(), map_limits.max().y(),
                                      map_limits.resolution(), cells.data()),
              "csm_discretize_points2d");
  size_t at = 0;
  for (size_t s = 0; s != scans.size(); ++s) {
    discrete_scans[s].reserve(scans[s].size());
    for (size_t p = 0; p != scans[s].size(); ++p, ++at)
      discrete_scans[s].emplace_back(cells[2 * at], cells[2 * at + 1]);
  }
  return discrete_scans;
}

}  // namespace scan_matching
}  // namespace mapping
}  // namespace cartographer
