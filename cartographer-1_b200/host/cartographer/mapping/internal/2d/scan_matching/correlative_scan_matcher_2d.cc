#include "cartographer/mapping/internal/2d/scan_matching/correlative_scan_matcher_2d.h"

#include <algorithm>
#include <cmath>

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

}  // namespace scan_matching
}  // namespace mapping
}  // namespace cartographer
