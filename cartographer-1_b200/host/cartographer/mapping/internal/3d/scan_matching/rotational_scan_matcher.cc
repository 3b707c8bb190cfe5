#include "cartographer/mapping/internal/3d/scan_matching/rotational_scan_matcher.h"

#include <cmath>

namespace cartographer {
namespace mapping {
namespace scan_matching {
namespace {

// Eigen's norm() / dot() restated as scalar sums in index order
// (rotational_scan_matcher.cc:121-133).
float Norm(const Eigen::VectorXf& h) {
  float sum = 0.f;
  for (const float v : h) sum += v * v;
  return std::sqrt(sum);
}

float MatchHistograms(const Eigen::VectorXf& submap_histogram, const Eigen::VectorXf& scan_histogram) {
  const float scan_histogram_norm = Norm(scan_histogram);
  const float submap_histogram_norm = Norm(submap_histogram);
  const float normalization = scan_histogram_norm * submap_histogram_norm;
  if (normalization < 1e-3f) return 1.f;
  float dot = 0.f;
  for (size_t i = 0; i != submap_histogram.size(); ++i) dot += submap_histogram[i] * scan_histogram[i];
  return dot / normalization;
}

}  // namespace

RotationalScanMatcher::RotationalScanMatcher(const Eigen::VectorXf* histogram) : histogram_(histogram) {}

// rotational_scan_matcher.cc:141-162
Eigen::VectorXf RotationalScanMatcher::RotateHistogram(const Eigen::VectorXf& histogram, const float angle) {
  if (histogram.empty()) return histogram;
  const int size = static_cast<int>(histogram.size());
  // float * Index -> float; / M_PI -> double; narrowed on assignment.
  const float rotate_by_buckets = static_cast<float>((-angle * static_cast<float>(size)) / M_PI);
  int full_buckets = common::RoundToInt(rotate_by_buckets - 0.5f);
  const float fraction = rotate_by_buckets - full_buckets;
  while (full_buckets < 0) full_buckets += size;
  Eigen::VectorXf rotated(size);
  for (int i = 0; i != size; ++i) {
    const float h0 = histogram[(i + full_buckets) % size];
    const float h1 = histogram[(i + 1 + full_buckets) % size];
    rotated[i] = fraction * h1 + (1.f - fraction) * h0;
  }
  return rotated;
}

// rotational_scan_matcher.cc:178-189
std::vector<float> RotationalScanMatcher::Match(const Eigen::VectorXf& histogram, const float initial_angle,
                                                const std::vector<float>& angles) const {
  std::vector<float> result;
  result.reserve(angles.size());
  for (const float angle : angles)
    result.push_back(MatchHistograms(*histogram_, RotateHistogram(histogram, initial_angle + angle)));
  return result;
}

}  // namespace scan_matching
}  // namespace mapping
}  // namespace cartographer
