// TEST INFRASTRUCTURE ONLY (see csm_oracle.h).  CPU restatement of
//   cartographer/mapping/internal/2d/tsdf_range_data_inserter_2d.cc:30-242
//   cartographer/mapping/internal/2d/normal_estimation_2d.cc:24-112
// Held bit for bit to the reference's own sources in oracle/_ref by tests/test_oracle_vs_ref.py.
// Float / double mix as in the reference: GaussianKernel promotes to double (its literals are
// doubles), everything else is float; std::sort with the reference's comparator.
#ifndef ORACLE_CSM_TSDF_INSERTER_H_
#define ORACLE_CSM_TSDF_INSERTER_H_

#include <algorithm>
#include <cmath>
#include <limits>
#include <vector>

#include "csm_oracle.h"

namespace oracle {

struct TSDFRangeDataInserterOptions2D {
  double truncation_distance = 0.3;
  double maximum_weight = 10.;
  bool update_free_space = false;
  int num_normal_samples = 4;     // normal_estimation_options
  float sample_radius = 0.5f;     // normal_estimation_options
  bool project_sdf_distance_to_scan_normal = true;
  int update_weight_range_exponent = 0;
  double update_weight_angle_scan_normal_to_ray_kernel_bandwidth = 0.5;
  double update_weight_distance_cell_to_hit_kernel_bandwidth = 0.5;
};

namespace tsdf_detail {

struct V2 {
  float x, y;
};
inline float Norm(const V2 v) { return std::sqrt(v.x * v.x + v.y * v.y); }
inline float NormV3(const Vec3f& v) { return std::sqrt((v.x * v.x + v.y * v.y) + v.z * v.z); }
// Eigen's normalized(): v / sqrt(squaredNorm) when squaredNorm > 0.
inline V2 Normalized(const V2 v) {
  const float z = v.x * v.x + v.y * v.y;
  if (z > 0.f) {
    const float n = std::sqrt(z);
    return V2{v.x / n, v.y / n};
  }
  return v;
}
inline Vec3f Normalized3(const Vec3f& v) {
  const float z = (v.x * v.x + v.y * v.y) + v.z * v.z;
  if (z > 0.f) {
    const float n = std::sqrt(z);
    return Vec3f{v.x / n, v.y / n, v.z / n};
  }
  return v;
}
// common/math.h:60-67 with T = float: the literal 2. promotes the update to double.
inline float NormalizeAngleDifference(float difference) {
  const float kPi = static_cast<float>(M_PI);
  while (difference > kPi) difference = static_cast<float>(difference - 2. * kPi);
  while (difference < -kPi) difference = static_cast<float>(difference + 2. * kPi);
  return difference;
}
// tsdf_range_data_inserter_2d.cc:31,51-53
inline float GaussianKernel(const float x, const float sigma) {
  static const float kSqrtTwoPi = static_cast<float>(std::sqrt(2.0 * M_PI));
  return static_cast<float>(1.0 / (kSqrtTwoPi * sigma) * std::exp(-0.5 * x * x / (sigma * sigma)));
}
// :98-104
inline float ComputeRangeWeightFactor(const float range, const int exponent) {
  constexpr float kMinRangeMeters = 1e-6f;
  float weight = 0.f;
  if (std::abs(range) > kMinRangeMeters) weight = static_cast<float>(1.f / (std::pow(range, exponent)));
  return weight;
}
// :71-90
struct RangeDataSorter {
  V2 origin;
  bool operator()(const Vec3f& lhs, const Vec3f& rhs) const {
    const V2 dl = Normalized(V2{lhs.x - origin.x, lhs.y - origin.y});
    const V2 dr = Normalized(V2{rhs.x - origin.x, rhs.y - origin.y});
    if ((dl.y < 0.f) != (dr.y < 0.f)) return dl.y < 0.f;
    if (dl.y < 0.f) return dl.x < dr.x;
    return dl.x > dr.x;
  }
};

// normal_estimation_2d.cc:24-66
inline float NormalTo2DAngle(const Vec3f& v) { return std::atan2(v.y, v.x); }
inline float EstimateNormal(const std::vector<Vec3f>& returns, const size_t estimation_point_index,
                            const size_t begin, const size_t end, const Vec3f& sensor_origin) {
  const Vec3f& p = returns[estimation_point_index];
  const Vec3f to_observation{sensor_origin.x - p.x, sensor_origin.y - p.y, sensor_origin.z - p.z};
  if (end - begin < 2) return NormalTo2DAngle(to_observation);
  Vec3f mean{0.f, 0.f, 0.f};
  for (size_t k = begin; k < end; ++k) {
    if (k == estimation_point_index) continue;
    const Vec3f& q = returns[k];
    const Vec3f tangent{p.x - q.x, p.y - q.y, p.z - q.z};
    Vec3f n{-tangent.y, tangent.x, 0.f};
    constexpr float kMinNormalLength = 1e-6f;
    if (NormV3(n) < kMinNormalLength) continue;
    if ((n.x * to_observation.x + n.y * to_observation.y) + n.z * to_observation.z < 0) n = Vec3f{-n.x, -n.y, -n.z};
    n = Normalized3(n);
    mean = Vec3f{mean.x + n.x, mean.y + n.y, mean.z + n.z};
  }
  return NormalTo2DAngle(mean);
}
// normal_estimation_2d.cc:81-110
inline std::vector<float> EstimateNormals(const Vec3f& origin, const std::vector<Vec3f>& returns,
                                          const int num_normal_samples, const float sample_radius) {
  std::vector<float> normals;
  const size_t max_num_samples = static_cast<size_t>(num_normal_samples);
  auto dist = [](const Vec3f& a, const Vec3f& b) { return NormV3(Vec3f{a.x - b.x, a.y - b.y, a.z - b.z}); };
  for (size_t current = 0; current < returns.size(); ++current) {
    const Vec3f& hit = returns[current];
    size_t begin = current;
    for (; begin > 0 && current - begin < max_num_samples / 2 && dist(hit, returns[begin - 1]) < sample_radius; --begin) {
    }
    size_t end = current;
    for (; end < returns.size() && end - current < std::ceil(max_num_samples / 2.0) + 1 &&
           dist(hit, returns[end]) < sample_radius;
         ++end) {
    }
    normals.push_back(EstimateNormal(returns, current, begin, end, origin));
  }
  return normals;
}

}  // namespace tsdf_detail

class TSDFRangeDataInserter2D {
 public:
  explicit TSDFRangeDataInserter2D(const TSDFRangeDataInserterOptions2D& options) : options_(options) {}

  // :131-165
  void Insert(const Vec3f& origin3, const std::vector<Vec3f>& returns_in, TSDF2D* tsdf) const {
    using namespace tsdf_detail;
    const float truncation_distance = static_cast<float>(options_.truncation_distance);
    GrowAsNeeded(origin3, returns_in, truncation_distance, tsdf);
    const bool scale_angle = options_.update_weight_angle_scan_normal_to_ray_kernel_bandwidth != 0.f;
    std::vector<Vec3f> returns = returns_in;
    std::vector<float> normals;
    if (options_.project_sdf_distance_to_scan_normal || scale_angle) {
      std::sort(returns.begin(), returns.end(), RangeDataSorter{V2{origin3.x, origin3.y}});
      normals = EstimateNormals(origin3, returns, options_.num_normal_samples, options_.sample_radius);
    }
    const V2 origin{origin3.x, origin3.y};
    for (size_t i = 0; i < returns.size(); ++i) {
      const float normal = normals.empty() ? std::numeric_limits<float>::quiet_NaN() : normals[i];
      InsertHit(V2{returns[i].x, returns[i].y}, origin, normal, tsdf);
    }
    tsdf->FinishUpdate();
  }

 private:
  // :33-48
  static void GrowAsNeeded(const Vec3f& origin, const std::vector<Vec3f>& returns, const float truncation_distance,
                           TSDF2D* tsdf) {
    using namespace tsdf_detail;
    float min_x = origin.x, max_x = origin.x, min_y = origin.y, max_y = origin.y;
    for (const Vec3f& hit : returns) {
      const Vec3f direction = Normalized3(Vec3f{hit.x - origin.x, hit.y - origin.y, hit.z - origin.z});
      const float ex = hit.x + truncation_distance * direction.x, ey = hit.y + truncation_distance * direction.y;
      min_x = std::min(min_x, ex);
      max_x = std::max(max_x, ex);
      min_y = std::min(min_y, ey);
      max_y = std::max(max_y, ey);
    }
    constexpr float kPadding = 1e-6f;
    tsdf->GrowLimits(min_x - kPadding * 1.f, min_y - kPadding * 1.f);
    tsdf->GrowLimits(max_x + kPadding * 1.f, max_y + kPadding * 1.f);
  }

  // :167-224
  void InsertHit(const tsdf_detail::V2 hit, const tsdf_detail::V2 origin, const float normal, TSDF2D* tsdf) const {
    using namespace tsdf_detail;
    constexpr int kSubpixelScale = 1000;
    const V2 ray{hit.x - origin.x, hit.y - origin.y};
    const float range = Norm(ray);
    const float truncation_distance = static_cast<float>(options_.truncation_distance);
    if (range < truncation_distance) return;
    const float truncation_ratio = truncation_distance / range;
    const V2 ray_begin = options_.update_free_space
                             ? origin
                             : V2{origin.x + (1.0f - truncation_ratio) * ray.x, origin.y + (1.0f - truncation_ratio) * ray.y};
    const V2 ray_end{origin.x + (1.0f + truncation_ratio) * ray.x, origin.y + (1.0f + truncation_ratio) * ray.y};
    const MapLimits& limits = tsdf->limits();
    MapLimits super;  // SuperscaleRay, :55-69
    super.resolution = limits.resolution / kSubpixelScale;
    super.max_x = limits.max_x;
    super.max_y = limits.max_y;
    super.cell_limits.num_x_cells = limits.cell_limits.num_x_cells * kSubpixelScale;
    super.cell_limits.num_y_cells = limits.cell_limits.num_y_cells * kSubpixelScale;
    const std::vector<Array2i> ray_mask = RayToPixelMask(super.GetCellIndex(ray_begin.x, ray_begin.y),
                                                         super.GetCellIndex(ray_end.x, ray_end.y), kSubpixelScale);
    float weight_factor_angle_ray_normal = 1.f;
    if (options_.update_weight_angle_scan_normal_to_ray_kernel_bandwidth != 0.f) {
      const float angle_ray_normal = NormalizeAngleDifference(normal - std::atan2(-ray.y, -ray.x));
      weight_factor_angle_ray_normal = GaussianKernel(
          angle_ray_normal, static_cast<float>(options_.update_weight_angle_scan_normal_to_ray_kernel_bandwidth));
    }
    float weight_factor_range = 1.f;
    if (options_.update_weight_range_exponent != 0) {
      weight_factor_range = ComputeRangeWeightFactor(range, options_.update_weight_range_exponent);
    }
    for (const Array2i& cell : ray_mask) {
      if (tsdf->CellIsUpdated(cell)) continue;
      // MapLimits::GetCellCenter (map_limits.h:79-83): doubles narrowed to float
      const V2 center{static_cast<float>(limits.max_x - limits.resolution * (cell.y + 0.5)),
                      static_cast<float>(limits.max_y - limits.resolution * (cell.x + 0.5))};
      const float distance_cell_to_origin = Norm(V2{center.x - origin.x, center.y - origin.y});
      float update_tsd = range - distance_cell_to_origin;
      if (options_.project_sdf_distance_to_scan_normal) {
        update_tsd = (center.x - hit.x) * std::cos(normal) + (center.y - hit.y) * std::sin(normal);
      }
      update_tsd = Clamp(update_tsd, -truncation_distance, truncation_distance);
      float update_weight = weight_factor_range * weight_factor_angle_ray_normal;
      if (options_.update_weight_distance_cell_to_hit_kernel_bandwidth != 0.f) {
        update_weight *= GaussianKernel(
            update_tsd, static_cast<float>(options_.update_weight_distance_cell_to_hit_kernel_bandwidth));
      }
      UpdateCell(cell, update_tsd, update_weight, tsdf);
    }
  }

  // :226-239
  void UpdateCell(const Array2i& cell, const float update_sdf, const float update_weight, TSDF2D* tsdf) const {
    if (update_weight == 0.f) return;
    const std::pair<float, float> tw = tsdf->GetTSDAndWeight(cell);
    float updated_weight = tw.second + update_weight;
    const float updated_sdf = (tw.first * tw.second + update_sdf * update_weight) / updated_weight;
    updated_weight = std::min(updated_weight, static_cast<float>(options_.maximum_weight));
    tsdf->SetCellMarked(cell, updated_sdf, updated_weight);
  }

  TSDFRangeDataInserterOptions2D options_;
};

}  // namespace oracle

#endif  // ORACLE_CSM_TSDF_INSERTER_H_
