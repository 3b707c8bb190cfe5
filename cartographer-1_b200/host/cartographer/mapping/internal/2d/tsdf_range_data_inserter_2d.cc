#include "cartographer/mapping/internal/2d/tsdf_range_data_inserter_2d.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <limits>

namespace cartographer {
namespace mapping {
namespace {

void CheckOk(int status, const char* what) {
  if (status != CSM_OK) {
    std::fprintf(stderr, "%s failed: %s\n", what, csm_last_error());
    std::abort();
  }
}

// The float vector arithmetic of the reference's Eigen expressions, in Eigen's operation order.
struct V2 {
  float x, y;
};
struct V3 {
  float x, y, z;
};
V3 Sub(const sensor::Vector3f& a, const sensor::Vector3f& b) { return V3{a.x() - b.x(), a.y() - b.y(), a.z() - b.z()}; }
float Norm(const V3& v) { return std::sqrt((v.x * v.x + v.y * v.y) + v.z * v.z); }
V3 Normalized(const V3& v) {  // MatrixBase::normalized(): v / sqrt(squaredNorm) when squaredNorm > 0
  const float z = (v.x * v.x + v.y * v.y) + v.z * v.z;
  if (z > 0.f) {
    const float n = std::sqrt(z);
    return V3{v.x / n, v.y / n, v.z / n};
  }
  return v;
}
V2 Normalized(const V2& v) {
  const float z = v.x * v.x + v.y * v.y;
  if (z > 0.f) {
    const float n = std::sqrt(z);
    return V2{v.x / n, v.y / n};
  }
  return v;
}

constexpr float kMinRangeMeters = 1e-6f;  // tsdf_range_data_inserter_2d.cc:30

// :51-53 -- float in, double inside (the literals), float out.
float GaussianKernel(const float x, const float sigma) {
  static const float kSqrtTwoPi = static_cast<float>(std::sqrt(2.0 * M_PI));
  return static_cast<float>(1.0 / (kSqrtTwoPi * sigma) * std::exp(-0.5 * x * x / (sigma * sigma)));
}
// :98-104
float ComputeRangeWeightFactor(const float range, const int exponent) {
  float weight = 0.f;
  if (std::abs(range) > kMinRangeMeters) weight = static_cast<float>(1.f / (std::pow(range, exponent)));
  return weight;
}
// common/math.h:60-67 with T = float (the literal 2. makes the update a double expression)
float NormalizeAngleDifference(float difference) {
  const float kPi = static_cast<float>(M_PI);
  while (difference > kPi) difference = static_cast<float>(difference - 2. * kPi);
  while (difference < -kPi) difference = static_cast<float>(difference + 2. * kPi);
  return difference;
}
// :71-90
struct RangeDataSorter {
  V2 origin;
  bool operator()(const sensor::RangefinderPoint& lhs, const sensor::RangefinderPoint& rhs) const {
    const V2 dl = Normalized(V2{lhs.position.x() - origin.x, lhs.position.y() - origin.y});
    const V2 dr = Normalized(V2{rhs.position.x() - origin.x, rhs.position.y() - origin.y});
    if ((dl.y < 0.f) != (dr.y < 0.f)) return dl.y < 0.f;
    if (dl.y < 0.f) return dl.x < dr.x;
    return dl.x > dr.x;
  }
};

// normal_estimation_2d.cc:24-66
float NormalTo2DAngle(const V3& v) { return std::atan2(v.y, v.x); }
float EstimateNormal(const sensor::PointCloud& returns, const size_t estimation_point_index, const size_t begin,
                     const size_t end, const sensor::Vector3f& sensor_origin) {
  const sensor::Vector3f& p = returns[estimation_point_index].position;
  const V3 to_observation = Sub(sensor_origin, p);
  if (end - begin < 2) return NormalTo2DAngle(to_observation);
  V3 mean{0.f, 0.f, 0.f};
  for (size_t k = begin; k < end; ++k) {
    if (k == estimation_point_index) continue;
    const V3 tangent = Sub(p, returns[k].position);
    V3 n{-tangent.y, tangent.x, 0.f};
    constexpr float kMinNormalLength = 1e-6f;
    if (Norm(n) < kMinNormalLength) continue;
    if ((n.x * to_observation.x + n.y * to_observation.y) + n.z * to_observation.z < 0) n = V3{-n.x, -n.y, -n.z};
    n = Normalized(n);
    mean = V3{mean.x + n.x, mean.y + n.y, mean.z + n.z};
  }
  return NormalTo2DAngle(mean);
}

}  // namespace

proto::NormalEstimationOptions2D CreateNormalEstimationOptions2D(
    common::LuaParameterDictionary* const parameter_dictionary) {
  proto::NormalEstimationOptions2D options;
  options.set_num_normal_samples(parameter_dictionary->GetInt("num_normal_samples"));
  options.set_sample_radius(static_cast<float>(parameter_dictionary->GetDouble("sample_radius")));
  CSM_CHECK(options.num_normal_samples() > 0);
  CSM_CHECK(options.sample_radius() > 0.0);
  return options;
}

// The stand-in dictionary is one flat table (common/lua_parameter_dictionary.h): the keys of the
// reference's nested `normal_estimation_options` table are read from the same table.
proto::TSDFRangeDataInserterOptions2D CreateTSDFRangeDataInserterOptions2D(
    common::LuaParameterDictionary* const parameter_dictionary) {
  proto::TSDFRangeDataInserterOptions2D options;
  options.set_truncation_distance(parameter_dictionary->GetDouble("truncation_distance"));
  options.set_maximum_weight(parameter_dictionary->GetDouble("maximum_weight"));
  options.set_update_free_space(parameter_dictionary->GetBool("update_free_space"));
  *options.mutable_normal_estimation_options() = CreateNormalEstimationOptions2D(parameter_dictionary);
  options.set_project_sdf_distance_to_scan_normal(parameter_dictionary->GetBool("project_sdf_distance_to_scan_normal"));
  options.set_update_weight_range_exponent(parameter_dictionary->GetInt("update_weight_range_exponent"));
  options.set_update_weight_angle_scan_normal_to_ray_kernel_bandwidth(
      parameter_dictionary->GetDouble("update_weight_angle_scan_normal_to_ray_kernel_bandwidth"));
  options.set_update_weight_distance_cell_to_hit_kernel_bandwidth(
      parameter_dictionary->GetDouble("update_weight_distance_cell_to_hit_kernel_bandwidth"));
  return options;
}

std::vector<float> EstimateNormals(const sensor::RangeData& range_data,
                                   const proto::NormalEstimationOptions2D& normal_estimation_options) {
  std::vector<float> normals;
  normals.reserve(range_data.returns.size());
  const size_t max_num_samples = normal_estimation_options.num_normal_samples();
  const float sample_radius = normal_estimation_options.sample_radius();
  for (size_t current = 0; current < range_data.returns.size(); ++current) {
    const sensor::Vector3f& hit = range_data.returns[current].position;
    size_t begin = current;
    for (; begin > 0 && current - begin < max_num_samples / 2 &&
           Norm(Sub(hit, range_data.returns[begin - 1].position)) < sample_radius;
         --begin) {
    }
    size_t end = current;
    for (; end < range_data.returns.size() && end - current < std::ceil(max_num_samples / 2.0) + 1 &&
           Norm(Sub(hit, range_data.returns[end].position)) < sample_radius;
         ++end) {
    }
    normals.push_back(EstimateNormal(range_data.returns, current, begin, end, range_data.origin));
  }
  return normals;
}

TSDFRangeDataInserter2D::TSDFRangeDataInserter2D(const proto::TSDFRangeDataInserterOptions2D& options)
    : TSDFRangeDataInserter2D(options, nullptr) {}

TSDFRangeDataInserter2D::TSDFRangeDataInserter2D(const proto::TSDFRangeDataInserterOptions2D& options,
                                                 DeviceContext* context)
    : options_(options), context_(context) {}

void TSDFRangeDataInserter2D::GrowAsNeeded(const sensor::RangeData& range_data, const float truncation_distance,
                                           TSDF2D* const tsdf) {
  float min_x = range_data.origin.x(), max_x = min_x, min_y = range_data.origin.y(), max_y = min_y;
  for (const sensor::RangefinderPoint& hit : range_data.returns) {
    const V3 direction = Normalized(Sub(hit.position, range_data.origin));
    const float ex = hit.position.x() + truncation_distance * direction.x;
    const float ey = hit.position.y() + truncation_distance * direction.y;
    min_x = std::min(min_x, ex);
    max_x = std::max(max_x, ex);
    min_y = std::min(min_y, ey);
    max_y = std::max(max_y, ey);
  }
  constexpr float kPadding = 1e-6f;  // avoids numerical issues at cell boundaries
  tsdf->GrowLimits(min_x - kPadding * 1.f, min_y - kPadding * 1.f);
  tsdf->GrowLimits(max_x + kPadding * 1.f, max_y + kPadding * 1.f);
}

void TSDFRangeDataInserter2D::Insert(const sensor::RangeData& range_data, Grid2D* const grid) const {
  CSM_CHECK(grid != nullptr && grid->GetGridType() == GridType::TSDF);
  TSDF2D* const tsdf = static_cast<TSDF2D*>(grid);
  GrowAsNeeded(range_data, static_cast<float>(options_.truncation_distance()), tsdf);
  DeviceContext* ctx = context_ != nullptr ? context_ : DeviceContext::ThreadDefault();
  DeviceGrid device_grid(ctx, *tsdf);
  Insert({&range_data}, *tsdf, &device_grid);
  device_grid.Download(tsdf);
}

void TSDFRangeDataInserter2D::Insert(const std::vector<const sensor::RangeData*>& range_data, const TSDF2D& layout,
                                     DeviceGrid* const grid) const {
  if (range_data.empty()) return;
  CSM_CHECK(grid != nullptr);
  DeviceContext* ctx = context_ != nullptr ? context_ : DeviceContext::ThreadDefault();
  const bool scale_angle = options_.update_weight_angle_scan_normal_to_ray_kernel_bandwidth() != 0.f;
  const bool need_normals = options_.project_sdf_distance_to_scan_normal() || scale_angle;
  const TSDValueConverter& converter = layout.value_converter();
  struct Packed {
    std::vector<float> hits, normal_cs, weight;
  };
  std::vector<Packed> packed(range_data.size());
  std::vector<csm_tsdf_range_data2d> scans(range_data.size());
  for (size_t s = 0; s != range_data.size(); ++s) {
    // :139-151: sort by bearing and estimate the normals only when something uses them.
    sensor::RangeData sorted = *range_data[s];
    std::vector<float> normals;
    if (need_normals) {
      std::vector<sensor::RangefinderPoint> returns = sorted.returns.points();
      std::sort(returns.begin(), returns.end(), RangeDataSorter{V2{sorted.origin.x(), sorted.origin.y()}});
      sorted.returns = sensor::PointCloud(std::move(returns));
      normals = EstimateNormals(sorted, options_.normal_estimation_options());
    }
    Packed& p = packed[s];
    const size_t n = sorted.returns.size();
    p.hits.resize(2 * n);
    p.normal_cs.resize(2 * n);
    p.weight.resize(n);
    for (size_t i = 0; i != n; ++i) {
      const sensor::Vector3f& hit = sorted.returns[i].position;
      p.hits[2 * i] = hit.x();
      p.hits[2 * i + 1] = hit.y();
      const float normal = normals.empty() ? std::numeric_limits<float>::quiet_NaN() : normals[i];
      p.normal_cs[2 * i] = std::cos(normal);      // Eigen::Vector2f{std::cos(normal_orientation),
      p.normal_cs[2 * i + 1] = std::sin(normal);  //                 std::sin(normal_orientation)}, :213-215
      // :171-173, 189-203: per-return weight factors
      const float rx = hit.x() - sorted.origin.x(), ry = hit.y() - sorted.origin.y();
      const float range = std::sqrt(rx * rx + ry * ry);
      float weight_factor_angle_ray_normal = 1.f;
      if (scale_angle) {
        const float angle_ray_normal = NormalizeAngleDifference(normal - std::atan2(-ry, -rx));
        weight_factor_angle_ray_normal = GaussianKernel(
            angle_ray_normal, static_cast<float>(options_.update_weight_angle_scan_normal_to_ray_kernel_bandwidth()));
      }
      float weight_factor_range = 1.f;
      if (options_.update_weight_range_exponent() != 0) {
        weight_factor_range = ComputeRangeWeightFactor(range, options_.update_weight_range_exponent());
      }
      p.weight[i] = weight_factor_range * weight_factor_angle_ray_normal;
    }
    csm_tsdf_range_data2d& out = scans[s];
    out.max_x = grid->limits().max().x();
    out.max_y = grid->limits().max().y();
    out.resolution = grid->limits().resolution();
    out.origin_x = sorted.origin.x();
    out.origin_y = sorted.origin.y();
    out.num_returns = static_cast<int32_t>(n);
    out.update_free_space = options_.update_free_space() ? 1 : 0;
    out.hits_xy = p.hits.data();
    out.normal_cos_sin = p.normal_cs.data();
    out.hit_weight = p.weight.data();
    out.truncation_distance = static_cast<float>(options_.truncation_distance());
    out.maximum_weight = static_cast<float>(options_.maximum_weight());
    out.distance_kernel_bandwidth = static_cast<float>(options_.update_weight_distance_cell_to_hit_kernel_bandwidth());
    out.project_sdf_distance_to_scan_normal = options_.project_sdf_distance_to_scan_normal() ? 1 : 0;
    out.min_tsd = converter.getMinTSD();
    out.max_tsd = converter.getMaxTSD();
    out.tsd_resolution = 32766.f / (converter.getMaxTSD() - converter.getMinTSD());
    out.max_weight_value = converter.getMaxWeight();
    out.weight_resolution = 32766.f / (converter.getMaxWeight() - converter.getMinWeight());
    out.pad = 0;
  }
  CheckOk(csm_tsdfgrid2d_insert(ctx->get(), grid->get(), scans.data(), static_cast<int>(scans.size())),
          "csm_tsdfgrid2d_insert");
}

TSDFRangeDataInserter2D::DeviceGrid::DeviceGrid(DeviceContext* const context, const TSDF2D& tsdf)
    : context_(context), limits_(tsdf.limits()) {
  const CellLimits& cells = limits_.cell_limits();
  CheckOk(csm_tsdfgrid2d_create(context_->get(), tsdf.correspondence_cost_cells().data(), tsdf.weight_cells().data(),
                                cells.num_x_cells, cells.num_y_cells, tsdf.value_converter().value_to_tsd_table().data(),
                                tsdf.value_converter().value_to_weight_table().data(), &grid_),
          "csm_tsdfgrid2d_create");
}

TSDFRangeDataInserter2D::DeviceGrid::~DeviceGrid() { csm_tsdfgrid2d_destroy(context_->get(), grid_); }

void TSDFRangeDataInserter2D::DeviceGrid::Download(TSDF2D* const tsdf) const {
  CSM_CHECK(tsdf != nullptr);
  const CellLimits &a = limits_.cell_limits(), &b = tsdf->limits().cell_limits();
  CSM_CHECK(a.num_x_cells == b.num_x_cells && a.num_y_cells == b.num_y_cells);
  CheckOk(csm_tsdfgrid2d_download(context_->get(), grid_, tsdf->mutable_correspondence_cost_cells()->data(),
                                  tsdf->mutable_weight_cells()->data()),
          "csm_tsdfgrid2d_download");
}

}  // namespace mapping
}  // namespace cartographer
