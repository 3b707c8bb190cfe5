#include "cartographer/mapping/3d/range_data_inserter_3d.h"

#include <cstdio>
#include <cstdlib>

namespace cartographer {
namespace mapping {
namespace {

void CheckOk(int status, const char* what) {
  if (status != CSM_OK) {
    std::fprintf(stderr, "%s failed: %s\n", what, csm_last_error());
    std::abort();
  }
}

constexpr uint16 kUpdateMarker = 1u << 15;
constexpr int kValueCount = 32768;

// probability_values.h:32-46,88-98 and probability_values.cc:28-57
float ClampProbability(const float p) { return p < kMinProbability ? kMinProbability : (p > kMaxProbability ? kMaxProbability : p); }
uint16 ProbabilityToValue(const float probability) {
  const int value = common::RoundToInt((ClampProbability(probability) - kMinProbability) *
                                       (32766.f / (kMaxProbability - kMinProbability))) +
                    1;
  return static_cast<uint16>(value);
}
float SlowValueToProbability(const int value) {  // value in [0, 32767]
  if (value == 0) return kMinProbability;  // kUnknownProbabilityValue
  const float kScale = (kMaxProbability - kMinProbability) / 32766.f;
  return value * kScale + (kMinProbability - kScale);
}

}  // namespace

proto::RangeDataInserterOptions3D CreateRangeDataInserterOptions3D(
    common::LuaParameterDictionary* const parameter_dictionary) {
  proto::RangeDataInserterOptions3D options;
  options.set_hit_probability(parameter_dictionary->GetDouble("hit_probability"));
  options.set_miss_probability(parameter_dictionary->GetDouble("miss_probability"));
  options.set_num_free_space_voxels(parameter_dictionary->GetInt("num_free_space_voxels"));
  options.set_intensity_threshold(parameter_dictionary->GetDouble("intensity_threshold"));
  CSM_CHECK(options.hit_probability() > 0.5);
  CSM_CHECK(options.miss_probability() < 0.5);
  return options;
}

std::vector<uint16> ComputeLookupTableToApplyOdds(const float odds) {
  std::vector<uint16> result;
  result.reserve(kValueCount);
  result.push_back(ProbabilityToValue(ProbabilityFromOdds(odds)) + kUpdateMarker);
  for (int cell = 1; cell != kValueCount; ++cell) {
    result.push_back(ProbabilityToValue(ProbabilityFromOdds(odds * Odds(SlowValueToProbability(cell)))) +
                     kUpdateMarker);
  }
  return result;
}

RangeDataInserter3D::RangeDataInserter3D(const proto::RangeDataInserterOptions3D& options)
    : RangeDataInserter3D(options, nullptr) {}

RangeDataInserter3D::RangeDataInserter3D(const proto::RangeDataInserterOptions3D& options, DeviceContext* context)
    : options_(options),
      hit_table_(ComputeLookupTableToApplyOdds(Odds(options_.hit_probability()))),
      miss_table_(ComputeLookupTableToApplyOdds(Odds(options_.miss_probability()))),
      context_(context) {}

void RangeDataInserter3D::Insert(const sensor::RangeData& range_data, HybridGrid* const hybrid_grid,
                                 IntensityHybridGrid* const intensity_hybrid_grid) const {
  CSM_CHECK(hybrid_grid != nullptr);              // reference: CHECK_NOTNULL
  CSM_CHECK(intensity_hybrid_grid == nullptr);    // not carried
  DeviceContext* ctx = context_ != nullptr ? context_ : DeviceContext::ThreadDefault();
  DeviceGrid device_grid(ctx, *hybrid_grid);
  Insert({&range_data}, &device_grid);
  device_grid.Download(hybrid_grid);
}

void RangeDataInserter3D::Insert(const std::vector<const sensor::RangeData*>& range_data,
                                 DeviceGrid* const grid) const {
  if (range_data.empty()) return;
  CSM_CHECK(grid != nullptr);
  DeviceContext* ctx = context_ != nullptr ? context_ : DeviceContext::ThreadDefault();
  std::vector<std::vector<float>> xyz(range_data.size());
  std::vector<csm_range_data3d> scans(range_data.size());
  for (size_t s = 0; s != range_data.size(); ++s) {
    const sensor::RangeData& rd = *range_data[s];
    xyz[s].reserve(3 * rd.returns.size());
    for (const sensor::RangefinderPoint& hit : rd.returns) {
      xyz[s].push_back(hit.position.x());
      xyz[s].push_back(hit.position.y());
      xyz[s].push_back(hit.position.z());
    }
    csm_range_data3d& out = scans[s];
    out.origin[0] = rd.origin.x();
    out.origin[1] = rd.origin.y();
    out.origin[2] = rd.origin.z();
    out.num_returns = static_cast<int32_t>(rd.returns.size());
    out.returns_xyz = xyz[s].data();
    out.hit_table = hit_table_.data();
    out.miss_table = miss_table_.data();
    out.num_free_space_voxels = options_.num_free_space_voxels();
    out.pad = 0;
  }
  CheckOk(csm_hgrid3d_insert(ctx->get(), grid->get(), scans.data(), static_cast<int>(scans.size())),
          "csm_hgrid3d_insert");
}

RangeDataInserter3D::DeviceGrid::DeviceGrid(DeviceContext* const context, const HybridGrid& hybrid_grid)
    : context_(context), resolution_(hybrid_grid.resolution()) {
  std::vector<int32_t> xyz;
  std::vector<uint16> values;
  hybrid_grid.Export(&xyz, &values);
  CheckOk(csm_hgrid3d_create(context_->get(), resolution_, xyz.data(), values.data(),
                             static_cast<int>(values.size()), &grid_),
          "csm_hgrid3d_create");
}

RangeDataInserter3D::DeviceGrid::~DeviceGrid() { csm_hgrid3d_destroy(context_->get(), grid_); }

void RangeDataInserter3D::DeviceGrid::Export(std::vector<int32_t>* const xyz, std::vector<uint16>* const values) const {
  int n = 0;
  CheckOk(csm_hgrid3d_export(context_->get(), grid_, nullptr, nullptr, 0, &n), "csm_hgrid3d_export");
  xyz->resize(3 * static_cast<size_t>(n));
  values->resize(n);
  if (n == 0) return;
  int m = 0;
  CheckOk(csm_hgrid3d_export(context_->get(), grid_, xyz->data(), values->data(), n, &m), "csm_hgrid3d_export");
  CSM_CHECK(m == n);
}

void RangeDataInserter3D::DeviceGrid::Download(HybridGrid* const hybrid_grid) const {
  CSM_CHECK(hybrid_grid != nullptr && hybrid_grid->resolution() == resolution_);
  std::vector<int32_t> xyz;
  std::vector<uint16> values;
  Export(&xyz, &values);
  // Insertion never clears a voxel, so writing the device's voxels over the host's is the update.
  for (size_t i = 0; i != values.size(); ++i) {
    *hybrid_grid->mutable_value(Eigen::Array3i(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2])) = values[i];
  }
}

}  // namespace mapping
}  // namespace cartographer
