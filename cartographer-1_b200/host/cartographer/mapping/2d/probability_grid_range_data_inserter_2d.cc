#include "cartographer/mapping/2d/probability_grid_range_data_inserter_2d.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace cartographer {
namespace mapping {
namespace {

void CheckOk(int status, const char* what) {
  if (status != CSM_OK) {
    std::fprintf(stderr, "%s failed: %s\n", what, csm_last_error());
    std::abort();
  }
}

struct Packed {
  std::vector<float> returns, misses;
  csm_range_data2d scan;
};

void Pack(const sensor::RangeData& rd, const MapLimits& limits, const std::vector<uint16>& hit_table,
          const std::vector<uint16>& miss_table, const bool insert_free_space, Packed* out) {
  for (const sensor::RangefinderPoint& p : rd.returns) {
    out->returns.push_back(p.position.x());
    out->returns.push_back(p.position.y());
  }
  for (const sensor::RangefinderPoint& p : rd.misses) {
    out->misses.push_back(p.position.x());
    out->misses.push_back(p.position.y());
  }
  csm_range_data2d& s = out->scan;
  std::memset(&s, 0, sizeof(s));
  s.max_x = limits.max().x();
  s.max_y = limits.max().y();
  s.resolution = limits.resolution();
  s.origin_x = rd.origin.x();
  s.origin_y = rd.origin.y();
  s.returns_xy = out->returns.data();
  s.num_returns = static_cast<int32_t>(rd.returns.size());
  s.misses_xy = out->misses.data();
  s.num_misses = static_cast<int32_t>(rd.misses.size());
  s.hit_table = hit_table.data();
  s.miss_table = miss_table.data();
  s.insert_free_space = insert_free_space ? 1 : 0;
}

}  // namespace

// GrowAsNeeded (reference .cc:33-45): bounding box of origin, returns and misses, padded.
void ProbabilityGridRangeDataInserter2D::GrowAsNeeded(const sensor::RangeData& range_data,
                                                      Grid2D* const grid) {
  float min_x = range_data.origin.x(), max_x = min_x, min_y = range_data.origin.y(), max_y = min_y;
  auto extend = [&](const sensor::Vector3f& p) {
    min_x = std::min(min_x, p.x());
    max_x = std::max(max_x, p.x());
    min_y = std::min(min_y, p.y());
    max_y = std::max(max_y, p.y());
  };
  for (const sensor::RangefinderPoint& hit : range_data.returns) extend(hit.position);
  for (const sensor::RangefinderPoint& miss : range_data.misses) extend(miss.position);
  constexpr float kPadding = 1e-6f;
  grid->GrowLimits(min_x - kPadding * 1.f, min_y - kPadding * 1.f);
  grid->GrowLimits(max_x + kPadding * 1.f, max_y + kPadding * 1.f);
}

std::vector<uint16> ComputeLookupTableToApplyCorrespondenceCostOdds(const float odds) {
  ValueConversionTables tables;
  const std::vector<float>& cost = *tables.GetConversionTable(kMaxCorrespondenceCost, kMinCorrespondenceCost,
                                                              kMaxCorrespondenceCost);
  constexpr uint16 kUpdateMarker = 1u << 15;
  std::vector<uint16> result;
  result.reserve(32768);
  result.push_back(CorrespondenceCostToValue(1.f - ProbabilityFromOdds(odds)) + kUpdateMarker);
  for (int cell = 1; cell != 32768; ++cell) {
    result.push_back(CorrespondenceCostToValue(1.f - ProbabilityFromOdds(odds * Odds(1.f - cost[cell]))) +
                     kUpdateMarker);
  }
  return result;
}

ProbabilityGridRangeDataInserter2D::ProbabilityGridRangeDataInserter2D(
    const proto::ProbabilityGridRangeDataInserterOptions2D& options)
    : ProbabilityGridRangeDataInserter2D(options, nullptr) {}

ProbabilityGridRangeDataInserter2D::ProbabilityGridRangeDataInserter2D(
    const proto::ProbabilityGridRangeDataInserterOptions2D& options, DeviceContext* context)
    : options_(options),
      hit_table_(ComputeLookupTableToApplyCorrespondenceCostOdds(
          Odds(static_cast<float>(options.hit_probability())))),
      miss_table_(ComputeLookupTableToApplyCorrespondenceCostOdds(
          Odds(static_cast<float>(options.miss_probability())))),
      context_(context) {}

void ProbabilityGridRangeDataInserter2D::Insert(const sensor::RangeData& range_data,
                                                Grid2D* const grid) const {
  CSM_CHECK(grid != nullptr);
  CSM_CHECK(grid->GetGridType() == GridType::PROBABILITY_GRID);
  // By not finishing the update after hits are inserted, we give hits priority (reference :97-99):
  // on the device the hit flag wins over the miss flag of the same cell.
  GrowAsNeeded(range_data, grid);
  DeviceContext* ctx = context_ != nullptr ? context_ : DeviceContext::ThreadDefault();
  DeviceGrid device_grid(ctx, *grid);
  Insert({&range_data}, &device_grid);
  device_grid.Download(grid);
}

void ProbabilityGridRangeDataInserter2D::Insert(const std::vector<const sensor::RangeData*>& range_data,
                                                DeviceGrid* const grid) const {
  if (range_data.empty()) return;
  DeviceContext* ctx = context_ != nullptr ? context_ : DeviceContext::ThreadDefault();
  std::vector<Packed> packed(range_data.size());
  std::vector<csm_range_data2d> scans(range_data.size());
  for (size_t i = 0; i != range_data.size(); ++i) {
    Pack(*range_data[i], grid->limits(), hit_table_, miss_table_, options_.insert_free_space(), &packed[i]);
    scans[i] = packed[i].scan;
  }
  CheckOk(csm_grid2d_insert(ctx->get(), grid->get(), scans.data(), static_cast<int>(scans.size())),
          "csm_grid2d_insert");
}

ProbabilityGridRangeDataInserter2D::DeviceGrid::DeviceGrid(DeviceContext* context, const Grid2D& grid)
    : context_(context), limits_(grid.limits()) {
  CheckOk(csm_grid2d_create(context_->get(), grid.correspondence_cost_cells().data(),
                            limits_.cell_limits().num_x_cells, limits_.cell_limits().num_y_cells, &grid_),
          "csm_grid2d_create");
}

ProbabilityGridRangeDataInserter2D::DeviceGrid::~DeviceGrid() { csm_grid2d_destroy(context_->get(), grid_); }

void ProbabilityGridRangeDataInserter2D::DeviceGrid::Download(Grid2D* const grid) const {
  CSM_CHECK(grid->limits().cell_limits().num_x_cells == limits_.cell_limits().num_x_cells &&
            grid->limits().cell_limits().num_y_cells == limits_.cell_limits().num_y_cells);
  CheckOk(csm_grid2d_download(context_->get(), grid_, grid->mutable_correspondence_cost_cells()->data()),
          "csm_grid2d_download");
}

}  // namespace mapping
}  // namespace cartographer
