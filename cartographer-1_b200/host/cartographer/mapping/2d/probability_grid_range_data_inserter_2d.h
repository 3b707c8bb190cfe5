// B200 implementation behind the reference's ProbabilityGridRangeDataInserter2D interface
// (cartographer/mapping/2d/probability_grid_range_data_inserter_2d.h:35-52).
//
// Insert(range_data, grid) keeps the reference's signature and result: the grid grows to
// hold the range data (on the host: GrowLimits only re-allocates), then the rays are cast
// on the device (csm_grid2d_insert) and the cells come back -- bit-identical to the
// reference's.  That round trip exists for drop-in parity; the path that pays is the
// device-resident one: DeviceGrid keeps the cells of a live submap in HBM, takes any number
// of Insert calls and hands its device cells to the precomputation-stack build of the
// finished submap.  No CPU fallback.
#ifndef CARTOGRAPHER_MAPPING_2D_PROBABILITY_GRID_RANGE_DATA_INSERTER_2D_H_
#define CARTOGRAPHER_MAPPING_2D_PROBABILITY_GRID_RANGE_DATA_INSERTER_2D_H_

#include <vector>

#include "cartographer/mapping/internal/2d/scan_matching/fast_correlative_scan_matcher_2d.h"
#include "cartographer_compat.h"
#include "csm_b200.h"

namespace cartographer {
namespace mapping {

// probability_values.cc:89-105
std::vector<uint16> ComputeLookupTableToApplyCorrespondenceCostOdds(float odds);

using scan_matching::DeviceContext;

class ProbabilityGridRangeDataInserter2D {
 public:
  explicit ProbabilityGridRangeDataInserter2D(
      const proto::ProbabilityGridRangeDataInserterOptions2D& options);
  ProbabilityGridRangeDataInserter2D(const proto::ProbabilityGridRangeDataInserterOptions2D& options,
                                     DeviceContext* context);

  ProbabilityGridRangeDataInserter2D(const ProbabilityGridRangeDataInserter2D&) = delete;
  ProbabilityGridRangeDataInserter2D& operator=(const ProbabilityGridRangeDataInserter2D&) = delete;

  // Inserts 'range_data' into 'grid' (reference :44-46; GridInterface* there).
  void Insert(const sensor::RangeData& range_data, Grid2D* grid) const;

  // GrowAsNeeded (reference .cc:33-45): grows 'grid' to hold the padded bounding box of the
  // range data.  Insert(range_data, grid) calls it; users of a DeviceGrid call it on the host
  // grid for every scan they are going to insert before creating the DeviceGrid.
  static void GrowAsNeeded(const sensor::RangeData& range_data, Grid2D* grid);

  // The cells of one grid kept on the device between insertions.
  class DeviceGrid {
   public:
    DeviceGrid(DeviceContext* context, const Grid2D& grid);
    ~DeviceGrid();
    DeviceGrid(const DeviceGrid&) = delete;
    DeviceGrid& operator=(const DeviceGrid&) = delete;
    csm_grid2d* get() const { return grid_; }
    const MapLimits& limits() const { return limits_; }
    const uint16* cells_dev() const { return csm_grid2d_cells_dev(grid_); }
    void Download(Grid2D* grid) const;  // same limits required

   private:
    DeviceContext* context_;
    MapLimits limits_;
    csm_grid2d* grid_ = nullptr;
  };
  // Inserts every range data into the device-resident grid (which must already contain them).
  void Insert(const std::vector<const sensor::RangeData*>& range_data, DeviceGrid* grid) const;

 private:
  const proto::ProbabilityGridRangeDataInserterOptions2D options_;
  const std::vector<uint16> hit_table_;
  const std::vector<uint16> miss_table_;
  DeviceContext* context_ = nullptr;
};

}  // namespace mapping
}  // namespace cartographer

#endif  // CARTOGRAPHER_MAPPING_2D_PROBABILITY_GRID_RANGE_DATA_INSERTER_2D_H_
