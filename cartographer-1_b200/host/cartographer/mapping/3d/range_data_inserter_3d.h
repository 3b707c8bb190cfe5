// B200 implementation behind the reference's RangeDataInserter3D interface
// (cartographer/mapping/3d/range_data_inserter_3d.h:28-50).
//
// Insert(range_data, hybrid_grid) keeps the reference's signature and result: hits take the hit
// table once per voxel, the last num_free_space_voxels samples of every ray take the miss table
// once per voxel unless the voxel also holds a hit, then the update markers are dropped
// (range_data_inserter_3d.cc:44-76,112-131).  The voxel updates run on the device
// (csm_hgrid3d_insert); for the reference-signature call the grid goes up and comes back, which
// exists for drop-in parity.  The path that pays is DeviceGrid: a live submap's voxels stay in HBM
// over any number of insertions and are exported once, for the matcher's stack, when the submap is
// finished.  The intensity grid this fork adds (IntensityHybridGrid, .cc:78-92) is not carried:
// passing one aborts.  No CPU fallback.
#ifndef CARTOGRAPHER_MAPPING_3D_RANGE_DATA_INSERTER_3D_H_
#define CARTOGRAPHER_MAPPING_3D_RANGE_DATA_INSERTER_3D_H_

#include <vector>

#include "cartographer/mapping/internal/2d/scan_matching/fast_correlative_scan_matcher_2d.h"
#include "cartographer_compat.h"
#include "cartographer_compat_3d.h"
#include "csm_b200.h"

namespace cartographer {
namespace mapping {

namespace proto {
// mapping/proto/range_data_inserter_options_3d.proto
class RangeDataInserterOptions3D {
 public:
  double hit_probability() const { return hit_probability_; }
  double miss_probability() const { return miss_probability_; }
  int num_free_space_voxels() const { return num_free_space_voxels_; }
  double intensity_threshold() const { return intensity_threshold_; }
  void set_hit_probability(double v) { hit_probability_ = v; }
  void set_miss_probability(double v) { miss_probability_ = v; }
  void set_num_free_space_voxels(int v) { num_free_space_voxels_ = v; }
  void set_intensity_threshold(double v) { intensity_threshold_ = v; }

 private:
  double hit_probability_ = 0., miss_probability_ = 0., intensity_threshold_ = 0.;
  int num_free_space_voxels_ = 0;
};
}  // namespace proto

// range_data_inserter_3d.cc:94-110
proto::RangeDataInserterOptions3D CreateRangeDataInserterOptions3D(
    common::LuaParameterDictionary* parameter_dictionary);

// probability_values.cc:72-87
std::vector<uint16> ComputeLookupTableToApplyOdds(float odds);

class IntensityHybridGrid;  // not carried (see above)

using scan_matching::DeviceContext;

class RangeDataInserter3D {
 public:
  explicit RangeDataInserter3D(const proto::RangeDataInserterOptions3D& options);
  RangeDataInserter3D(const proto::RangeDataInserterOptions3D& options, DeviceContext* context);

  RangeDataInserter3D(const RangeDataInserter3D&) = delete;
  RangeDataInserter3D& operator=(const RangeDataInserter3D&) = delete;

  // Inserts 'range_data' into 'hybrid_grid' (reference :39-43).
  void Insert(const sensor::RangeData& range_data, HybridGrid* hybrid_grid,
              IntensityHybridGrid* intensity_hybrid_grid = nullptr) const;

  // The voxels of one HybridGrid kept on the device between insertions.
  class DeviceGrid {
   public:
    DeviceGrid(DeviceContext* context, const HybridGrid& hybrid_grid);
    ~DeviceGrid();
    DeviceGrid(const DeviceGrid&) = delete;
    DeviceGrid& operator=(const DeviceGrid&) = delete;
    csm_hgrid3d* get() const { return grid_; }
    float resolution() const { return resolution_; }
    // Non-default voxels as flat lists (what HybridGrid::Iterator visits), unordered.
    void Export(std::vector<int32_t>* xyz, std::vector<uint16>* values) const;
    // Replaces the contents of 'hybrid_grid' (same resolution required) by the device's.
    void Download(HybridGrid* hybrid_grid) const;

   private:
    DeviceContext* context_;
    float resolution_;
    csm_hgrid3d* grid_ = nullptr;
  };
  // Inserts every range data, in order, into the device-resident grid.
  void Insert(const std::vector<const sensor::RangeData*>& range_data, DeviceGrid* grid) const;

 private:
  const proto::RangeDataInserterOptions3D options_;
  const std::vector<uint16> hit_table_;
  const std::vector<uint16> miss_table_;
  DeviceContext* context_ = nullptr;
};

}  // namespace mapping
}  // namespace cartographer

#endif  // CARTOGRAPHER_MAPPING_3D_RANGE_DATA_INSERTER_3D_H_
