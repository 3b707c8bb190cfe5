// Voxel filtering of the clouds a node carries (what produces
// filtered_gravity_aligned_point_cloud and the 3D high / low resolution clouds).  Mirrors this
// fork's sensor/internal/voxel_filter.{h,cc}:
//   VoxelFilter(points, resolution)            cc:139-161,206-232   one point per voxel, chosen by
//                                              reservoir sampling in point order
//   AdaptiveVoxelFilter(cloud, options)        cc:32-76,264-269     max-range crop, then the largest
//                                              voxel edge that keeps min_num_points
//   VoxelFilterEdge(cloud, rad, ratio)         cc:89-137,165-204    this fork's addition: keeps the
//                                              points of sparsely hit voxels
// Host code by nature: ONE std::minstd_rand0 advances only when a point falls into an occupied voxel,
// through std::uniform_int_distribution (which may draw more than once per call), so every choice
// depends on all earlier points -- a few thousand points per node, tens of microseconds.  The same
// standard-library generator and distribution are used here, so with the same libstdc++ the kept
// points are the reference's, index for index (tests/test_voxel_filter_cpu.py holds them to the
// reference's own source compiled into oracle/_ref).  Point intensities and the TimedPointCloud /
// RangeMeasurement overloads are not carried: the host layer's PointCloud has positions only.
#ifndef CARTOGRAPHER_SENSOR_INTERNAL_VOXEL_FILTER_H_
#define CARTOGRAPHER_SENSOR_INTERNAL_VOXEL_FILTER_H_

#include <vector>

#include "cartographer/common/lua_parameter_dictionary.h"
#include "cartographer_compat.h"

namespace cartographer {
namespace sensor {
namespace proto {

// sensor/proto/adaptive_voxel_filter_options.proto (all three fields are floats there)
class AdaptiveVoxelFilterOptions {
 public:
  float max_length() const { return max_length_; }
  float min_num_points() const { return min_num_points_; }
  float max_range() const { return max_range_; }
  void set_max_length(float v) { max_length_ = v; }
  void set_min_num_points(float v) { min_num_points_ = v; }
  void set_max_range(float v) { max_range_ = v; }

 private:
  float max_length_ = 0.f, min_num_points_ = 0.f, max_range_ = 0.f;
};

}  // namespace proto

PointCloud VoxelFilterEdge(const PointCloud& point_cloud, float rad, float voxeledgeratio);

std::vector<RangefinderPoint> VoxelFilter(const std::vector<RangefinderPoint>& points, float resolution);
PointCloud VoxelFilter(const PointCloud& point_cloud, float resolution);

proto::AdaptiveVoxelFilterOptions CreateAdaptiveVoxelFilterOptions(
    common::LuaParameterDictionary* parameter_dictionary);

PointCloud AdaptiveVoxelFilter(const PointCloud& point_cloud, const proto::AdaptiveVoxelFilterOptions& options);

// The flags behind VoxelFilter: kept[i] for every input point (RandomizedVoxelFilterIndices, cc:139-161).
std::vector<bool> VoxelFilterIndices(const std::vector<RangefinderPoint>& points, float resolution);

}  // namespace sensor
}  // namespace cartographer

#endif  // CARTOGRAPHER_SENSOR_INTERNAL_VOXEL_FILTER_H_
