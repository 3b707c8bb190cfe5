// TEST INFRASTRUCTURE (oracle/_ref build only): stand-in for protoc's output of
// cartographer/sensor/proto/adaptive_voxel_filter_options.proto.
#ifndef ORACLE_REF_SHIM_ADAPTIVE_VOXEL_FILTER_OPTIONS_PB_H_
#define ORACLE_REF_SHIM_ADAPTIVE_VOXEL_FILTER_OPTIONS_PB_H_
#include "shim_proto.h"
namespace cartographer {
namespace sensor {
namespace proto {
class AdaptiveVoxelFilterOptions {
  SHIM_SCALAR(float, max_length)
  SHIM_SCALAR(float, min_num_points)
  SHIM_SCALAR(float, max_range)
};
}  // namespace proto
}  // namespace sensor
}  // namespace cartographer
#endif
