// TEST INFRASTRUCTURE (oracle/_ref build only): stand-in for protoc's output of
// range_data_inserter_options_3d.proto.
#ifndef ORACLE_REF_SHIM_RANGE_DATA_INSERTER_OPTIONS_3D_PB_H_
#define ORACLE_REF_SHIM_RANGE_DATA_INSERTER_OPTIONS_3D_PB_H_
#include "shim_proto.h"
namespace cartographer {
namespace mapping {
namespace proto {
class RangeDataInserterOptions3D {
  SHIM_SCALAR(double, hit_probability)
  SHIM_SCALAR(double, miss_probability)
  SHIM_SCALAR(int32_t, num_free_space_voxels)
  SHIM_SCALAR(float, intensity_threshold)
};
}  // namespace proto
}  // namespace mapping
}  // namespace cartographer
#endif
