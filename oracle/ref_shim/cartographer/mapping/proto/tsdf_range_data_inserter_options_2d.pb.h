// TEST INFRASTRUCTURE (oracle/_ref build only): stand-in for protoc's output of
// tsdf_range_data_inserter_options_2d.proto.
#ifndef ORACLE_REF_SHIM_TSDF_INSERTER_OPTIONS_2D_PB_H_
#define ORACLE_REF_SHIM_TSDF_INSERTER_OPTIONS_2D_PB_H_
#include "cartographer/mapping/proto/normal_estimation_options_2d.pb.h"
namespace cartographer {
namespace mapping {
namespace proto {
class TSDFRangeDataInserterOptions2D {
  SHIM_SCALAR(double, truncation_distance)
  SHIM_SCALAR(double, maximum_weight)
  SHIM_SCALAR(bool, update_free_space)
  SHIM_MESSAGE(NormalEstimationOptions2D, normal_estimation_options)
  SHIM_SCALAR(bool, project_sdf_distance_to_scan_normal)
  SHIM_SCALAR(int32_t, update_weight_range_exponent)
  SHIM_SCALAR(double, update_weight_angle_scan_normal_to_ray_kernel_bandwidth)
  SHIM_SCALAR(double, update_weight_distance_cell_to_hit_kernel_bandwidth)
};
}  // namespace proto
}  // namespace mapping
}  // namespace cartographer
#endif
