// TEST INFRASTRUCTURE (oracle/_ref build only): stand-in for protoc's output of
// fast_correlative_scan_matcher_options_3d.proto.
#ifndef ORACLE_REF_SHIM_FCSM_OPTIONS_3D_PB_H_
#define ORACLE_REF_SHIM_FCSM_OPTIONS_3D_PB_H_
#include "shim_proto.h"
namespace cartographer {
namespace mapping {
namespace scan_matching {
namespace proto {
class FastCorrelativeScanMatcherOptions3D {
  SHIM_SCALAR(int32_t, branch_and_bound_depth)
  SHIM_SCALAR(int32_t, full_resolution_depth)
  SHIM_SCALAR(double, min_rotational_score)
  SHIM_SCALAR(double, min_low_resolution_score)
  SHIM_SCALAR(double, linear_xy_search_window)
  SHIM_SCALAR(double, linear_z_search_window)
  SHIM_SCALAR(double, angular_search_window)
};
}  // namespace proto
}  // namespace scan_matching
}  // namespace mapping
}  // namespace cartographer
#endif
