// TEST INFRASTRUCTURE (oracle/_ref build only): stand-in for protoc's output of
// fast_correlative_scan_matcher_options_2d.proto.
#ifndef ORACLE_REF_SHIM_FCSM_OPTIONS_2D_PB_H_
#define ORACLE_REF_SHIM_FCSM_OPTIONS_2D_PB_H_
#include "shim_proto.h"
namespace cartographer {
namespace mapping {
namespace scan_matching {
namespace proto {
class FastCorrelativeScanMatcherOptions2D {
  SHIM_SCALAR(double, linear_search_window)
  SHIM_SCALAR(double, angular_search_window)
  SHIM_SCALAR(int32_t, branch_and_bound_depth)
};
}  // namespace proto
}  // namespace scan_matching
}  // namespace mapping
}  // namespace cartographer
#endif
