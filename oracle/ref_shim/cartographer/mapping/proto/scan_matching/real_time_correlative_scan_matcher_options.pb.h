// TEST INFRASTRUCTURE (oracle/_ref build only): stand-in for protoc's output of
// real_time_correlative_scan_matcher_options.proto.
#ifndef ORACLE_REF_SHIM_RTCSM_OPTIONS_PB_H_
#define ORACLE_REF_SHIM_RTCSM_OPTIONS_PB_H_
#include "shim_proto.h"
namespace cartographer {
namespace mapping {
namespace scan_matching {
namespace proto {
class RealTimeCorrelativeScanMatcherOptions {
  SHIM_SCALAR(double, linear_search_window)
  SHIM_SCALAR(double, angular_search_window)
  SHIM_SCALAR(double, translation_delta_cost_weight)
  SHIM_SCALAR(double, rotation_delta_cost_weight)
};
}  // namespace proto
}  // namespace scan_matching
}  // namespace mapping
}  // namespace cartographer
#endif
