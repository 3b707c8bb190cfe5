// TEST INFRASTRUCTURE (oracle/_ref build only): stand-in for protoc's output of
// normal_estimation_options_2d.proto.
#ifndef ORACLE_REF_SHIM_NORMAL_ESTIMATION_OPTIONS_2D_PB_H_
#define ORACLE_REF_SHIM_NORMAL_ESTIMATION_OPTIONS_2D_PB_H_
#include "shim_proto.h"
namespace cartographer {
namespace mapping {
namespace proto {
class NormalEstimationOptions2D {
  SHIM_SCALAR(int32_t, num_normal_samples)
  SHIM_SCALAR(float, sample_radius)
};
}  // namespace proto
}  // namespace mapping
}  // namespace cartographer
#endif
