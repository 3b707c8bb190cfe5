// TEST INFRASTRUCTURE (oracle/_ref build only): stand-in for protoc's output of hybrid_grid.proto.
#ifndef ORACLE_REF_SHIM_HYBRID_GRID_PB_H_
#define ORACLE_REF_SHIM_HYBRID_GRID_PB_H_
#include "shim_proto.h"
namespace cartographer {
namespace mapping {
namespace proto {
class HybridGrid {
  SHIM_SCALAR(float, resolution)
  SHIM_REPEATED(int32_t, x_indices)
  SHIM_REPEATED(int32_t, y_indices)
  SHIM_REPEATED(int32_t, z_indices)
  SHIM_REPEATED(int32_t, values)
};
}  // namespace proto
}  // namespace mapping
}  // namespace cartographer
#endif
