// TEST INFRASTRUCTURE (oracle/_ref build only): stand-in for protoc's output of cell_limits_2d.proto.
#ifndef ORACLE_REF_SHIM_CELL_LIMITS_PB_H_
#define ORACLE_REF_SHIM_CELL_LIMITS_PB_H_
#include "shim_proto.h"
namespace cartographer {
namespace mapping {
namespace proto {
class CellLimits { SHIM_SCALAR(int32_t, num_x_cells) SHIM_SCALAR(int32_t, num_y_cells) };
}  // namespace proto
}  // namespace mapping
}  // namespace cartographer
#endif
