// TEST INFRASTRUCTURE (oracle/_ref build only): stand-in for protoc's output of map_limits.proto.
#ifndef ORACLE_REF_SHIM_MAP_LIMITS_PB_H_
#define ORACLE_REF_SHIM_MAP_LIMITS_PB_H_
#include "cartographer/mapping/proto/cell_limits_2d.pb.h"
#include "cartographer/transform/proto/transform.pb.h"
namespace cartographer {
namespace mapping {
namespace proto {
class MapLimits {
  SHIM_SCALAR(double, resolution)
  SHIM_MESSAGE(transform::proto::Vector2d, max)
  SHIM_MESSAGE(CellLimits, cell_limits)
};
}  // namespace proto
}  // namespace mapping
}  // namespace cartographer
#endif
