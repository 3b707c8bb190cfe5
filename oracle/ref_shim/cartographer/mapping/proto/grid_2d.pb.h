// TEST INFRASTRUCTURE (oracle/_ref build only): stand-in for protoc's output of grid_2d.proto
// (with probability_grid.proto and tsdf_2d.proto, which it imports).
#ifndef ORACLE_REF_SHIM_GRID_2D_PB_H_
#define ORACLE_REF_SHIM_GRID_2D_PB_H_
#include "cartographer/mapping/proto/map_limits.pb.h"
namespace cartographer {
namespace mapping {
namespace proto {
class ProbabilityGrid {};
class TSDF2D {
  SHIM_SCALAR(float, truncation_distance)
  SHIM_SCALAR(float, max_weight)
  SHIM_REPEATED(int32_t, weight_cells)
};
class Grid2D_CellBox {
  SHIM_SCALAR(int32_t, max_x) SHIM_SCALAR(int32_t, max_y) SHIM_SCALAR(int32_t, min_x) SHIM_SCALAR(int32_t, min_y)
};
class Grid2D {
 public:
  typedef Grid2D_CellBox CellBox;
  SHIM_MESSAGE(MapLimits, limits)
  SHIM_REPEATED(int32_t, cells)
  SHIM_MESSAGE(CellBox, known_cells_box)
  SHIM_MESSAGE(ProbabilityGrid, probability_grid_2d)
  SHIM_MESSAGE(TSDF2D, tsdf_2d)
  SHIM_SCALAR(float, min_correspondence_cost)
  SHIM_SCALAR(float, max_correspondence_cost)
};
}  // namespace proto
}  // namespace mapping
}  // namespace cartographer
#endif
