// TEST INFRASTRUCTURE (oracle/_ref build only): stand-in for protoc's output of
// submaps_options_2d.proto and grid_2d_options.proto (only the grid options are touched).
#ifndef ORACLE_REF_SHIM_SUBMAPS_OPTIONS_2D_PB_H_
#define ORACLE_REF_SHIM_SUBMAPS_OPTIONS_2D_PB_H_
#include "shim_proto.h"
namespace cartographer {
namespace mapping {
namespace proto {
enum GridOptions2D_GridType {
  GridOptions2D_GridType_INVALID_GRID = 0,
  GridOptions2D_GridType_PROBABILITY_GRID = 1,
  GridOptions2D_GridType_TSDF = 2
};
inline bool GridOptions2D_GridType_Parse(const std::string& name, GridOptions2D_GridType* value) {
  if (name == "INVALID_GRID") return *value = GridOptions2D_GridType_INVALID_GRID, true;
  if (name == "PROBABILITY_GRID") return *value = GridOptions2D_GridType_PROBABILITY_GRID, true;
  if (name == "TSDF") return *value = GridOptions2D_GridType_TSDF, true;
  return false;
}
class GridOptions2D {
 public:
  typedef GridOptions2D_GridType GridType;
  static constexpr GridType INVALID_GRID = GridOptions2D_GridType_INVALID_GRID;
  static constexpr GridType PROBABILITY_GRID = GridOptions2D_GridType_PROBABILITY_GRID;
  static constexpr GridType TSDF = GridOptions2D_GridType_TSDF;
  SHIM_SCALAR(GridType, grid_type)
  SHIM_SCALAR(float, resolution)
};
class RangeDataInserterOptions {};
class SubmapsOptions2D {
  SHIM_SCALAR(int32_t, num_range_data)
  SHIM_MESSAGE(GridOptions2D, grid_options_2d)
  SHIM_MESSAGE(RangeDataInserterOptions, range_data_inserter_options)
};
}  // namespace proto
}  // namespace mapping
}  // namespace cartographer
#endif
