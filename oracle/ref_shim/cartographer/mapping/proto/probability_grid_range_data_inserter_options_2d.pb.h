// TEST INFRASTRUCTURE (oracle/_ref build only): stand-in for protoc's output of
// probability_grid_range_data_inserter_options_2d.proto.
#ifndef ORACLE_REF_SHIM_PG_INSERTER_OPTIONS_2D_PB_H_
#define ORACLE_REF_SHIM_PG_INSERTER_OPTIONS_2D_PB_H_
#include "shim_proto.h"
namespace cartographer {
namespace mapping {
namespace proto {
class ProbabilityGridRangeDataInserterOptions2D {
  SHIM_SCALAR(double, hit_probability)
  SHIM_SCALAR(double, miss_probability)
  SHIM_SCALAR(bool, insert_free_space)
};
}  // namespace proto
}  // namespace mapping
}  // namespace cartographer
#endif
