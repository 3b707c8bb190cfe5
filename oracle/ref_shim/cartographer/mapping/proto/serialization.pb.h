// TEST INFRASTRUCTURE (oracle/_ref build only): mapping/submaps.h names proto::Submap in pure
// virtual signatures only.
#ifndef ORACLE_REF_SHIM_SERIALIZATION_PB_H_
#define ORACLE_REF_SHIM_SERIALIZATION_PB_H_
#include "shim_proto.h"
namespace cartographer {
namespace mapping {
namespace proto {
class Submap {};
}  // namespace proto
}  // namespace mapping
}  // namespace cartographer
#endif
