// TEST INFRASTRUCTURE (oracle/_ref build only): stand-in for the two messages of pose_graph.proto
// that mapping/id.h touches.
#ifndef ORACLE_REF_SHIM_POSE_GRAPH_PB_H_
#define ORACLE_REF_SHIM_POSE_GRAPH_PB_H_
#include "shim_proto.h"
namespace cartographer {
namespace mapping {
namespace proto {
class SubmapId { SHIM_SCALAR(int32_t, trajectory_id) SHIM_SCALAR(int32_t, submap_index) };
class NodeId { SHIM_SCALAR(int32_t, trajectory_id) SHIM_SCALAR(int32_t, node_index) };
}  // namespace proto
}  // namespace mapping
}  // namespace cartographer
#endif
