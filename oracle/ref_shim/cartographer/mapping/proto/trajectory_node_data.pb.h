// TEST INFRASTRUCTURE (oracle/_ref build only): stand-in for protoc's output of
// trajectory_node_data.proto (pulled in through mapping/trajectory_node.h).
#ifndef ORACLE_REF_SHIM_TRAJECTORY_NODE_DATA_PB_H_
#define ORACLE_REF_SHIM_TRAJECTORY_NODE_DATA_PB_H_
#include "cartographer/sensor/proto/sensor.pb.h"
#include "cartographer/transform/proto/transform.pb.h"
namespace cartographer {
namespace mapping {
namespace proto {
class TrajectoryNodeData {
  SHIM_SCALAR(int64_t, timestamp)
  SHIM_MESSAGE(transform::proto::Quaterniond, gravity_alignment)
  SHIM_MESSAGE(sensor::proto::CompressedPointCloud, filtered_gravity_aligned_point_cloud)
  SHIM_MESSAGE(sensor::proto::CompressedPointCloud, high_resolution_point_cloud)
  SHIM_MESSAGE(sensor::proto::CompressedPointCloud, low_resolution_point_cloud)
  SHIM_REPEATED(float, rotational_scan_matcher_histogram)
  SHIM_MESSAGE(transform::proto::Rigid3d, local_pose)
};
}  // namespace proto
}  // namespace mapping
}  // namespace cartographer
#endif
