// A saved mapping state (.pbstream) -> the submaps and nodes the constraint search takes
// ("lets the bench run on real .pbstream maps").  Mirrors
//   io::ProtoStreamDeserializer                       io/proto_stream_deserializer.{h,cc}
//   the kSubmap / kNode cases of MapBuilder::LoadState mapping/map_builder.cc:227-328
//   Submap2D(const proto::Submap2D&, tables)          mapping/2d/submap_2d.cc:78-94
//   Submap3D(const proto::Submap3D&)                  mapping/3d/submap_3d.cc:259-262,285-300
//   FromProto(const proto::TrajectoryNodeData&)       mapping/trajectory_node.cc:50-69
//   transform::ToRigid3 / ToEigen                     transform/transform.cc:39-46,71-74
// on the wire format (io/proto_wire.h; libprotobuf is not in this image).  Stream order
// (proto_stream_deserializer.cc:46-71): SerializationHeader, then SerializedData{pose_graph},
// SerializedData{all_trajectory_builder_options}, then any number of SerializedData records.
// Message layouts: mapping/proto/serialization.proto, pose_graph.proto, trajectory.proto,
// submap.proto, trajectory_node_data.proto, sensor/proto/sensor.proto (CompressedPointCloud),
// transform/proto/transform.proto.  Records the scan-matching path does not read (IMU, odometry,
// fixed-frame and landmark data, trajectory data, the builder options' contents) are skipped.
#ifndef CARTOGRAPHER_IO_PROTO_STREAM_DESERIALIZER_H_
#define CARTOGRAPHER_IO_PROTO_STREAM_DESERIALIZER_H_

#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#include "cartographer/io/proto_stream.h"
#include "cartographer_compat.h"
#include "cartographer_compat_3d.h"

namespace cartographer {
namespace io {

// io/internal/mapping_state_serialization.h:28-29
constexpr int kMappingStateSerializationFormatVersion = 2;
constexpr int kFormatVersionWithoutSubmapHistograms = 1;

// mapping/proto/pose_graph.proto PoseGraph + trajectory.proto Trajectory: the fields LoadState reads.
struct PoseGraphProto {
  struct Trajectory {
    struct Node {
      int node_index = 0;
      int64_t timestamp = 0;
      transform::Rigid3d pose;  // tracking -> global map frame
    };
    struct Submap {
      int submap_index = 0;
      transform::Rigid3d pose;  // submap -> global map frame
    };
    int trajectory_id = 0;
    std::vector<Node> node;
    std::vector<Submap> submap;
  };
  std::vector<mapping::PoseGraphInterface::Constraint> constraint;
  std::vector<Trajectory> trajectory;
};

// serialization.proto Submap, unpacked the way the Submap2D / Submap3D proto constructors do.
struct SubmapProto {
  mapping::SubmapId submap_id{0, 0};
  bool has_submap_2d = false, has_submap_3d = false;
  transform::Rigid3d local_pose;
  int num_range_data = 0;
  bool finished = false;
  std::unique_ptr<mapping::Grid2D> grid;  // 2D; null when the message carries no grid
  std::unique_ptr<mapping::HybridGrid> high_resolution_hybrid_grid, low_resolution_hybrid_grid;  // 3D
  std::vector<float> rotational_scan_matcher_histogram;                                          // 3D
};

// serialization.proto Node; node_data is TrajectoryNode::Data as trajectory_node.cc:50-69 rebuilds it.
struct NodeProto {
  mapping::NodeId node_id{0, 0};
  int64_t timestamp = 0;  // universal time, 100 ns ticks (common/time.h)
  mapping::TrajectoryNode::Data node_data;
  transform::Rigid3d local_pose;
};

struct SerializedData {
  // The oneof's field numbers (serialization.proto:79-91), named like protoc's DataCase.
  enum DataCase {
    DATA_NOT_SET = 0,
    kPoseGraph = 1,
    kAllTrajectoryBuilderOptions = 2,
    kSubmap = 3,
    kNode = 4,
    kTrajectoryData = 5,
    kImuData = 6,
    kOdometryData = 7,
    kFixedFramePoseData = 8,
    kLandmarkData = 9
  };
  DataCase data_case() const { return data_case_; }
  bool has_pose_graph() const { return data_case_ == kPoseGraph; }
  bool has_all_trajectory_builder_options() const { return data_case_ == kAllTrajectoryBuilderOptions; }
  const PoseGraphProto& pose_graph() const { return pose_graph_; }
  int options_with_sensor_ids_size() const { return options_with_sensor_ids_size_; }
  const SubmapProto& submap() const { return submap_; }
  SubmapProto* mutable_submap() { return &submap_; }
  const NodeProto& node() const { return node_; }
  NodeProto* mutable_node() { return &node_; }

  DataCase data_case_ = DATA_NOT_SET;
  PoseGraphProto pose_graph_;
  int options_with_sensor_ids_size_ = 0;
  SubmapProto submap_;
  NodeProto node_;
};

// One serialized SerializedData message -> `out`.  Grids of 2D submaps borrow `conversion_tables`.
void ParseSerializedData(const std::string& serialized, mapping::ValueConversionTables* conversion_tables,
                         SerializedData* out);

class ProtoStreamDeserializer {
 public:
  // Reads the header, the pose graph and the builder options; aborts on an unsupported
  // format version or a stream out of order (the reference's CHECKs).
  ProtoStreamDeserializer(ProtoStreamReader* reader, mapping::ValueConversionTables* conversion_tables);
  ProtoStreamDeserializer(const ProtoStreamDeserializer&) = delete;
  ProtoStreamDeserializer& operator=(const ProtoStreamDeserializer&) = delete;

  uint32_t format_version() const { return format_version_; }  // header().format_version()
  const PoseGraphProto& pose_graph() const { return pose_graph_.pose_graph(); }
  int options_with_sensor_ids_size() const {
    return all_trajectory_builder_options_.options_with_sensor_ids_size();
  }
  // The next record; false at the end of the stream.  `raw`, if given, receives its serialized bytes.
  bool ReadNextSerializedData(SerializedData* data, std::string* raw = nullptr);
  // The serialized PoseGraph and AllTrajectoryBuilderOptions records as read.
  const std::string& raw_pose_graph_record() const { return raw_pose_graph_; }
  const std::string& raw_options_record() const { return raw_options_; }

 private:
  ProtoStreamReader* reader_;
  mapping::ValueConversionTables* conversion_tables_;
  uint32_t format_version_ = 0;
  SerializedData pose_graph_;
  SerializedData all_trajectory_builder_options_;
  std::string raw_pose_graph_, raw_options_;
};

// What the kSubmap / kNode cases of MapBuilder::LoadState hand to the pose graph
// (AddSubmapFromProto(global pose, submap), AddNodeFromProto(global pose, node)), kept as lists in
// stream order; the poses come from the PoseGraph record and a submap or node without one aborts
// (`submap_poses.at`, map_builder.cc:316,325).
struct MappingState {
  struct Submap {
    transform::Rigid3d global_pose;
    SubmapProto proto;
  };
  struct Node {
    transform::Rigid3d global_pose;
    NodeProto proto;
  };
  // Every record after the builder options, in stream order: `index` points into submaps / nodes
  // for those two kinds; `raw` holds the serialized record of every other kind (and of the first
  // two records of the stream below), so that the state can be written back (grid_proto_writer.h).
  struct Record {
    SerializedData::DataCase data_case = SerializedData::DATA_NOT_SET;
    int index = -1;
    std::string raw;
  };
  uint32_t format_version = 0;
  std::unique_ptr<mapping::ValueConversionTables> conversion_tables;  // the 2D grids borrow it
  std::vector<Submap> submaps;
  std::vector<Node> nodes;
  std::vector<mapping::PoseGraphInterface::Constraint> constraints;
  std::vector<Record> records;
  std::string raw_pose_graph_record, raw_options_record;
};

// `expect_3d` applies LoadState's refusal of format version 1 (submaps without rotational
// histograms) for a 3D map (map_builder.cc:290-296).
MappingState LoadState(ProtoStreamReader* reader, bool expect_3d = false);
MappingState LoadStateFromFile(const std::string& state_filename, bool expect_3d = false);

}  // namespace io
}  // namespace cartographer

#endif  // CARTOGRAPHER_IO_PROTO_STREAM_DESERIALIZER_H_
