#include "cartographer/io/proto_stream_deserializer.h"

#include <map>
#include <utility>

#include "cartographer/io/grid_proto_reader.h"
#include "cartographer/io/proto_wire.h"
#include "cartographer/sensor/compressed_point_cloud.h"

namespace cartographer {
namespace io {
namespace {

using wire::Reader;
using wire::Span;

// Every parser below walks one message; unknown fields are skipped, a malformed one aborts
// (ParseFromString failing is a CHECK in the reference's readers).
#define CSM_PARSE_LOOP(span)   \
  Reader r(span);              \
  int field = 0, type = 0;     \
  while (r.Next(&field, &type))

// transform.proto Vector3d {double x = 1, y = 2, z = 3}
transform::Vector3d ParseVector3d(const Span& s) {
  double v[3] = {0., 0., 0.};
  CSM_PARSE_LOOP(s) {
    if (field >= 1 && field <= 3 && type == 1) v[field - 1] = r.Double();
    else r.Skip(type);
  }
  CSM_CHECK(r.ok());
  return transform::Vector3d(v[0], v[1], v[2]);
}
// transform.proto Quaterniond {double x = 1, y = 2, z = 3, w = 4}; ToEigen(w, x, y, z), no
// normalisation (transform.cc:43-46).  An absent message is all zeros, as protobuf's default.
transform::Quaterniond ParseQuaterniond(const Span& s) {
  double v[4] = {0., 0., 0., 0.};
  CSM_PARSE_LOOP(s) {
    if (field >= 1 && field <= 4 && type == 1) v[field - 1] = r.Double();
    else r.Skip(type);
  }
  CSM_CHECK(r.ok());
  return transform::Quaterniond(v[3], v[0], v[1], v[2]);
}
// transform.proto Rigid3d {Vector3d translation = 1; Quaterniond rotation = 2}
transform::Rigid3d ParseRigid3d(const Span& s) {
  transform::Vector3d translation(0., 0., 0.);
  transform::Quaterniond rotation(0., 0., 0., 0.);
  CSM_PARSE_LOOP(s) {
    if (field == 1 && type == 2) translation = ParseVector3d(r.Bytes());
    else if (field == 2 && type == 2) rotation = ParseQuaterniond(r.Bytes());
    else r.Skip(type);
  }
  CSM_CHECK(r.ok());
  return transform::Rigid3d(translation, rotation);
}
// pose_graph.proto SubmapId / NodeId {int32 trajectory_id = 1; int32 index = 2}
std::pair<int, int> ParseId(const Span& s) {
  int v[2] = {0, 0};
  CSM_PARSE_LOOP(s) {
    if (field >= 1 && field <= 2 && type == 0) v[field - 1] = static_cast<int32_t>(r.Varint());
    else r.Skip(type);
  }
  CSM_CHECK(r.ok());
  return {v[0], v[1]};
}
// A repeated float: packed (wire type 2) or one unpacked element (wire type 5).
void RepeatedFloat(Reader* r, int type, std::vector<float>* out) {
  if (type == 2) {
    Reader packed(r->Bytes());
    while (!packed.done()) out->push_back(packed.Float());
    CSM_CHECK(packed.ok());
  } else if (type == 5) {
    out->push_back(r->Float());
  } else {
    r->Skip(type);
  }
}

// sensor.proto CompressedPointCloud {int32 num_points = 1; repeated int32 point_data = 3}
sensor::PointCloud ParseCompressedPointCloud(const Span& s) {
  int32_t num_points = 0;
  std::vector<int32_t> point_data;
  CSM_PARSE_LOOP(s) {
    if (field == 1 && type == 0) num_points = static_cast<int32_t>(r.Varint());
    else if (field == 3) wire::RepeatedVarint(&r, type, [&](uint64_t v) { point_data.push_back(static_cast<int32_t>(v)); });
    else r.Skip(type);
  }
  CSM_CHECK(r.ok() && num_points >= 0);
  return sensor::CompressedPointCloud(static_cast<size_t>(num_points), std::move(point_data)).Decompress();
}

// trajectory_node_data.proto, rebuilt as mapping::FromProto does (trajectory_node.cc:50-69).
void ParseTrajectoryNodeData(const Span& s, NodeProto* out) {
  out->node_data.gravity_alignment = transform::Quaterniond(0., 0., 0., 0.);
  out->local_pose = transform::Rigid3d(transform::Vector3d(0., 0., 0.), transform::Quaterniond(0., 0., 0., 0.));
  CSM_PARSE_LOOP(s) {
    if (field == 1 && type == 0) out->timestamp = static_cast<int64_t>(r.Varint());
    else if (field == 2 && type == 2) out->node_data.gravity_alignment = ParseQuaterniond(r.Bytes());
    else if (field == 3 && type == 2) out->node_data.filtered_gravity_aligned_point_cloud = ParseCompressedPointCloud(r.Bytes());
    else if (field == 4 && type == 2) out->node_data.high_resolution_point_cloud = ParseCompressedPointCloud(r.Bytes());
    else if (field == 5 && type == 2) out->node_data.low_resolution_point_cloud = ParseCompressedPointCloud(r.Bytes());
    else if (field == 6) RepeatedFloat(&r, type, &out->node_data.rotational_scan_matcher_histogram);
    else if (field == 7 && type == 2) out->local_pose = ParseRigid3d(r.Bytes());
    else r.Skip(type);
  }
  CSM_CHECK(r.ok());
}
// serialization.proto Node {NodeId node_id = 1; TrajectoryNodeData node_data = 5}
void ParseNode(const Span& s, NodeProto* out) {
  *out = NodeProto();
  CSM_PARSE_LOOP(s) {
    if (field == 1 && type == 2) {
      const std::pair<int, int> id = ParseId(r.Bytes());
      out->node_id = mapping::NodeId(id.first, id.second);
    } else if (field == 5 && type == 2) {
      ParseTrajectoryNodeData(r.Bytes(), out);
    } else {
      r.Skip(type);
    }
  }
  CSM_CHECK(r.ok());
}

// submap.proto Submap2D {Rigid3d local_pose = 1; int32 num_range_data = 2; bool finished = 3; Grid2D grid = 4}
void ParseSubmap2D(const Span& s, mapping::ValueConversionTables* tables, SubmapProto* out) {
  out->has_submap_2d = true;
  CSM_PARSE_LOOP(s) {
    if (field == 1 && type == 2) out->local_pose = ParseRigid3d(r.Bytes());
    else if (field == 2 && type == 0) out->num_range_data = static_cast<int32_t>(r.Varint());
    else if (field == 3 && type == 0) out->finished = r.Varint() != 0;
    else if (field == 4 && type == 2) {
      const Span grid = r.Bytes();
      // submap_2d.cc:82-91: a grid of neither kind is LOG(FATAL); ReadGrid2DProto aborts on it.
      if (r.ok()) out->grid = ReadGrid2DProto(grid.data, grid.size, tables);
    } else {
      r.Skip(type);
    }
  }
  CSM_CHECK(r.ok());
}
// submap.proto Submap3D {local_pose = 1; num_range_data = 2; finished = 3; HybridGrid high = 4, low = 5;
// repeated float rotational_scan_matcher_histogram = 6}; submap_3d.cc:290-307.
void ParseSubmap3D(const Span& s, SubmapProto* out) {
  out->has_submap_3d = true;
  CSM_PARSE_LOOP(s) {
    if (field == 1 && type == 2) out->local_pose = ParseRigid3d(r.Bytes());
    else if (field == 2 && type == 0) out->num_range_data = static_cast<int32_t>(r.Varint());
    else if (field == 3 && type == 0) out->finished = r.Varint() != 0;
    else if ((field == 4 || field == 5) && type == 2) {
      const Span grid = r.Bytes();
      if (r.ok())
        (field == 4 ? out->high_resolution_hybrid_grid : out->low_resolution_hybrid_grid) =
            ReadHybridGridProto(grid.data, grid.size);
    } else if (field == 6) {
      RepeatedFloat(&r, type, &out->rotational_scan_matcher_histogram);
    } else {
      r.Skip(type);
    }
  }
  CSM_CHECK(r.ok());
}
// serialization.proto Submap {SubmapId submap_id = 1; Submap2D submap_2d = 2; Submap3D submap_3d = 3}
void ParseSubmap(const Span& s, mapping::ValueConversionTables* tables, SubmapProto* out) {
  *out = SubmapProto();
  out->local_pose = transform::Rigid3d(transform::Vector3d(0., 0., 0.), transform::Quaterniond(0., 0., 0., 0.));
  CSM_PARSE_LOOP(s) {
    if (field == 1 && type == 2) {
      const std::pair<int, int> id = ParseId(r.Bytes());
      out->submap_id = mapping::SubmapId(id.first, id.second);
    } else if (field == 2 && type == 2) {
      ParseSubmap2D(r.Bytes(), tables, out);
    } else if (field == 3 && type == 2) {
      ParseSubmap3D(r.Bytes(), out);
    } else {
      r.Skip(type);
    }
  }
  CSM_CHECK(r.ok());
}

// pose_graph.proto PoseGraph.Constraint {submap_id = 1; node_id = 2; Rigid3d relative_pose = 3;
// Tag tag = 5; double translation_weight = 6; double rotation_weight = 7}
mapping::PoseGraphInterface::Constraint ParseConstraint(const Span& s) {
  using Constraint = mapping::PoseGraphInterface::Constraint;
  Constraint c{mapping::SubmapId(0, 0), mapping::NodeId(0, 0),
               Constraint::Pose{transform::Rigid3d(transform::Vector3d(0., 0., 0.), transform::Quaterniond(0., 0., 0., 0.)), 0., 0.},
               Constraint::INTRA_SUBMAP};
  CSM_PARSE_LOOP(s) {
    if (field == 1 && type == 2) {
      const std::pair<int, int> id = ParseId(r.Bytes());
      c.submap_id = mapping::SubmapId(id.first, id.second);
    } else if (field == 2 && type == 2) {
      const std::pair<int, int> id = ParseId(r.Bytes());
      c.node_id = mapping::NodeId(id.first, id.second);
    } else if (field == 3 && type == 2) {
      c.pose.zbar_ij = ParseRigid3d(r.Bytes());
    } else if (field == 5 && type == 0) {
      c.tag = r.Varint() == 1 ? Constraint::INTER_SUBMAP : Constraint::INTRA_SUBMAP;
    } else if (field == 6 && type == 1) {
      c.pose.translation_weight = r.Double();
    } else if (field == 7 && type == 1) {
      c.pose.rotation_weight = r.Double();
    } else {
      r.Skip(type);
    }
  }
  CSM_CHECK(r.ok());
  return c;
}
// trajectory.proto Trajectory.Node {int64 timestamp = 1; Rigid3d pose = 5; int32 node_index = 7}
PoseGraphProto::Trajectory::Node ParseTrajectoryNode(const Span& s) {
  PoseGraphProto::Trajectory::Node n;
  n.pose = transform::Rigid3d(transform::Vector3d(0., 0., 0.), transform::Quaterniond(0., 0., 0., 0.));
  CSM_PARSE_LOOP(s) {
    if (field == 1 && type == 0) n.timestamp = static_cast<int64_t>(r.Varint());
    else if (field == 5 && type == 2) n.pose = ParseRigid3d(r.Bytes());
    else if (field == 7 && type == 0) n.node_index = static_cast<int32_t>(r.Varint());
    else r.Skip(type);
  }
  CSM_CHECK(r.ok());
  return n;
}
// trajectory.proto Trajectory.Submap {Rigid3d pose = 1; int32 submap_index = 2}
PoseGraphProto::Trajectory::Submap ParseTrajectorySubmap(const Span& s) {
  PoseGraphProto::Trajectory::Submap m;
  m.pose = transform::Rigid3d(transform::Vector3d(0., 0., 0.), transform::Quaterniond(0., 0., 0., 0.));
  CSM_PARSE_LOOP(s) {
    if (field == 1 && type == 2) m.pose = ParseRigid3d(r.Bytes());
    else if (field == 2 && type == 0) m.submap_index = static_cast<int32_t>(r.Varint());
    else r.Skip(type);
  }
  CSM_CHECK(r.ok());
  return m;
}
// trajectory.proto Trajectory {repeated Node node = 1; repeated Submap submap = 2; int32 trajectory_id = 3}
PoseGraphProto::Trajectory ParseTrajectory(const Span& s) {
  PoseGraphProto::Trajectory t;
  CSM_PARSE_LOOP(s) {
    if (field == 1 && type == 2) t.node.push_back(ParseTrajectoryNode(r.Bytes()));
    else if (field == 2 && type == 2) t.submap.push_back(ParseTrajectorySubmap(r.Bytes()));
    else if (field == 3 && type == 0) t.trajectory_id = static_cast<int32_t>(r.Varint());
    else r.Skip(type);
  }
  CSM_CHECK(r.ok());
  return t;
}
// pose_graph.proto PoseGraph {repeated Constraint constraint = 2; repeated Trajectory trajectory = 4;
// repeated LandmarkPose landmark_poses = 5 (skipped)}
void ParsePoseGraph(const Span& s, PoseGraphProto* out) {
  *out = PoseGraphProto();
  CSM_PARSE_LOOP(s) {
    if (field == 2 && type == 2) out->constraint.push_back(ParseConstraint(r.Bytes()));
    else if (field == 4 && type == 2) out->trajectory.push_back(ParseTrajectory(r.Bytes()));
    else r.Skip(type);
  }
  CSM_CHECK(r.ok());
}
// trajectory_builder_options.proto AllTrajectoryBuilderOptions {repeated ... options_with_sensor_ids = 1}:
// only the count is read (proto_stream_deserializer.cc:68-70).
int CountOptionsWithSensorIds(const Span& s) {
  int count = 0;
  CSM_PARSE_LOOP(s) {
    if (field == 1 && type == 2) ++count;
    r.Skip(type);
  }
  CSM_CHECK(r.ok());
  return count;
}

#undef CSM_PARSE_LOOP

}  // namespace

void ParseSerializedData(const std::string& serialized, mapping::ValueConversionTables* conversion_tables,
                         SerializedData* out) {
  out->data_case_ = SerializedData::DATA_NOT_SET;
  Reader r(serialized.data(), serialized.size());
  int field = 0, type = 0;
  while (r.Next(&field, &type)) {
    // A oneof: the last member on the wire wins, as in protobuf's parser.
    if (type == 2 && field >= SerializedData::kPoseGraph && field <= SerializedData::kLandmarkData) {
      const Span body = r.Bytes();
      if (!r.ok()) break;
      out->data_case_ = static_cast<SerializedData::DataCase>(field);
      switch (field) {
        case SerializedData::kPoseGraph: ParsePoseGraph(body, &out->pose_graph_); break;
        case SerializedData::kAllTrajectoryBuilderOptions:
          out->options_with_sensor_ids_size_ = CountOptionsWithSensorIds(body);
          break;
        case SerializedData::kSubmap: ParseSubmap(body, conversion_tables, &out->submap_); break;
        case SerializedData::kNode: ParseNode(body, &out->node_); break;
        default: break;  // sensor and trajectory data: not read by the scan-matching path
      }
    } else {
      r.Skip(type);
    }
  }
  CSM_CHECK(r.ok());
}

ProtoStreamDeserializer::ProtoStreamDeserializer(ProtoStreamReader* const reader,
                                                 mapping::ValueConversionTables* const conversion_tables)
    : reader_(reader), conversion_tables_(conversion_tables) {
  // proto_stream_deserializer.cc:27-32: SerializationHeader {uint32 format_version = 1}
  std::string header;
  CSM_CHECK(reader_->Read(&header) && "Failed to read SerializationHeader.");
  {
    Reader r(header.data(), header.size());
    int field = 0, type = 0;
    while (r.Next(&field, &type)) {
      if (field == 1 && type == 0) format_version_ = static_cast<uint32_t>(r.Varint());
      else r.Skip(type);
    }
    CSM_CHECK(r.ok());
  }
  // :34-37,53-54
  CSM_CHECK((format_version_ == static_cast<uint32_t>(kMappingStateSerializationFormatVersion) ||
             format_version_ == static_cast<uint32_t>(kFormatVersionWithoutSubmapHistograms)) &&
            "Unsupported serialization format");
  // :56-70
  CSM_CHECK(ReadNextSerializedData(&pose_graph_, &raw_pose_graph_) && "Serialized stream misses PoseGraph.");
  CSM_CHECK(pose_graph_.has_pose_graph() && "Serialized stream order corrupt. Expecting `PoseGraph` after `SerializationHeader`");
  CSM_CHECK(ReadNextSerializedData(&all_trajectory_builder_options_, &raw_options_) &&
            "Serialized stream misses `AllTrajectoryBuilderOptions`.");
  CSM_CHECK(all_trajectory_builder_options_.has_all_trajectory_builder_options() &&
            "Serialized stream order corrupt. Expecting `AllTrajectoryBuilderOptions` after PoseGraph");
  CSM_CHECK(static_cast<int>(pose_graph_.pose_graph().trajectory.size()) ==
            all_trajectory_builder_options_.options_with_sensor_ids_size());
}

bool ProtoStreamDeserializer::ReadNextSerializedData(SerializedData* data, std::string* raw) {
  std::string serialized;
  if (!reader_->Read(&serialized)) return false;
  ParseSerializedData(serialized, conversion_tables_, data);
  if (raw != nullptr) raw->swap(serialized);
  return true;
}

MappingState LoadState(ProtoStreamReader* const reader, const bool expect_3d) {
  MappingState state;
  state.conversion_tables.reset(new mapping::ValueConversionTables);
  ProtoStreamDeserializer deserializer(reader, state.conversion_tables.get());
  state.format_version = deserializer.format_version();
  const PoseGraphProto& pose_graph = deserializer.pose_graph();
  state.constraints = pose_graph.constraint;
  state.raw_pose_graph_record = deserializer.raw_pose_graph_record();
  state.raw_options_record = deserializer.raw_options_record();
  // map_builder.cc:262-281 (trajectory ids are kept: there is one map builder state here, no remapping)
  std::map<mapping::SubmapId, transform::Rigid3d> submap_poses;
  std::map<mapping::NodeId, transform::Rigid3d> node_poses;
  for (const PoseGraphProto::Trajectory& trajectory : pose_graph.trajectory) {
    for (const PoseGraphProto::Trajectory::Submap& submap : trajectory.submap)
      submap_poses.emplace(mapping::SubmapId(trajectory.trajectory_id, submap.submap_index), submap.pose);
    for (const PoseGraphProto::Trajectory::Node& node : trajectory.node)
      node_poses.emplace(mapping::NodeId(trajectory.trajectory_id, node.node_index), node.pose);
  }
  // :290-296
  if (expect_3d) CSM_CHECK(state.format_version != static_cast<uint32_t>(kFormatVersionWithoutSubmapHistograms));
  SerializedData proto;
  std::string raw;
  while (deserializer.ReadNextSerializedData(&proto, &raw)) {
    MappingState::Record record;
    record.data_case = proto.data_case();
    switch (proto.data_case()) {
      case SerializedData::kPoseGraph:
        std::fprintf(stderr, "Found multiple serialized `PoseGraph`. Serialized stream likely corrupt!.\n");
        break;
      case SerializedData::kAllTrajectoryBuilderOptions:
        std::fprintf(stderr,
                     "Found multiple serialized `AllTrajectoryBuilderOptions`. Serialized stream likely corrupt!.\n");
        break;
      case SerializedData::kSubmap: {
        const auto pose = submap_poses.find(proto.submap().submap_id);
        CSM_CHECK(pose != submap_poses.end());
        record.index = static_cast<int>(state.submaps.size());
        state.submaps.push_back(MappingState::Submap{pose->second, std::move(*proto.mutable_submap())});
        break;
      }
      case SerializedData::kNode: {
        const auto pose = node_poses.find(proto.node().node_id);
        CSM_CHECK(pose != node_poses.end());
        record.index = static_cast<int>(state.nodes.size());
        state.nodes.push_back(MappingState::Node{pose->second, std::move(*proto.mutable_node())});
        break;
      }
      default:
        break;
    }
    if (record.index < 0) record.raw.swap(raw);  // everything but submaps and nodes is kept as read
    state.records.push_back(std::move(record));
  }
  // map_builder.cc:393: CHECK(reader->eof())
  CSM_CHECK(reader->eof());
  return state;
}

MappingState LoadStateFromFile(const std::string& state_filename, const bool expect_3d) {
  ProtoStreamReader stream(state_filename);
  return LoadState(&stream, expect_3d);
}

}  // namespace io
}  // namespace cartographer
