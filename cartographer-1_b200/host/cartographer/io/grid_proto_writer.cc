#include "cartographer/io/grid_proto_writer.h"

#include <algorithm>
#include <limits>
#include <vector>

#include "cartographer/io/proto_wire.h"
#include "cartographer/sensor/compressed_point_cloud.h"

namespace cartographer {
namespace io {
namespace {

using wire::Writer;

// transform.proto Vector3d {x = 1, y = 2, z = 3}, Quaterniond {x = 1, y = 2, z = 3, w = 4},
// Rigid3d {translation = 1, rotation = 2} (transform.cc:48-74: both members are always set)
std::string Rigid3dBytes(const transform::Rigid3d& pose) {
  Writer t, q, out;
  t.Double(1, pose.translation().x());
  t.Double(2, pose.translation().y());
  t.Double(3, pose.translation().z());
  q.Double(1, pose.rotation().x());
  q.Double(2, pose.rotation().y());
  q.Double(3, pose.rotation().z());
  q.Double(4, pose.rotation().w());
  out.Message(1, t.str());
  out.Message(2, q.str());
  return out.str();
}
std::string QuaterniondBytes(const transform::Quaterniond& rotation) {
  Writer q;
  q.Double(1, rotation.x());
  q.Double(2, rotation.y());
  q.Double(3, rotation.z());
  q.Double(4, rotation.w());
  return q.str();
}
// pose_graph.proto SubmapId / NodeId {trajectory_id = 1, index = 2}
std::string IdBytes(const int trajectory_id, const int index) {
  Writer out;
  out.Int32(1, trajectory_id);
  out.Int32(2, index);
  return out.str();
}
// sensor.proto CompressedPointCloud {num_points = 1, repeated int32 point_data = 3}
std::string CompressedPointCloudBytes(const sensor::PointCloud& cloud) {
  const sensor::CompressedPointCloud compressed(cloud);
  const std::vector<int32_t>& words = compressed.point_data();
  Writer out;
  out.Int32(1, static_cast<int32_t>(compressed.size()));
  out.PackedInt32(3, words.size(), [&](size_t i) { return words[i]; });
  return out.str();
}

// The place of a voxel in the reference's iteration (hybrid_grid.h:40-52,255-345): the grid of
// `grid_size` voxels per axis is 2^bits meta cells of 64^3 per axis, a meta cell 8^3 bricks of 8^3
// voxels; every level is walked in flat z-major order.
uint64_t IterationKey(const int32_t x, const int32_t y, const int32_t z, const int grid_size) {
  int meta_bits = 0;
  while ((64 << meta_bits) < grid_size) ++meta_bits;
  const uint32_t u[3] = {static_cast<uint32_t>(x + (grid_size >> 1)), static_cast<uint32_t>(y + (grid_size >> 1)),
                         static_cast<uint32_t>(z + (grid_size >> 1))};
  auto flat = [](const uint32_t a, const uint32_t b, const uint32_t c, const int bits) {
    return static_cast<uint64_t>((((c << bits) + b) << bits) + a);
  };
  return (flat(u[0] >> 6, u[1] >> 6, u[2] >> 6, meta_bits) << 18) |
         (flat((u[0] >> 3) & 7, (u[1] >> 3) & 7, (u[2] >> 3) & 7, 3) << 9) | flat(u[0] & 7, u[1] & 7, u[2] & 7, 3);
}

}  // namespace

std::string WriteGrid2DProto(const mapping::Grid2D& grid) {
  const mapping::MapLimits& limits = grid.limits();
  const int num_x = limits.cell_limits().num_x_cells, num_y = limits.cell_limits().num_y_cells;
  const std::vector<uint16>& cells = grid.correspondence_cost_cells();
  Writer out;
  {  // limits = 1: MapLimits {resolution = 1, Vector2d max = 2, CellLimits cell_limits = 3}
    Writer max, cell_limits, l;
    max.Double(1, limits.max().x());
    max.Double(2, limits.max().y());
    cell_limits.Int32(1, num_x);
    cell_limits.Int32(2, num_y);
    l.Double(1, limits.resolution());
    l.Message(2, max.str());
    l.Message(3, cell_limits.str());
    out.Message(1, l.str());
  }
  out.PackedInt32(2, cells.size(), [&](size_t i) { return static_cast<int32_t>(cells[i]); });
  {  // known_cells_box = 3 {max_x = 1, max_y = 2, min_x = 3, min_y = 4}, only when some cell is known
    int min_x = std::numeric_limits<int>::max(), min_y = min_x, max_x = -1, max_y = -1;
    for (int y = 0; y < num_y; ++y)
      for (int x = 0; x < num_x; ++x)
        if (cells[static_cast<size_t>(y) * num_x + x] != 0) {
          min_x = std::min(min_x, x);
          max_x = std::max(max_x, x);
          min_y = std::min(min_y, y);
          max_y = std::max(max_y, y);
        }
    if (max_x >= 0) {
      Writer box;
      box.Int32(1, max_x);
      box.Int32(2, max_y);
      box.Int32(3, min_x);
      box.Int32(4, min_y);
      out.Message(3, box.str());
    }
  }
  if (grid.GetGridType() == mapping::GridType::TSDF) {
    // tsdf_2d = 5: TSDF2D {truncation_distance = 1, max_weight = 2, repeated int32 weight_cells = 3}
    const mapping::TSDF2D& tsdf = static_cast<const mapping::TSDF2D&>(grid);
    const std::vector<uint16>& weights = tsdf.weight_cells();
    Writer t;
    t.Float(1, tsdf.value_converter().getMaxTSD());
    t.Float(2, tsdf.value_converter().getMaxWeight());
    t.PackedInt32(3, weights.size(), [&](size_t i) { return static_cast<int32_t>(weights[i]); });
    out.Message(5, t.str());
  } else {
    out.Message(4, "");  // probability_grid_2d = 4: ProbabilityGrid {}
  }
  out.Float(6, grid.GetMinCorrespondenceCost());
  out.Float(7, grid.GetMaxCorrespondenceCost());
  return out.str();
}

std::string WriteHybridGridProto(const mapping::HybridGrid& grid) {
  std::vector<int32_t> xyz;
  std::vector<uint16> values;
  grid.Export(&xyz, &values);
  std::vector<size_t> order(values.size());
  std::vector<uint64_t> keys(values.size());
  for (size_t i = 0; i != values.size(); ++i) {
    order[i] = i;
    keys[i] = IterationKey(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], grid.grid_size());
  }
  std::sort(order.begin(), order.end(), [&](size_t a, size_t b) { return keys[a] < keys[b]; });
  // hybrid_grid.proto: float resolution = 1; repeated sint32 x / y / z_indices = 3 / 4 / 5; repeated int32 values = 6
  Writer out;
  out.Float(1, grid.resolution());
  out.PackedSInt32(3, order.size(), [&](size_t i) { return xyz[3 * order[i]]; });
  out.PackedSInt32(4, order.size(), [&](size_t i) { return xyz[3 * order[i] + 1]; });
  out.PackedSInt32(5, order.size(), [&](size_t i) { return xyz[3 * order[i] + 2]; });
  out.PackedInt32(6, order.size(), [&](size_t i) { return static_cast<int32_t>(values[order[i]]); });
  return out.str();
}

std::string WriteSubmapRecord(const SubmapProto& submap) {
  // serialization.proto Submap {SubmapId submap_id = 1; Submap2D submap_2d = 2; Submap3D submap_3d = 3}
  Writer s;
  s.Message(1, IdBytes(submap.submap_id.trajectory_id, submap.submap_id.submap_index));
  if (submap.has_submap_2d) {
    // submap.proto Submap2D {local_pose = 1; num_range_data = 2; finished = 3; Grid2D grid = 4}
    Writer m;
    m.Message(1, Rigid3dBytes(submap.local_pose));
    m.Int32(2, submap.num_range_data);
    m.Bool(3, submap.finished);
    if (submap.grid != nullptr) m.Message(4, WriteGrid2DProto(*submap.grid));
    s.Message(2, m.str());
  }
  if (submap.has_submap_3d) {
    // Submap3D {local_pose = 1; num_range_data = 2; finished = 3; HybridGrid high = 4, low = 5; repeated float
    // rotational_scan_matcher_histogram = 6}
    Writer m;
    m.Message(1, Rigid3dBytes(submap.local_pose));
    m.Int32(2, submap.num_range_data);
    m.Bool(3, submap.finished);
    if (submap.high_resolution_hybrid_grid != nullptr) m.Message(4, WriteHybridGridProto(*submap.high_resolution_hybrid_grid));
    if (submap.low_resolution_hybrid_grid != nullptr) m.Message(5, WriteHybridGridProto(*submap.low_resolution_hybrid_grid));
    const std::vector<float>& h = submap.rotational_scan_matcher_histogram;
    m.PackedFloat(6, h.size(), [&](size_t i) { return h[i]; });
    s.Message(3, m.str());
  }
  Writer out;
  out.Message(SerializedData::kSubmap, s.str());
  return out.str();
}

std::string WriteNodeRecord(const NodeProto& node) {
  // trajectory_node_data.proto (trajectory_node.cc:27-48: every member message is set)
  Writer d;
  d.Int64(1, node.timestamp);
  d.Message(2, QuaterniondBytes(node.node_data.gravity_alignment));
  d.Message(3, CompressedPointCloudBytes(node.node_data.filtered_gravity_aligned_point_cloud));
  d.Message(4, CompressedPointCloudBytes(node.node_data.high_resolution_point_cloud));
  d.Message(5, CompressedPointCloudBytes(node.node_data.low_resolution_point_cloud));
  const std::vector<float>& h = node.node_data.rotational_scan_matcher_histogram;
  d.PackedFloat(6, h.size(), [&](size_t i) { return h[i]; });
  d.Message(7, Rigid3dBytes(node.local_pose));
  // serialization.proto Node {NodeId node_id = 1; TrajectoryNodeData node_data = 5}
  Writer n;
  n.Message(1, IdBytes(node.node_id.trajectory_id, node.node_id.node_index));
  n.Message(5, d.str());
  Writer out;
  out.Message(SerializedData::kNode, n.str());
  return out.str();
}

std::string WritePoseGraphRecord(const PoseGraphProto& pose_graph) {
  Writer g;
  for (const mapping::PoseGraphInterface::Constraint& c : pose_graph.constraint) {
    // PoseGraph.Constraint {submap_id = 1; node_id = 2; relative_pose = 3; tag = 5; translation_weight = 6; rotation_weight = 7}
    Writer m;
    m.Message(1, IdBytes(c.submap_id.trajectory_id, c.submap_id.submap_index));
    m.Message(2, IdBytes(c.node_id.trajectory_id, c.node_id.node_index));
    m.Message(3, Rigid3dBytes(c.pose.zbar_ij));
    m.Int32(5, c.tag == mapping::PoseGraphInterface::Constraint::INTER_SUBMAP ? 1 : 0);
    m.Double(6, c.pose.translation_weight);
    m.Double(7, c.pose.rotation_weight);
    g.Message(2, m.str());
  }
  for (const PoseGraphProto::Trajectory& t : pose_graph.trajectory) {
    // Trajectory {repeated Node node = 1; repeated Submap submap = 2; trajectory_id = 3}
    Writer m;
    for (const PoseGraphProto::Trajectory::Node& n : t.node) {
      Writer e;  // Node {timestamp = 1; pose = 5; node_index = 7}
      e.Int64(1, n.timestamp);
      e.Message(5, Rigid3dBytes(n.pose));
      e.Int32(7, n.node_index);
      m.Message(1, e.str());
    }
    for (const PoseGraphProto::Trajectory::Submap& s : t.submap) {
      Writer e;  // Submap {pose = 1; submap_index = 2}
      e.Message(1, Rigid3dBytes(s.pose));
      e.Int32(2, s.submap_index);
      m.Message(2, e.str());
    }
    m.Int32(3, t.trajectory_id);
    g.Message(4, m.str());
  }
  Writer out;
  out.Message(SerializedData::kPoseGraph, g.str());
  return out.str();
}

std::string WriteSerializationHeader(const uint32_t format_version) {
  Writer out;
  out.UInt32(1, format_version);
  return out.str();
}

void WriteStateToFile(const MappingState& state, const std::string& filename) {
  ProtoStreamWriter writer(filename);
  writer.Write(WriteSerializationHeader(state.format_version));
  writer.Write(state.raw_pose_graph_record);
  writer.Write(state.raw_options_record);
  for (const MappingState::Record& record : state.records) {
    if (record.data_case == SerializedData::kSubmap && record.index >= 0) {
      writer.Write(WriteSubmapRecord(state.submaps[record.index].proto));
    } else if (record.data_case == SerializedData::kNode && record.index >= 0) {
      writer.Write(WriteNodeRecord(state.nodes[record.index].proto));
    } else {
      writer.Write(record.raw);
    }
  }
  CSM_CHECK(writer.Close());
}

}  // namespace io
}  // namespace cartographer
