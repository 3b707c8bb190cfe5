// Flat C entry points over the .pbstream readers (cartographer/io/proto_stream*.h), for ctypes:
// a saved mapping state -> grid handles, poses and point clouds the matcher entry points of
// host_capi.cc / host_capi_3d.cc take.  Product code: never touches oracle/.
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "cartographer/io/grid_proto_writer.h"
#include "cartographer/io/proto_stream.h"
#include "cartographer/io/proto_stream_deserializer.h"
#include "host_capi_holders.h"

using namespace cartographer;

namespace {

using StateHandle = std::shared_ptr<io::MappingState>;

void StorePose(const transform::Rigid3d& pose, double* out7) {
  out7[0] = pose.translation().x();
  out7[1] = pose.translation().y();
  out7[2] = pose.translation().z();
  out7[3] = pose.rotation().w();
  out7[4] = pose.rotation().x();
  out7[5] = pose.rotation().y();
  out7[6] = pose.rotation().z();
}

const sensor::PointCloud& NodeCloud(const io::NodeProto& node, int which) {
  return which == 0 ? node.node_data.filtered_gravity_aligned_point_cloud
                    : (which == 1 ? node.node_data.high_resolution_point_cloud
                                  : node.node_data.low_resolution_point_cloud);
}

}  // namespace

extern "C" {

// The container alone: `count` serialized messages (concatenated in `bytes`, message i =
// [offsets[i], offsets[i + 1])) -> a .pbstream file written by io::ProtoStreamWriter.
int csmh_pbstream_write(const char* filename, int count, const uint8_t* bytes, const int64_t* offsets) {
  io::ProtoStreamWriter writer(filename);
  for (int i = 0; i < count; ++i)
    writer.Write(std::string(reinterpret_cast<const char*>(bytes) + offsets[i],
                             static_cast<size_t>(offsets[i + 1] - offsets[i])));
  return writer.Close() ? 0 : -1;
}
// Every message of a .pbstream, decompressed, concatenated into `bytes` (capacity `cap`) with
// offsets[count + 1]; returns the message count, or -(bytes needed) - 1 when `cap` (or
// `max_count`) is too small.  Call with cap = 0 to size.
int64_t csmh_pbstream_read(const char* filename, uint8_t* bytes, int64_t cap, int64_t* offsets, int max_count) {
  io::ProtoStreamReader reader(filename);
  std::vector<std::string> messages;
  std::string message;
  int64_t total = 0;
  while (reader.Read(&message)) {
    total += static_cast<int64_t>(message.size());
    messages.push_back(message);
  }
  if (!reader.eof()) return INT64_MIN;  // a record that ends before its size says so
  if (total > cap || static_cast<int64_t>(messages.size()) > max_count) return -total - 1;
  int64_t at = 0;
  for (size_t i = 0; i < messages.size(); ++i) {
    offsets[i] = at;
    std::memcpy(bytes + at, messages[i].data(), messages[i].size());
    at += static_cast<int64_t>(messages[i].size());
  }
  offsets[messages.size()] = at;
  return static_cast<int64_t>(messages.size());
}
int64_t csmh_pbstream_count(const char* filename) {
  io::ProtoStreamReader reader(filename);
  std::string message;
  int64_t count = 0;
  while (reader.Read(&message)) ++count;
  return reader.eof() ? count : INT64_MIN;
}

// A saved mapping state.  counts4 = {submaps, nodes, constraints, format version}.
void* csmh_state_load(const char* filename, int expect_3d, int32_t* counts4) {
  StateHandle* h = new StateHandle(new io::MappingState(io::LoadStateFromFile(filename, expect_3d != 0)));
  counts4[0] = static_cast<int32_t>((*h)->submaps.size());
  counts4[1] = static_cast<int32_t>((*h)->nodes.size());
  counts4[2] = static_cast<int32_t>((*h)->constraints.size());
  counts4[3] = static_cast<int32_t>((*h)->format_version);
  return h;
}
void csmh_state_destroy(void* st) { delete static_cast<StateHandle*>(st); }

// ints8 = {trajectory id, submap index, has submap_2d, has submap_3d, num_range_data, finished,
// grids present (2D: 1 if a grid; 3D: bit 0 high, bit 1 low), histogram size};
// poses14 = local pose, global pose, each {tx, ty, tz, qw, qx, qy, qz}.
void csmh_state_submap(void* st, int i, int32_t* ints8, double* poses14) {
  const io::MappingState::Submap& s = (**static_cast<StateHandle*>(st)).submaps[i];
  ints8[0] = s.proto.submap_id.trajectory_id;
  ints8[1] = s.proto.submap_id.submap_index;
  ints8[2] = s.proto.has_submap_2d;
  ints8[3] = s.proto.has_submap_3d;
  ints8[4] = s.proto.num_range_data;
  ints8[5] = s.proto.finished;
  ints8[6] = s.proto.has_submap_3d ? ((s.proto.high_resolution_hybrid_grid ? 1 : 0) |
                                      (s.proto.low_resolution_hybrid_grid ? 2 : 0))
                                   : (s.proto.grid ? 1 : 0);
  ints8[7] = static_cast<int32_t>(s.proto.rotational_scan_matcher_histogram.size());
  StorePose(s.proto.local_pose, poses14);
  StorePose(s.global_pose, poses14 + 7);
}
// The grid of 2D submap i as a grid handle (csmh_grid_destroy frees it; the state lives as long
// as any of its grid handles).  NULL when the submap has none.  info6 as csmh_grid_from_proto.
void* csmh_state_submap_grid(void* st, int i, double* info6) {
  const StateHandle& state = *static_cast<StateHandle*>(st);
  mapping::Grid2D* grid = state->submaps[i].proto.grid.get();
  if (grid == nullptr) return nullptr;
  GridHolder* h = new GridHolder;
  h->grid = std::shared_ptr<mapping::Grid2D>(state, grid);
  const mapping::MapLimits& l = grid->limits();
  info6[0] = l.cell_limits().num_x_cells;
  info6[1] = l.cell_limits().num_y_cells;
  info6[2] = l.resolution();
  info6[3] = l.max().x();
  info6[4] = l.max().y();
  info6[5] = grid->GetGridType() == mapping::GridType::TSDF ? 1. : 0.;
  return h;
}
// Non-default voxels of the high (which = 0) or low (1) resolution grid of 3D submap i; xyz /
// values may be NULL to count.  Returns -1 when the submap has no such grid.
int csmh_state_submap_hybrid_grid(void* st, int i, int which, float* resolution, int* grid_size, int32_t* xyz,
                                  uint16_t* values) {
  const io::SubmapProto& s = (**static_cast<StateHandle*>(st)).submaps[i].proto;
  const mapping::HybridGrid* grid =
      which == 0 ? s.high_resolution_hybrid_grid.get() : s.low_resolution_hybrid_grid.get();
  if (grid == nullptr) return -1;
  *resolution = grid->resolution();
  *grid_size = grid->grid_size();
  std::vector<int32_t> cells;
  std::vector<uint16_t> vals;
  grid->Export(&cells, &vals);
  if (xyz != nullptr) {
    std::memcpy(xyz, cells.data(), sizeof(int32_t) * cells.size());
    std::memcpy(values, vals.data(), sizeof(uint16_t) * vals.size());
  }
  return static_cast<int>(vals.size());
}
void csmh_state_submap_histogram(void* st, int i, float* out) {
  const std::vector<float>& h = (**static_cast<StateHandle*>(st)).submaps[i].proto.rotational_scan_matcher_histogram;
  std::memcpy(out, h.data(), sizeof(float) * h.size());
}

// ints8 = {trajectory id, node index, points in filtered_gravity_aligned / high_resolution /
// low_resolution point cloud, histogram size, 0, 0}; doubles18 = gravity_alignment {w, x, y, z},
// local pose, global pose (7 each).
void csmh_state_node(void* st, int i, int32_t* ints8, int64_t* timestamp, double* doubles18) {
  const io::MappingState::Node& n = (**static_cast<StateHandle*>(st)).nodes[i];
  ints8[0] = n.proto.node_id.trajectory_id;
  ints8[1] = n.proto.node_id.node_index;
  for (int which = 0; which < 3; ++which) ints8[2 + which] = static_cast<int32_t>(NodeCloud(n.proto, which).size());
  ints8[5] = static_cast<int32_t>(n.proto.node_data.rotational_scan_matcher_histogram.size());
  ints8[6] = ints8[7] = 0;
  *timestamp = n.proto.timestamp;
  const transform::Quaterniond& g = n.proto.node_data.gravity_alignment;
  doubles18[0] = g.w();
  doubles18[1] = g.x();
  doubles18[2] = g.y();
  doubles18[3] = g.z();
  StorePose(n.proto.local_pose, doubles18 + 4);
  StorePose(n.global_pose, doubles18 + 11);
}
void csmh_state_node_cloud(void* st, int i, int which, float* xyz) {
  const sensor::PointCloud& cloud = NodeCloud((**static_cast<StateHandle*>(st)).nodes[i].proto, which);
  for (size_t p = 0; p < cloud.size(); ++p) {
    xyz[3 * p] = cloud[p].position.x();
    xyz[3 * p + 1] = cloud[p].position.y();
    xyz[3 * p + 2] = cloud[p].position.z();
  }
}
void csmh_state_node_histogram(void* st, int i, float* out) {
  const std::vector<float>& h = (**static_cast<StateHandle*>(st)).nodes[i].proto.node_data.rotational_scan_matcher_histogram;
  std::memcpy(out, h.data(), sizeof(float) * h.size());
}
// Per constraint: ints5 = {submap trajectory, submap index, node trajectory, node index, tag};
// doubles9 = relative pose {tx, ty, tz, qw, qx, qy, qz}, translation weight, rotation weight.
void csmh_state_constraints(void* st, int32_t* ints5, double* doubles9) {
  const std::vector<mapping::PoseGraphInterface::Constraint>& cs = (**static_cast<StateHandle*>(st)).constraints;
  for (size_t c = 0; c < cs.size(); ++c) {
    ints5[5 * c] = cs[c].submap_id.trajectory_id;
    ints5[5 * c + 1] = cs[c].submap_id.submap_index;
    ints5[5 * c + 2] = cs[c].node_id.trajectory_id;
    ints5[5 * c + 3] = cs[c].node_id.node_index;
    ints5[5 * c + 4] = cs[c].tag == mapping::PoseGraphInterface::Constraint::INTER_SUBMAP ? 1 : 0;
    StorePose(cs[c].pose.zbar_ij, doubles9 + 9 * c);
    doubles9[9 * c + 7] = cs[c].pose.translation_weight;
    doubles9[9 * c + 8] = cs[c].pose.rotation_weight;
  }
}

// ---- writing (cartographer/io/grid_proto_writer.h) ----
static int64_t CopyOut(const std::string& bytes, uint8_t* out, int64_t cap) {
  if (out != nullptr) {
    if (static_cast<int64_t>(bytes.size()) > cap) return -1;
    std::memcpy(out, bytes.data(), bytes.size());
  }
  return static_cast<int64_t>(bytes.size());
}
// A grid handle -> its serialized proto::Grid2D (out may be NULL to size).
int64_t csmh_grid_to_proto(void* grid, uint8_t* out, int64_t cap) {
  return CopyOut(io::WriteGrid2DProto(*static_cast<GridHolder*>(grid)->grid), out, cap);
}
// A voxel list -> the serialized proto::HybridGrid of the HybridGrid holding it.
int64_t csmh_hybrid_grid_to_proto(float resolution, const int32_t* xyz, const uint16_t* values, int n, uint8_t* out,
                                  int64_t cap) {
  mapping::HybridGrid grid(resolution);
  for (int i = 0; i < n; ++i) *grid.mutable_value(Eigen::Array3i(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2])) = values[i];
  return CopyOut(io::WriteHybridGridProto(grid), out, cap);
}
// A serialized SerializedData{pose_graph} -> parsed and written again (WritePoseGraphRecord).
int64_t csmh_reencode_pose_graph(const uint8_t* bytes, int64_t size, uint8_t* out, int64_t cap) {
  mapping::ValueConversionTables tables;
  io::SerializedData data;
  io::ParseSerializedData(std::string(reinterpret_cast<const char*>(bytes), static_cast<size_t>(size)), &tables, &data);
  if (!data.has_pose_graph()) return -2;
  return CopyOut(io::WritePoseGraphRecord(data.pose_graph()), out, cap);
}
// The loaded state back into a .pbstream (WriteStateToFile).
void csmh_state_write(void* st, const char* filename) { io::WriteStateToFile(**static_cast<StateHandle*>(st), filename); }
// Replaces the cells of 2D submap i's grid (same limits), e.g. with cells that were updated on the device.
int csmh_state_set_submap_cells(void* st, int i, const uint16_t* cells, int64_t count) {
  mapping::Grid2D* grid = (**static_cast<StateHandle*>(st)).submaps[i].proto.grid.get();
  if (grid == nullptr || static_cast<int64_t>(grid->correspondence_cost_cells().size()) != count) return -1;
  std::memcpy(grid->mutable_correspondence_cost_cells()->data(), cells, sizeof(uint16_t) * static_cast<size_t>(count));
  return 0;
}

}  // extern "C"
