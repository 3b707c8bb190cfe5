// The writing half of grid ingestion and of the saved-state records: what
//   Grid2D::ToProto / ProbabilityGrid::ToProto / TSDF2D::ToProto   mapping/2d/grid_2d.cc:166-183,
//                                                mapping/2d/probability_grid.cc:84-89, mapping/internal/2d/tsdf_2d.cc:107-116
//   ToProto(MapLimits), ToProto(CellLimits)      mapping/2d/map_limits.h:98-104, xy_index.h:47-52
//   HybridGrid::ToProto                          mapping/3d/hybrid_grid.h:528-540
//   Submap2D::ToProto / Submap3D::ToProto        mapping/2d/submap_2d.cc:96-110, mapping/3d/submap_3d.cc:264-283
//   ToProto(TrajectoryNode::Data)                mapping/trajectory_node.cc:27-48
// followed by SerializeAsString() produce in a Cartographer process, written here straight in the
// wire format (io/proto_wire.h) in libprotobuf's canonical form, so that a grid born or updated on
// the device can go back into a .pbstream the reference loads.  The voxels of a HybridGrid leave in
// the order the reference's iterator visits them (meta cells, bricks, cells, each z-major).
#ifndef CARTOGRAPHER_IO_GRID_PROTO_WRITER_H_
#define CARTOGRAPHER_IO_GRID_PROTO_WRITER_H_

#include <string>

#include "cartographer/io/proto_stream_deserializer.h"
#include "cartographer_compat.h"
#include "cartographer_compat_3d.h"

namespace cartographer {
namespace io {

// A serialized proto::Grid2D.  known_cells_box is the bounding box of the cells that are not
// unknown (value 0), which is what the reference's box has grown to after the same updates.
std::string WriteGrid2DProto(const mapping::Grid2D& grid);
// A serialized proto::HybridGrid.
std::string WriteHybridGridProto(const mapping::HybridGrid& grid);
// Serialized SerializedData records (one oneof member each).
std::string WriteSubmapRecord(const SubmapProto& submap);
std::string WriteNodeRecord(const NodeProto& node);
std::string WritePoseGraphRecord(const PoseGraphProto& pose_graph);
std::string WriteSerializationHeader(uint32_t format_version);

// The state back into a .pbstream: header, the pose graph and builder-options records as they were
// read, then every record in its original place -- submap and node records re-encoded from `state`
// (so a replaced grid or cloud is what gets saved), all others passed through unchanged.
void WriteStateToFile(const MappingState& state, const std::string& filename);

}  // namespace io
}  // namespace cartographer

#endif  // CARTOGRAPHER_IO_GRID_PROTO_WRITER_H_
