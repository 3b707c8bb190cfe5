// Grid ingestion: serialized proto::Grid2D / proto::HybridGrid messages ->
// the grids the device path takes.  Mirrors the reference's proto constructors
//   Grid2D::Grid2D(const proto::Grid2D&, ValueConversionTables*)      mapping/2d/grid_2d.cc:22-96
//   ProbabilityGrid(const proto::Grid2D&, ...)                         mapping/2d/probability_grid.cc:33-37
//   TSDF2D(const proto::Grid2D&, ...)                                  mapping/internal/2d/tsdf_2d.cc:35-47
//   MapLimits(const proto::MapLimits&), CellLimits(const proto::CellLimits&)
//                                                    mapping/2d/map_limits.h:50-55, xy_index.h:39-41
//   HybridGrid(const proto::HybridGrid&)                               mapping/3d/hybrid_grid.h:473-485
// on the wire format itself (io/proto_wire.h), with the reference's CHECKs as aborts.  The bytes
// are what `grid.ToProto().SerializeAsString()` produces in a Cartographer process, i.e. the
// `grid` field of a proto::Submap2D / the `high_resolution_hybrid_grid` of a proto::Submap3D
// inside a .pbstream; the stream container and the records around the grids are read by
// io/proto_stream.h and io/proto_stream_deserializer.h.
#ifndef CARTOGRAPHER_IO_GRID_PROTO_READER_H_
#define CARTOGRAPHER_IO_GRID_PROTO_READER_H_

#include <memory>
#include <string>

#include "cartographer_compat.h"
#include "cartographer_compat_3d.h"

namespace cartographer {
namespace io {

// A ProbabilityGrid or TSDF2D, whichever the message's oneof holds.
std::unique_ptr<mapping::Grid2D> ReadGrid2DProto(const void* data, size_t size,
                                                 mapping::ValueConversionTables* conversion_tables);
inline std::unique_ptr<mapping::Grid2D> ReadGrid2DProto(const std::string& serialized,
                                                        mapping::ValueConversionTables* conversion_tables) {
  return ReadGrid2DProto(serialized.data(), serialized.size(), conversion_tables);
}

std::unique_ptr<mapping::HybridGrid> ReadHybridGridProto(const void* data, size_t size);
inline std::unique_ptr<mapping::HybridGrid> ReadHybridGridProto(const std::string& serialized) {
  return ReadHybridGridProto(serialized.data(), serialized.size());
}

}  // namespace io
}  // namespace cartographer

#endif  // CARTOGRAPHER_IO_GRID_PROTO_READER_H_
