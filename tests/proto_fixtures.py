"""Real protobuf wire bytes for the grid-ingestion tests: the reference's message definitions
(cartographer/mapping/proto/{grid_2d,map_limits,cell_limits_2d,probability_grid,tsdf_2d,hybrid_grid}.proto,
cartographer/transform/proto/transform.proto: same field names, numbers and types) declared through
descriptor_pb2 -- protoc is not in this image, the protobuf runtime is -- and serialized by the
protobuf library itself."""
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

_T = descriptor_pb2.FieldDescriptorProto


def _field(msg, name, number, ftype, label=_T.LABEL_OPTIONAL, type_name=None, oneof=None):
    f = msg.field.add()
    f.name, f.number, f.type, f.label = name, number, ftype, label
    if type_name:
        f.type_name = type_name
    if oneof is not None:
        f.oneof_index = oneof
    return f


def _build():
    fd = descriptor_pb2.FileDescriptorProto()
    fd.name = "csm_test_grids.proto"
    fd.package = "cartographer.mapping.proto"
    fd.syntax = "proto3"
    m = fd.message_type.add()
    m.name = "Vector2d"  # transform.proto
    _field(m, "x", 1, _T.TYPE_DOUBLE)
    _field(m, "y", 2, _T.TYPE_DOUBLE)
    m = fd.message_type.add()
    m.name = "CellLimits"
    _field(m, "num_x_cells", 1, _T.TYPE_INT32)
    _field(m, "num_y_cells", 2, _T.TYPE_INT32)
    m = fd.message_type.add()
    m.name = "MapLimits"
    _field(m, "resolution", 1, _T.TYPE_DOUBLE)
    _field(m, "max", 2, _T.TYPE_MESSAGE, type_name=".cartographer.mapping.proto.Vector2d")
    _field(m, "cell_limits", 3, _T.TYPE_MESSAGE, type_name=".cartographer.mapping.proto.CellLimits")
    m = fd.message_type.add()
    m.name = "ProbabilityGrid"
    m = fd.message_type.add()
    m.name = "TSDF2D"
    _field(m, "truncation_distance", 1, _T.TYPE_FLOAT)
    _field(m, "max_weight", 2, _T.TYPE_FLOAT)
    _field(m, "weight_cells", 3, _T.TYPE_INT32, label=_T.LABEL_REPEATED)
    m = fd.message_type.add()
    m.name = "Grid2D"
    box = m.nested_type.add()
    box.name = "CellBox"
    _field(box, "max_x", 1, _T.TYPE_INT32)
    _field(box, "max_y", 2, _T.TYPE_INT32)
    _field(box, "min_x", 3, _T.TYPE_INT32)
    _field(box, "min_y", 4, _T.TYPE_INT32)
    m.oneof_decl.add().name = "grid"
    _field(m, "limits", 1, _T.TYPE_MESSAGE, type_name=".cartographer.mapping.proto.MapLimits")
    _field(m, "cells", 2, _T.TYPE_INT32, label=_T.LABEL_REPEATED)
    _field(m, "known_cells_box", 3, _T.TYPE_MESSAGE, type_name=".cartographer.mapping.proto.Grid2D.CellBox")
    _field(m, "probability_grid_2d", 4, _T.TYPE_MESSAGE, type_name=".cartographer.mapping.proto.ProbabilityGrid", oneof=0)
    _field(m, "tsdf_2d", 5, _T.TYPE_MESSAGE, type_name=".cartographer.mapping.proto.TSDF2D", oneof=0)
    _field(m, "min_correspondence_cost", 6, _T.TYPE_FLOAT)
    _field(m, "max_correspondence_cost", 7, _T.TYPE_FLOAT)
    m = fd.message_type.add()
    m.name = "HybridGrid"
    _field(m, "resolution", 1, _T.TYPE_FLOAT)
    _field(m, "x_indices", 3, _T.TYPE_SINT32, label=_T.LABEL_REPEATED)
    _field(m, "y_indices", 4, _T.TYPE_SINT32, label=_T.LABEL_REPEATED)
    _field(m, "z_indices", 5, _T.TYPE_SINT32, label=_T.LABEL_REPEATED)
    _field(m, "values", 6, _T.TYPE_INT32, label=_T.LABEL_REPEATED)
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    get = lambda n: message_factory.GetMessageClass(pool.FindMessageTypeByName("cartographer.mapping.proto." + n))
    return get("Grid2D"), get("HybridGrid")


Grid2D, HybridGrid = _build()


def grid2d_bytes(cells, res, max_x, max_y, min_cost=0.1, max_cost=0.9, tsdf=None, box=None):
    """cells: uint16 [ny, nx]; tsdf = (weight_cells, truncation_distance, max_weight) for a TSDF2D."""
    g = Grid2D()
    g.limits.resolution = res
    g.limits.max.x, g.limits.max.y = max_x, max_y
    g.limits.cell_limits.num_x_cells, g.limits.cell_limits.num_y_cells = cells.shape[1], cells.shape[0]
    g.cells.extend(int(v) for v in cells.reshape(-1))
    if box is not None:
        g.known_cells_box.min_x, g.known_cells_box.min_y, g.known_cells_box.max_x, g.known_cells_box.max_y = box
    if tsdf is None:
        g.probability_grid_2d.SetInParent()
    else:
        g.tsdf_2d.truncation_distance, g.tsdf_2d.max_weight = tsdf[1], tsdf[2]
        g.tsdf_2d.weight_cells.extend(int(v) for v in tsdf[0].reshape(-1))
    g.min_correspondence_cost, g.max_correspondence_cost = min_cost, max_cost
    return g.SerializeToString()


def hybrid_grid_bytes(resolution, xyz, values):
    h = HybridGrid()
    h.resolution = resolution
    h.x_indices.extend(int(v) for v in xyz[:, 0])
    h.y_indices.extend(int(v) for v in xyz[:, 1])
    h.z_indices.extend(int(v) for v in xyz[:, 2])
    h.values.extend(int(v) for v in values)
    return h.SerializeToString()
