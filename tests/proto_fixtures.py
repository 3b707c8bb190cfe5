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


def _build_serialization(fd):
    """The messages of a saved mapping state: mapping/proto/{serialization,pose_graph,trajectory,submap,
    trajectory_node_data}.proto, sensor/proto/sensor.proto (CompressedPointCloud, ImuData),
    transform/proto/transform.proto -- same field names, numbers and types (the package differs,
    which the wire format does not carry)."""
    P = ".cartographer.mapping.proto."
    m = fd.message_type.add()
    m.name = "Vector3d"
    for i, n in enumerate("xyz"):
        _field(m, n, i + 1, _T.TYPE_DOUBLE)
    m = fd.message_type.add()
    m.name = "Quaterniond"
    for i, n in enumerate("xyzw"):
        _field(m, n, i + 1, _T.TYPE_DOUBLE)
    m = fd.message_type.add()
    m.name = "Rigid3d"
    _field(m, "translation", 1, _T.TYPE_MESSAGE, type_name=P + "Vector3d")
    _field(m, "rotation", 2, _T.TYPE_MESSAGE, type_name=P + "Quaterniond")
    for name, index in (("SubmapId", "submap_index"), ("NodeId", "node_index")):
        m = fd.message_type.add()
        m.name = name
        _field(m, "trajectory_id", 1, _T.TYPE_INT32)
        _field(m, index, 2, _T.TYPE_INT32)
    m = fd.message_type.add()
    m.name = "Trajectory"
    n = m.nested_type.add()
    n.name = "Node"
    _field(n, "node_index", 7, _T.TYPE_INT32)
    _field(n, "timestamp", 1, _T.TYPE_INT64)
    _field(n, "pose", 5, _T.TYPE_MESSAGE, type_name=P + "Rigid3d")
    n = m.nested_type.add()
    n.name = "Submap"
    _field(n, "submap_index", 2, _T.TYPE_INT32)
    _field(n, "pose", 1, _T.TYPE_MESSAGE, type_name=P + "Rigid3d")
    _field(m, "trajectory_id", 3, _T.TYPE_INT32)
    _field(m, "node", 1, _T.TYPE_MESSAGE, label=_T.LABEL_REPEATED, type_name=P + "Trajectory.Node")
    _field(m, "submap", 2, _T.TYPE_MESSAGE, label=_T.LABEL_REPEATED, type_name=P + "Trajectory.Submap")
    m = fd.message_type.add()
    m.name = "PoseGraph"
    c = m.nested_type.add()
    c.name = "Constraint"
    _field(c, "submap_id", 1, _T.TYPE_MESSAGE, type_name=P + "SubmapId")
    _field(c, "node_id", 2, _T.TYPE_MESSAGE, type_name=P + "NodeId")
    _field(c, "relative_pose", 3, _T.TYPE_MESSAGE, type_name=P + "Rigid3d")
    _field(c, "translation_weight", 6, _T.TYPE_DOUBLE)
    _field(c, "rotation_weight", 7, _T.TYPE_DOUBLE)
    _field(c, "tag", 5, _T.TYPE_INT32)  # an enum on the wire is a varint
    lm = m.nested_type.add()
    lm.name = "LandmarkPose"
    _field(lm, "landmark_id", 1, _T.TYPE_STRING)
    _field(lm, "global_pose", 2, _T.TYPE_MESSAGE, type_name=P + "Rigid3d")
    _field(m, "constraint", 2, _T.TYPE_MESSAGE, label=_T.LABEL_REPEATED, type_name=P + "PoseGraph.Constraint")
    _field(m, "trajectory", 4, _T.TYPE_MESSAGE, label=_T.LABEL_REPEATED, type_name=P + "Trajectory")
    _field(m, "landmark_poses", 5, _T.TYPE_MESSAGE, label=_T.LABEL_REPEATED, type_name=P + "PoseGraph.LandmarkPose")
    m = fd.message_type.add()
    m.name = "CompressedPointCloud"
    _field(m, "num_points", 1, _T.TYPE_INT32)
    _field(m, "point_data", 3, _T.TYPE_INT32, label=_T.LABEL_REPEATED)
    m = fd.message_type.add()
    m.name = "TrajectoryNodeData"
    _field(m, "timestamp", 1, _T.TYPE_INT64)
    _field(m, "gravity_alignment", 2, _T.TYPE_MESSAGE, type_name=P + "Quaterniond")
    _field(m, "filtered_gravity_aligned_point_cloud", 3, _T.TYPE_MESSAGE, type_name=P + "CompressedPointCloud")
    _field(m, "high_resolution_point_cloud", 4, _T.TYPE_MESSAGE, type_name=P + "CompressedPointCloud")
    _field(m, "low_resolution_point_cloud", 5, _T.TYPE_MESSAGE, type_name=P + "CompressedPointCloud")
    _field(m, "rotational_scan_matcher_histogram", 6, _T.TYPE_FLOAT, label=_T.LABEL_REPEATED)
    _field(m, "local_pose", 7, _T.TYPE_MESSAGE, type_name=P + "Rigid3d")
    m = fd.message_type.add()
    m.name = "Submap2D"
    _field(m, "local_pose", 1, _T.TYPE_MESSAGE, type_name=P + "Rigid3d")
    _field(m, "num_range_data", 2, _T.TYPE_INT32)
    _field(m, "finished", 3, _T.TYPE_BOOL)
    _field(m, "grid", 4, _T.TYPE_MESSAGE, type_name=P + "Grid2D")
    m = fd.message_type.add()
    m.name = "Submap3D"
    _field(m, "local_pose", 1, _T.TYPE_MESSAGE, type_name=P + "Rigid3d")
    _field(m, "num_range_data", 2, _T.TYPE_INT32)
    _field(m, "finished", 3, _T.TYPE_BOOL)
    _field(m, "high_resolution_hybrid_grid", 4, _T.TYPE_MESSAGE, type_name=P + "HybridGrid")
    _field(m, "low_resolution_hybrid_grid", 5, _T.TYPE_MESSAGE, type_name=P + "HybridGrid")
    _field(m, "rotational_scan_matcher_histogram", 6, _T.TYPE_FLOAT, label=_T.LABEL_REPEATED)
    m = fd.message_type.add()
    m.name = "Submap"
    _field(m, "submap_id", 1, _T.TYPE_MESSAGE, type_name=P + "SubmapId")
    _field(m, "submap_2d", 2, _T.TYPE_MESSAGE, type_name=P + "Submap2D")
    _field(m, "submap_3d", 3, _T.TYPE_MESSAGE, type_name=P + "Submap3D")
    m = fd.message_type.add()
    m.name = "Node"
    _field(m, "node_id", 1, _T.TYPE_MESSAGE, type_name=P + "NodeId")
    _field(m, "node_data", 5, _T.TYPE_MESSAGE, type_name=P + "TrajectoryNodeData")
    m = fd.message_type.add()
    m.name = "OptionsWithSensorIds"  # TrajectoryBuilderOptionsWithSensorIds: contents are not read
    _field(m, "sensor_id", 1, _T.TYPE_STRING, label=_T.LABEL_REPEATED)
    m = fd.message_type.add()
    m.name = "AllTrajectoryBuilderOptions"
    _field(m, "options_with_sensor_ids", 1, _T.TYPE_MESSAGE, label=_T.LABEL_REPEATED,
           type_name=P + "OptionsWithSensorIds")
    m = fd.message_type.add()
    m.name = "ImuData"  # serialization.proto ImuData {trajectory_id = 1; sensor ImuData = 2}: skipped by the reader
    _field(m, "trajectory_id", 1, _T.TYPE_INT32)
    _field(m, "payload", 2, _T.TYPE_BYTES)
    m = fd.message_type.add()
    m.name = "SerializationHeader"
    _field(m, "format_version", 1, _T.TYPE_UINT32)
    m = fd.message_type.add()
    m.name = "SerializedData"
    m.oneof_decl.add().name = "data"
    _field(m, "pose_graph", 1, _T.TYPE_MESSAGE, type_name=P + "PoseGraph", oneof=0)
    _field(m, "all_trajectory_builder_options", 2, _T.TYPE_MESSAGE, type_name=P + "AllTrajectoryBuilderOptions", oneof=0)
    _field(m, "submap", 3, _T.TYPE_MESSAGE, type_name=P + "Submap", oneof=0)
    _field(m, "node", 4, _T.TYPE_MESSAGE, type_name=P + "Node", oneof=0)
    _field(m, "imu_data", 6, _T.TYPE_MESSAGE, type_name=P + "ImuData", oneof=0)


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
    _build_serialization(fd)
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    get = lambda n: message_factory.GetMessageClass(pool.FindMessageTypeByName("cartographer.mapping.proto." + n))
    return get


_get = _build()
Grid2D, HybridGrid = _get("Grid2D"), _get("HybridGrid")
SerializationHeader, SerializedData = _get("SerializationHeader"), _get("SerializedData")


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


# ---------------------------------------------------------------------------------------------
# A saved mapping state (.pbstream) made WITHOUT the product: Python's gzip + struct for the
# container (io/proto_stream.cc:27-57), the protobuf runtime for the messages.
# ---------------------------------------------------------------------------------------------
import gzip
import struct

import numpy as np

PBSTREAM_MAGIC = 0x7B1D1F7B5BF501DB


def pbstream_bytes(messages, compresslevel=1):
    out = [struct.pack("<Q", PBSTREAM_MAGIC)]
    for m in messages:
        z = gzip.compress(m, compresslevel=compresslevel, mtime=0)
        out += [struct.pack("<Q", len(z)), z]
    return b"".join(out)


def pbstream_parse(blob):
    assert struct.unpack_from("<Q", blob, 0)[0] == PBSTREAM_MAGIC
    at, out = 8, []
    while at < len(blob):
        (n,) = struct.unpack_from("<Q", blob, at)
        out.append(gzip.decompress(blob[at + 8:at + 8 + n]))
        at += 8 + n
    return out


def compress_point_cloud(points):
    """sensor::CompressedPointCloud's layout (compressed_point_cloud.cc:30-36,103-146) written
    independently: millimetre raster, 1024^3-mm blocks, [count, bx, by, bz, points...] per block.
    Returns (num_points, int32 words, order) with `order` the input indices in encoded order
    (decoding returns the points block by block)."""
    pts = np.asarray(points, np.float32).reshape(-1, 3)
    raster = np.array([[int(np.floor(np.float32(v) / np.float32(0.001) + np.float32(0.5)))
                        if v >= 0 else -int(np.floor(-np.float32(v) / np.float32(0.001) + np.float32(0.5)))
                        for v in p] for p in pts], dtype=np.int64).reshape(-1, 3)
    blocks = {}
    for i, r in enumerate(raster):
        blocks.setdefault(tuple(int(c) >> 10 for c in r), []).append(i)
    words, order = [], []
    for key in sorted(blocks):
        idx = blocks[key]
        words += [len(idx), key[0], key[1], key[2]]
        for i in idx:
            x, y, z = (int(c) & 1023 for c in raster[i])
            words.append((((z << 10) + y) << 10) + x)
        order += idx
    words = np.array(words, dtype=np.int64)
    words = np.where(words >= 2 ** 31, words - 2 ** 32, words).astype(np.int32)
    return len(pts), words, np.array(order, dtype=np.int64)


def decode_point_cloud_reference_math(num_points, words):
    """What ConstIterator::ReadNextPoint computes (cc:80-101), in numpy float32."""
    out, at, left, block = [], 0, 0, (0, 0, 0)
    for _ in range(num_points):
        if left == 0:
            left = int(words[at])
            block = tuple(int(words[at + 1 + k]) << 10 for k in range(3))
            at += 4
        left -= 1
        p = int(words[at])
        at += 1
        c = (block[0] + (p & 1023), block[1] + ((p >> 10) & 1023), block[2] + (p >> 20))
        out.append([np.float32(np.float32(v) * np.float32(0.001)) for v in c])
    return np.array(out, dtype=np.float32).reshape(-1, 3)


def set_pose(msg, pose7):
    """pose7 = (tx, ty, tz, qw, qx, qy, qz) -> a transform.proto Rigid3d."""
    msg.translation.x, msg.translation.y, msg.translation.z = (float(v) for v in pose7[:3])
    msg.rotation.w, msg.rotation.x, msg.rotation.y, msg.rotation.z = (float(v) for v in pose7[3:])


def fill_cloud(msg, points):
    n, words, order = compress_point_cloud(points)
    msg.num_points = n
    msg.point_data.extend(int(w) for w in words)
    return order


def mapping_state_messages(submaps, nodes, constraints=(), format_version=2, extra_imu=True):
    """Serialized messages of a saved state, in stream order.  submaps: dicts {id: (traj, index),
    local_pose, global_pose, num_range_data, finished, and either grid_bytes (a serialized Grid2D) or
    high_bytes / low_bytes / histogram}; nodes: dicts {id, timestamp, gravity_alignment (w, x, y, z),
    local_pose, global_pose, filtered / high / low (point arrays), histogram}; constraints: dicts
    {submap_id, node_id, relative_pose, translation_weight, rotation_weight, tag}."""
    header = SerializationHeader()
    header.format_version = format_version
    graph = SerializedData()
    trajectories = {}
    for s in submaps:
        trajectories.setdefault(s["id"][0], ([], []))[0].append(s)
    for n in nodes:
        trajectories.setdefault(n["id"][0], ([], []))[1].append(n)
    for tid in sorted(trajectories):
        t = graph.pose_graph.trajectory.add()
        t.trajectory_id = tid
        for s in trajectories[tid][0]:
            e = t.submap.add()
            e.submap_index = s["id"][1]
            set_pose(e.pose, s["global_pose"])
        for n in trajectories[tid][1]:
            e = t.node.add()
            e.node_index, e.timestamp = n["id"][1], n["timestamp"]
            set_pose(e.pose, n["global_pose"])
    for c in constraints:
        e = graph.pose_graph.constraint.add()
        e.submap_id.trajectory_id, e.submap_id.submap_index = c["submap_id"]
        e.node_id.trajectory_id, e.node_id.node_index = c["node_id"]
        set_pose(e.relative_pose, c["relative_pose"])
        e.translation_weight, e.rotation_weight, e.tag = c["translation_weight"], c["rotation_weight"], c["tag"]
    lm = graph.pose_graph.landmark_poses.add()  # a field the reader skips
    lm.landmark_id = "door"
    set_pose(lm.global_pose, (1, 2, 3, 1, 0, 0, 0))
    options = SerializedData()
    for _ in trajectories:
        options.all_trajectory_builder_options.options_with_sensor_ids.add().sensor_id.append("scan")
    if not trajectories:
        options.all_trajectory_builder_options.SetInParent()
    messages = [header.SerializeToString(), graph.SerializeToString(), options.SerializeToString()]
    for s in submaps:
        d = SerializedData()
        d.submap.submap_id.trajectory_id, d.submap.submap_id.submap_index = s["id"]
        if "grid_bytes" in s:
            sub = d.submap.submap_2d
            if s["grid_bytes"] is not None:
                sub.grid.ParseFromString(s["grid_bytes"])
        else:
            sub = d.submap.submap_3d
            sub.high_resolution_hybrid_grid.ParseFromString(s["high_bytes"])
            sub.low_resolution_hybrid_grid.ParseFromString(s["low_bytes"])
            sub.rotational_scan_matcher_histogram.extend(float(v) for v in s["histogram"])
        set_pose(sub.local_pose, s["local_pose"])
        sub.num_range_data, sub.finished = s["num_range_data"], s["finished"]
        messages.append(d.SerializeToString())
    if extra_imu:
        d = SerializedData()
        d.imu_data.trajectory_id, d.imu_data.payload = 0, b"\x01\x02\x03"
        messages.append(d.SerializeToString())
    for n in nodes:
        d = SerializedData()
        d.node.node_id.trajectory_id, d.node.node_id.node_index = n["id"]
        nd = d.node.node_data
        nd.timestamp = n["timestamp"]
        g = n["gravity_alignment"]
        nd.gravity_alignment.w, nd.gravity_alignment.x, nd.gravity_alignment.y, nd.gravity_alignment.z = (float(v) for v in g)
        n["filtered_order"] = fill_cloud(nd.filtered_gravity_aligned_point_cloud, n.get("filtered", np.zeros((0, 3))))
        n["high_order"] = fill_cloud(nd.high_resolution_point_cloud, n.get("high", np.zeros((0, 3))))
        n["low_order"] = fill_cloud(nd.low_resolution_point_cloud, n.get("low", np.zeros((0, 3))))
        nd.rotational_scan_matcher_histogram.extend(float(v) for v in n.get("histogram", ()))
        set_pose(nd.local_pose, n["local_pose"])
        messages.append(d.SerializeToString())
    return messages
