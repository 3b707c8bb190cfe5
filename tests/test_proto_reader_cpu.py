"""Grid ingestion (host/cartographer/io/grid_proto_reader.h) on real protobuf wire bytes: the grids
that come out hold exactly what went in, under the rules of the reference's proto constructors
(grid_2d.cc:22-96, tsdf_2d.cc:35-47, hybrid_grid.h:473-485).  Host only: no GPU needed."""
import numpy as np

import oracle_binding as ob
import proto_fixtures as pf
import synth


def test_probability_grid_round_trip(csm):
    w = synth.World(150, 120, seed=5)
    g = csm.ProtoGrid(pf.grid2d_bytes(w.cells, w.res, w.max_x, w.max_y, box=(3, 4, 140, 110)))
    assert (g.nx, g.ny, g.res, g.max_x, g.max_y, g.is_tsdf) == (150, 120, w.res, w.max_x, w.max_y, False)
    assert g.cells().tobytes() == np.ascontiguousarray(w.cells).tobytes()
    # the bounds of the message select the value table: identical quantisation to a grid made directly
    assert g.quantization_table().tobytes() == csm.Grid(w.cells, w.res, w.max_x, w.max_y).quantization_table().tobytes()


def test_legacy_message_without_cost_bounds_loads_the_defaults(csm):
    w = synth.World(40, 30, seed=6)
    g = csm.ProtoGrid(pf.grid2d_bytes(w.cells, w.res, w.max_x, w.max_y, min_cost=0.0, max_cost=0.0))
    assert g.quantization_table().tobytes() == csm.Grid(w.cells, w.res, w.max_x, w.max_y).quantization_table().tobytes()


def test_tsdf_round_trip(csm):
    rng = np.random.default_rng(2)
    tsd = rng.integers(0, 32768, (25, 35)).astype(np.uint16)
    weight = rng.integers(0, 32768, (25, 35)).astype(np.uint16)
    g = csm.ProtoGrid(pf.grid2d_bytes(tsd, 0.05, 1.0, 2.0, min_cost=-0.3, max_cost=0.3, tsdf=(weight, 0.3, 10.0)))
    assert g.is_tsdf and (g.nx, g.ny) == (35, 25)
    assert g.cells().tobytes() == tsd.tobytes() and g.weight_cells().tobytes() == weight.tobytes()


def test_unpacked_repeated_fields_are_accepted(csm):
    # proto3 writers pack `cells`; a proto2-style writer emits one (tag, varint) record per element.
    packed = pf.grid2d_bytes(np.array([[1, 300, 32767], [0, 7, 128]], np.uint16), 0.1, 1.0, 1.0)
    msg = pf.Grid2D.FromString(packed)
    def varint(v):
        out = bytearray()
        while True:
            b = v & 0x7F
            v >>= 7
            out.append(b | (0x80 if v else 0))
            if not v:
                return bytes(out)
    cells = list(msg.cells)
    del msg.cells[:]
    unpacked = msg.SerializeToString() + b"".join(b"\x10" + varint(c) for c in cells)  # field 2, wire type 0
    g = csm.ProtoGrid(unpacked)
    assert g.cells().reshape(-1).tolist() == cells


def test_hybrid_grid_round_trip(csm):
    rng = np.random.default_rng(4)
    xyz = np.unique(rng.integers(-90, 90, (500, 3)), axis=0).astype(np.int32)
    values = rng.integers(1, 32768, len(xyz)).astype(np.uint16)
    res, size, got_xyz, got_val = csm.hybrid_grid_from_proto(pf.hybrid_grid_bytes(0.1, xyz, values))
    o = ob.OracleGrid3D(0.1, xyz, values)
    want_xyz, want_val = o.cells()
    assert abs(res - 0.1) < 1e-7 and size == o.grid_size == 256
    assert got_xyz.tobytes() == want_xyz.tobytes() and got_val.tobytes() == want_val.tobytes()
    assert (np.sort(got_val) == np.sort(values)).all()  # ProbabilityToValue(ValueToProbability(v)) == v
