"""A saved mapping state (.pbstream) through the host layer's readers
(host/cartographer/io/proto_stream*.h, sensor/compressed_point_cloud.h).  Three independent makers
of the bytes check them: Python's gzip + struct + protobuf runtime, the product's own writer, and --
where oracle/_ref exists -- the reference's own ProtoStreamWriter / CompressedPointCloud compiled
from /root/reference.  Host only: no GPU needed."""
import os
import struct

import numpy as np
import pytest

import proto_fixtures as pf
import ref_binding as rb
import synth

needs_ref = pytest.mark.skipif(not rb.available(), reason="oracle/_ref is not built and /root/reference is absent")


def _messages(rng, count=7):
    sizes = [0, 1, 17, 4096, 70000, 300000, 5][:count]
    return [bytes(rng.integers(0, 4, n, dtype=np.uint8)) for n in sizes]  # low entropy: really compresses


def test_container_written_by_python_is_read(csm, tmp_path):
    msgs = _messages(np.random.default_rng(1))
    path = tmp_path / "py.pbstream"
    path.write_bytes(pf.pbstream_bytes(msgs, compresslevel=6))
    assert csm.pbstream_read(str(path)) == msgs


def test_container_written_by_the_product_is_read_by_python(csm, tmp_path):
    msgs = _messages(np.random.default_rng(2))
    path = tmp_path / "product.pbstream"
    csm.pbstream_write(str(path), msgs)
    blob = path.read_bytes()
    assert struct.unpack_from("<Q", blob, 0)[0] == pf.PBSTREAM_MAGIC
    assert blob[16:18] == b"\x1f\x8b"  # a gzip member follows the first size
    assert pf.pbstream_parse(blob) == msgs
    assert csm.pbstream_read(str(path)) == msgs


def test_empty_stream_and_truncated_stream(csm, tmp_path):
    path = tmp_path / "empty.pbstream"
    csm.pbstream_write(str(path), [])
    assert path.read_bytes() == struct.pack("<Q", pf.PBSTREAM_MAGIC)
    assert csm.pbstream_read(str(path)) == []
    whole = pf.pbstream_bytes([b"abc" * 100, b"xyz" * 100])
    cut = tmp_path / "cut.pbstream"
    cut.write_bytes(whole[:-7])  # the last record ends before its size says so
    with pytest.raises(IOError):
        csm.pbstream_read(str(cut))


@needs_ref
def test_container_against_the_reference_writer_and_reader(csm, tmp_path):
    msgs = _messages(np.random.default_rng(3))
    ref_path, own_path = tmp_path / "ref.pbstream", tmp_path / "own.pbstream"
    rb.pbstream_write(str(ref_path), msgs)
    csm.pbstream_write(str(own_path), msgs)
    assert csm.pbstream_read(str(ref_path)) == msgs          # product reads the reference's file
    assert rb.pbstream_read(str(own_path)) == msgs           # reference reads the product's file
    assert pf.pbstream_parse(ref_path.read_bytes()) == msgs  # and Python agrees on the layout
    # same zlib, same level: the two writers produce the same file byte for byte
    assert own_path.read_bytes() == ref_path.read_bytes()


def _cloud(rng, n, spread):
    return (rng.uniform(-spread, spread, (n, 3))).astype(np.float32)


@needs_ref
@pytest.mark.parametrize("n,spread", [(0, 1.0), (1, 0.3), (500, 0.9), (3000, 30.0), (400, 4000.0)])
def test_compressed_point_cloud_decodes_like_the_reference(csm, tmp_path, n, spread):
    rng = np.random.default_rng(100 + n)
    pts = _cloud(rng, n, spread)
    words = rb.compress_point_cloud(pts)            # the reference's encoder
    want = rb.decompress_point_cloud(n, words)      # the reference's decoder
    # product: the words inside a node record of a saved state
    node = {"id": (0, 0), "timestamp": 1, "gravity_alignment": (1, 0, 0, 0), "local_pose": (0, 0, 0, 1, 0, 0, 0),
            "global_pose": (0, 0, 0, 1, 0, 0, 0)}
    msgs = pf.mapping_state_messages([], [node])
    d = pf.SerializedData.FromString(msgs[-1])
    d.node.node_data.filtered_gravity_aligned_point_cloud.num_points = n
    del d.node.node_data.filtered_gravity_aligned_point_cloud.point_data[:]
    d.node.node_data.filtered_gravity_aligned_point_cloud.point_data.extend(int(w) for w in words)
    msgs[-1] = d.SerializeToString()
    path = tmp_path / "node.pbstream"
    path.write_bytes(pf.pbstream_bytes(msgs))
    state = csm.MappingState(str(path))
    got = state.nodes[0]["filtered_gravity_aligned_point_cloud"]
    assert got.tobytes() == want.tobytes()
    # millimetre quantisation and nothing else; the Python statement of the decode agrees too
    assert got.tobytes() == pf.decode_point_cloud_reference_math(n, words).tobytes()
    if n:
        # decoding returns the points block by block; every original has its decoded point within
        # half a millimetre per axis (plus float32 spacing at the coordinate's size)
        from scipy.spatial import cKDTree
        dist, idx = cKDTree(got.astype(np.float64)).query(pts.astype(np.float64), p=np.inf)
        assert (dist <= 0.0005 + 2e-7 * spread).all() and len(np.unique(idx)) >= n - 2


@needs_ref
def test_python_encoder_matches_the_reference_decoder(csm):
    # the fixture's own encoder (used by the GPU test on the box) is held to the reference's decoder
    rng = np.random.default_rng(9)
    pts = _cloud(rng, 800, 12.0)
    n, words, order = pf.compress_point_cloud(pts)
    got = rb.decompress_point_cloud(n, words)
    want = (np.round(pts[order].astype(np.float64) * 1000)).astype(np.int64)
    assert (np.round(got.astype(np.float64) * 1000).astype(np.int64) == want).all()


def _state_2d(rng, num_submaps=3, nodes_per_submap=2):
    submaps, nodes = [], []
    for s in range(num_submaps):
        w = synth.World(120 + 10 * s, 100, seed=40 + s)
        yaw = 0.3 * s
        submaps.append({"id": (s % 2, s // 2), "world": w,
                        "grid_bytes": pf.grid2d_bytes(w.cells, w.res, w.max_x, w.max_y),
                        "local_pose": (1.0 + s, -2.0, 0.0, np.cos(yaw / 2), 0.0, 0.0, np.sin(yaw / 2)),
                        "global_pose": (1.5 + s, -2.25, 0.0, np.cos(yaw / 2 + 0.01), 0.0, 0.0, np.sin(yaw / 2 + 0.01)),
                        "num_range_data": 90 + s, "finished": s != num_submaps - 1})
        for k in range(nodes_per_submap):
            i = len(nodes)
            nodes.append({"id": (s % 2, i), "timestamp": 636_000_000_000_000_000 + 1000 * i,
                          "gravity_alignment": (0.999, 0.01, -0.02, 0.03),
                          "local_pose": (0.1 * i, 0.2, 0.0, 1.0, 0.0, 0.0, 0.0),
                          "global_pose": (0.1 * i + 0.05, 0.25, 0.0, 0.995, 0.0, 0.0, 0.0998),
                          "filtered": _cloud(rng, 150 + 20 * k, 8.0) * np.float32([1, 1, 0]),
                          "histogram": rng.random(0 if k else 5).astype(np.float32)})
    constraints = [{"submap_id": submaps[0]["id"], "node_id": nodes[-1]["id"],
                    "relative_pose": (0.5, -0.25, 0.0, 0.9, 0.0, 0.0, 0.4359),
                    "translation_weight": 1.1e4, "rotation_weight": 1e5, "tag": 1},
                   {"submap_id": submaps[1]["id"], "node_id": nodes[0]["id"],
                    "relative_pose": (0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0),
                    "translation_weight": 5e2, "rotation_weight": 1.6e3, "tag": 0}]
    return submaps, nodes, constraints


@pytest.mark.parametrize("writer", ["python", "product"])
def test_saved_2d_state_round_trip(csm, tmp_path, writer):
    rng = np.random.default_rng(4)
    submaps, nodes, constraints = _state_2d(rng)
    msgs = pf.mapping_state_messages(submaps, nodes, constraints)
    path = tmp_path / "state.pbstream"
    if writer == "python":
        path.write_bytes(pf.pbstream_bytes(msgs))
    else:
        csm.pbstream_write(str(path), msgs)
    state = csm.MappingState(str(path))
    assert state.format_version == 2
    assert len(state.submaps) == len(submaps) and len(state.nodes) == len(nodes)
    for got, want in zip(state.submaps, submaps):
        assert (got["trajectory_id"], got["submap_index"]) == want["id"] and not got["is_3d"]
        assert got["num_range_data"] == want["num_range_data"] and got["finished"] == want["finished"]
        assert got["local_pose"].tolist() == [float(v) for v in want["local_pose"]]
        assert got["global_pose"].tolist() == [float(v) for v in want["global_pose"]]
        w = want["world"]
        g = got["grid"]
        assert (g.nx, g.ny, g.res, g.max_x, g.max_y, g.is_tsdf) == (w.nx, w.ny, w.res, w.max_x, w.max_y, False)
        assert g.cells().tobytes() == np.ascontiguousarray(w.cells).tobytes()
    for got, want in zip(state.nodes, nodes):
        assert (got["trajectory_id"], got["node_index"]) == want["id"] and got["timestamp"] == want["timestamp"]
        assert got["gravity_alignment"].tolist() == [float(v) for v in want["gravity_alignment"]]
        assert got["local_pose"].tolist() == [float(v) for v in want["local_pose"]]
        assert got["global_pose"].tolist() == [float(v) for v in want["global_pose"]]
        n, words, order = pf.compress_point_cloud(want["filtered"])
        assert got["filtered_gravity_aligned_point_cloud"].tobytes() == pf.decode_point_cloud_reference_math(n, words).tobytes()
        assert len(got["high_resolution_point_cloud"]) == 0 and len(got["low_resolution_point_cloud"]) == 0
        assert got["histogram"].tobytes() == np.asarray(want["histogram"], np.float32).tobytes()
    assert len(state.constraints) == 2
    for got, want in zip(state.constraints, constraints):
        assert got["submap_id"] == want["submap_id"] and got["node_id"] == want["node_id"]
        assert got["tag"] == ("INTER_SUBMAP" if want["tag"] else "INTRA_SUBMAP")
        assert got["relative_pose"].tolist() == [float(v) for v in want["relative_pose"]]
        assert (got["translation_weight"], got["rotation_weight"]) == (want["translation_weight"], want["rotation_weight"])
    # a grid handle outlives the state object it came from
    grid = state.submaps[0]["grid"]
    state.close()
    assert grid.cells().tobytes() == np.ascontiguousarray(submaps[0]["world"].cells).tobytes()


def test_saved_3d_state_round_trip(csm, tmp_path):
    rng = np.random.default_rng(5)
    def voxels(n, extent):
        xyz = np.unique(rng.integers(-extent, extent, (n, 3)), axis=0).astype(np.int32)
        return xyz, rng.integers(1, 32768, len(xyz)).astype(np.uint16)
    hi_xyz, hi_val = voxels(4000, 90)   # beyond +-64: the grid must have grown to 256
    lo_xyz, lo_val = voxels(600, 20)
    hist = rng.random(120).astype(np.float32)
    submaps = [{"id": (0, 0), "high_bytes": pf.hybrid_grid_bytes(0.1, hi_xyz, hi_val),
                "low_bytes": pf.hybrid_grid_bytes(0.45, lo_xyz, lo_val), "histogram": hist,
                "local_pose": (0.5, 1.5, -0.25, 0.9, 0.1, -0.2, 0.3741657), "global_pose": (0, 0, 0, 1, 0, 0, 0),
                "num_range_data": 160, "finished": True}]
    nodes = [{"id": (0, 0), "timestamp": 7, "gravity_alignment": (1, 0, 0, 0), "local_pose": (0, 0, 0, 1, 0, 0, 0),
              "global_pose": (3, 2, 1, 1, 0, 0, 0), "high": _cloud(rng, 900, 15.0), "low": _cloud(rng, 200, 15.0),
              "histogram": rng.random(120).astype(np.float32)}]
    path = tmp_path / "state3d.pbstream"
    path.write_bytes(pf.pbstream_bytes(pf.mapping_state_messages(submaps, nodes)))
    state = csm.MappingState(str(path), expect_3d=True)
    s = state.submaps[0]
    assert s["is_3d"] and s["grid"] is None and s["histogram"].tobytes() == hist.tobytes()
    for key, res, xyz, val in (("high", 0.1, hi_xyz, hi_val), ("low", 0.45, lo_xyz, lo_val)):
        got_res, size, got_xyz, got_val = s[key]
        order = np.lexsort((xyz[:, 0], xyz[:, 1], xyz[:, 2]))
        assert got_res == np.float32(res) and size == (256 if key == "high" else 128)
        assert (got_xyz == xyz[order]).all() and (got_val == val[order]).all()
    n = state.nodes[0]
    for key, src in (("high_resolution_point_cloud", "high"), ("low_resolution_point_cloud", "low")):
        cnt, words, order = pf.compress_point_cloud(nodes[0][src])
        assert n[key].tobytes() == pf.decode_point_cloud_reference_math(cnt, words).tobytes()
    assert n["histogram"].tobytes() == nodes[0]["histogram"].tobytes()
    assert len(n["filtered_gravity_aligned_point_cloud"]) == 0


def test_reader_refuses_what_the_reference_refuses(csm, tmp_path):
    """The reference CHECK-fails (aborts) on these; so does the reader -- in a child process."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    def load_aborts(blob, expect_3d=False):
        path = tmp_path / "bad.pbstream"
        path.write_bytes(blob)
        code = ("import sys; sys.path.insert(0, %r); import conftest; csm = conftest.load_package(); "
                "csm.MappingState(%r, expect_3d=%r)" % (os.path.join(root, "tests"), str(path), expect_3d))
        return subprocess.run([sys.executable, "-c", code], capture_output=True).returncode != 0
    good = pf.mapping_state_messages([], [])
    assert not load_aborts(pf.pbstream_bytes(good))
    assert load_aborts(b"\0" * 8 + pf.pbstream_bytes(good)[8:])                     # wrong magic
    assert load_aborts(pf.pbstream_bytes(pf.mapping_state_messages([], [], format_version=3)))
    assert load_aborts(pf.pbstream_bytes([good[0], good[2], good[1]]))               # stream out of order
    assert load_aborts(pf.pbstream_bytes(good[:2]))                                  # builder options missing
    assert not load_aborts(pf.pbstream_bytes(pf.mapping_state_messages([], [], format_version=1)))
    assert load_aborts(pf.pbstream_bytes(pf.mapping_state_messages([], [], format_version=1)), expect_3d=True)
    # a submap record whose id the pose graph does not know (`submap_poses.at`)
    orphan = pf.SerializedData()
    orphan.submap.submap_id.submap_index = 9
    orphan.submap.submap_2d.finished = True
    assert load_aborts(pf.pbstream_bytes(good + [orphan.SerializeToString()]))


def test_mutated_messages_never_crash_the_reader(csm, tmp_path):
    """Truncated and bit-flipped records (inside valid gzip members, so the damage reaches the wire
    reader): the loader either returns or stops on one of its checks (SIGABRT, as the reference's
    CHECKs do) -- never a segmentation fault, a hang or a runaway allocation."""
    import signal
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    rng = np.random.default_rng(11)
    submaps, nodes, constraints = _state_2d(rng, num_submaps=2, nodes_per_submap=1)
    good = pf.mapping_state_messages(submaps, nodes, constraints)
    paths = []
    for case in range(24):
        msgs = list(good)
        k = int(rng.integers(1, len(msgs)))  # not the header: the version check would stop everything
        m = bytearray(msgs[k])
        if case % 3 == 0 and len(m) > 2:
            m = m[: int(rng.integers(1, len(m)))]                       # truncated record
        elif case % 3 == 1:
            for _ in range(int(rng.integers(1, 6))):
                m[int(rng.integers(0, len(m)))] ^= 1 << int(rng.integers(0, 8))  # flipped bits
        else:
            at = int(rng.integers(0, len(m)))
            m[at:at] = bytes(rng.integers(0, 256, int(rng.integers(1, 12)), dtype=np.uint8))  # inserted garbage
        msgs[k] = bytes(m)
        path = tmp_path / ("mutated_%d.pbstream" % case)
        path.write_bytes(pf.pbstream_bytes(msgs))
        paths.append(str(path))
    code = ("import sys, os, resource; sys.path.insert(0, %r); import conftest; csm = conftest.load_package()\n"
            "resource.setrlimit(resource.RLIMIT_AS, (8 << 30, 8 << 30))\n"
            "for p in sys.argv[1:]:\n"
            "    pid = os.fork()\n"
            "    if pid == 0:\n"
            "        s = csm.MappingState(p); os._exit(0)\n"
            "    _, status = os.waitpid(pid, 0)\n"
            "    print(p, os.WTERMSIG(status) if os.WIFSIGNALED(status) else 0, flush=True)\n"
            % os.path.join(root, "tests"))
    out = subprocess.run([sys.executable, "-c", code] + paths, capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr[-2000:]
    ends = [int(line.split()[-1]) for line in out.stdout.strip().splitlines()]
    assert len(ends) == len(paths)
    assert set(ends) <= {0, int(signal.SIGABRT)}, ends
    assert int(signal.SIGABRT) in ends  # at least one mutation was caught by a check
