"""A saved mapping state (.pbstream) -> loop-closure searches on the device: the bytes are made
with Python's gzip and the protobuf runtime (tests/proto_fixtures.py), read by the host layer's
ProtoStreamDeserializer, searched through ConstraintBuilder2D, and compared with the CPU oracle run
on the grids and point clouds the fixture was made from (decoded with the reference's arithmetic in
numpy, not with the product's reader)."""
import os
import sys

import numpy as np
import pytest

import oracle_binding as ob
import proto_fixtures as pf
import synth

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
import pbstream_job  # noqa: E402

pytestmark = pytest.mark.gpu


def _pose7(x, y, yaw):
    return (x, y, 0.0, np.cos(yaw / 2), 0.0, 0.0, np.sin(yaw / 2))


def _state(seed, n_sub=3, nodes_per_sub=4, size=240):
    rng = np.random.default_rng(seed)
    worlds = [synth.World(size, size, seed=seed * 10 + k) for k in range(n_sub)]
    submaps, nodes = [], []
    for k, w in enumerate(worlds):
        local = (rng.uniform(-2, 2), rng.uniform(-2, 2), rng.uniform(-1.5, 1.5))
        drift = (local[0] + rng.uniform(-0.2, 0.2), local[1] + rng.uniform(-0.2, 0.2), local[2] + rng.uniform(-0.05, 0.05))
        submaps.append({"id": (k % 2, k // 2), "world": w, "local2d": local, "global2d": drift,
                        "grid_bytes": pf.grid2d_bytes(w.cells, w.res, w.max_x, w.max_y),
                        "local_pose": _pose7(*local), "global_pose": _pose7(*drift),
                        "num_range_data": 90, "finished": True})
        for j in range(nodes_per_sub):
            pose, pts = synth.make_scan(w, seed * 1000 + 10 * k + j, max_range=6.0)
            # the node's pose in the submap frame is `pose` relative to the submap's local pose; put it
            # in the global frame through the submap's (drifted) global pose, plus an odometry error
            c, s = np.cos(-local[2]), np.sin(-local[2])
            dx, dy = pose[0] - local[0], pose[1] - local[1]
            rel = (c * dx - s * dy + rng.uniform(-0.4, 0.4), s * dx + c * dy + rng.uniform(-0.4, 0.4),
                   pose[2] - local[2] + rng.uniform(-0.15, 0.15))
            cg, sg = np.cos(drift[2]), np.sin(drift[2])
            glob = (drift[0] + cg * rel[0] - sg * rel[1], drift[1] + sg * rel[0] + cg * rel[1], drift[2] + rel[2])
            nodes.append({"id": (k % 2, len(nodes)), "timestamp": 1000 + len(nodes), "gravity_alignment": (1, 0, 0, 0),
                          "local_pose": _pose7(*pose), "global_pose": _pose7(*glob), "filtered": pts, "home": k})
    # the pairs a node was inserted into are INTRA_SUBMAP constraints of the saved graph: not searched again
    constraints = [{"submap_id": submaps[n["home"]]["id"], "node_id": n["id"], "relative_pose": _pose7(0, 0, 0),
                    "translation_weight": 1.0, "rotation_weight": 1.0, "tag": 0} for n in nodes[::2]]
    return worlds, submaps, nodes, constraints


def test_searches_over_a_saved_state_match_the_oracle(csm, tmp_path):
    worlds, submaps, nodes, constraints = _state(7)
    path = tmp_path / "map.pbstream"
    path.write_bytes(pf.pbstream_bytes(pf.mapping_state_messages(submaps, nodes, constraints)))
    state = csm.MappingState(str(path))
    grids, poses, sub, full, rel, clouds, pairs = pbstream_job.job_from_state(state)
    assert len(grids) == 3 and len(clouds) == 3 * len(nodes) - len(constraints)
    assert 0 < sum(full) < len(full)  # both kinds of search
    # the oracle's inputs come from the fixture, not from the reader under test
    ogrids = [ob.OracleGrid(w.cells, w.res, w.max_x, w.max_y) for w in worlds]
    by_id = {n["id"]: n for n in nodes}
    o_clouds = []
    for node_id, _ in pairs:
        cnt, words, _ = pf.compress_point_cloud(by_id[node_id]["filtered"])
        o_clouds.append(pf.decode_point_cloud_reference_math(cnt, words))
    for a, b in zip(clouds, o_clouds):
        assert a.tobytes() == b.tobytes()
    o_poses = np.array([s["local2d"] for s in submaps])
    assert np.abs(poses - o_poses).max() < 1e-12
    kw = dict(max_constraint_distance=15.0, min_score=0.45, global_min_score=0.5, linear_window=3.0,
              angular_window=0.5, depth=5)
    _, o_out, _ = ob.constraint_batch(4, ogrids, poses, sub, full, rel, o_clouds, **kw)
    _, found, tf, stats = csm.constraint_builder_run(grids, poses, sub, full, rel, clouds, **kw)
    assert found.sum() >= len(nodes)  # every node finds at least its own neighbourhood again
    for i in range(len(clouds)):
        assert found[i] == o_out[i].found, i
        if found[i]:
            assert abs(tf[i][0] - o_out[i].pose_x) <= 1e-9 * max(1.0, abs(o_out[i].pose_x))
            assert abs(tf[i][1] - o_out[i].pose_y) <= 1e-9 * max(1.0, abs(o_out[i].pose_y))
            d = (tf[i][2] - o_out[i].pose_theta + np.pi) % (2 * np.pi) - np.pi
            assert abs(d) <= 1e-9
    assert stats[1] == found.sum()


def test_3d_state_searches_match_the_oracle(csm, gpu_ctx, tmp_path):
    """The same for a 3D map: hybrid grids, histograms and the nodes' two point clouds travel through
    the saved state; ConstraintBuilder3D on what the reader returns against the oracle's matcher on
    the fixture's own data (clouds decoded with the reference's arithmetic in numpy)."""
    import synth3d
    worlds = [synth3d.World3D(61 + k, size_xy=8.0) for k in range(2)]
    opt = ob.options3d(6, 3, 0.3, 0.4, 1.2, 0.4, np.radians(10))
    rng = np.random.default_rng(8)
    submaps, nodes, raw = [], [], []
    for k, w in enumerate(worlds):
        xyz, val = w.voxels(w.res)
        lxyz, lval = w.voxels(w.low_res)
        submaps.append({"id": (0, k), "high_bytes": pf.hybrid_grid_bytes(0.1, xyz, val),
                        "low_bytes": pf.hybrid_grid_bytes(0.45, lxyz, lval), "histogram": w.submap_histogram(),
                        "local_pose": (0, 0, 0, 1, 0, 0, 0), "global_pose": (0, 0, 0, 1, 0, 0, 0),
                        "num_range_data": 160, "finished": True, "voxels": (xyz, val, lxyz, lval)})
    for i in range(10):
        s = i % 2
        hi, lo, hist, grav, true = worlds[s].node(np.append(rng.uniform(-2, 2, 2), 0.2), rng.uniform(-1, 1),
                                                  n_high=200, n_low=50, max_range=5.0)
        init = true.copy()
        init[:3] += rng.uniform(-0.4, 0.4, 3) * [1, 1, 0.3]
        init = init.astype(np.float32).astype(np.float64)
        nodes.append({"id": (0, i), "timestamp": 10 + i, "gravity_alignment": grav, "local_pose": init,
                      "global_pose": init, "high": hi, "low": lo, "histogram": hist, "submap": s})
    path = tmp_path / "map3d.pbstream"
    path.write_bytes(pf.pbstream_bytes(pf.mapping_state_messages(submaps, nodes)))
    state = csm.MappingState(str(path), expect_3d=True)
    assert len(state.submaps) == 2 and len(state.nodes) == 10
    gpu_m, oracle_m = [], []
    for got, want in zip(state.submaps, submaps):
        res, _, xyz, val = got["high"]
        lres, _, lxyz, lval = got["low"]
        gpu_m.append(csm.Matcher3D(gpu_ctx, res, xyz, val, lres, lxyz, lval, got["histogram"], opt))
        wx, wv, wlx, wlv = want["voxels"]
        oracle_m.append(ob.OracleFCSM3D(ob.OracleGrid3D(0.1, wx, wv), ob.OracleGrid3D(0.45, wlx, wlv),
                                        np.asarray(want["histogram"], np.float32), opt))
    gpu_nodes, node_poses, search_submap, search_full = [], [], [], []
    for got, want in zip(state.nodes, nodes):
        gpu_nodes.append(csm.Node3D(got["high_resolution_point_cloud"], got["low_resolution_point_cloud"],
                                    got["histogram"], got["gravity_alignment"]))
        node_poses.append(got["global_pose"])
        search_submap.append(want["submap"])
        search_full.append(0)
        assert got["global_pose"].tolist() == [float(v) for v in want["global_pose"]]
    ident = np.tile(ob.IDENTITY_POSE7.astype(np.float64), (len(nodes), 1))
    found, transform, stats, _ = csm.constraint_builder3d_run(
        gpu_m, search_submap, search_full, gpu_nodes, np.array(node_poses), ident, opt, min_score=0.45,
        global_localization_min_score=0.5, pool_threads=3, flush_threshold=4)
    assert stats["searches"] == len(nodes)
    for i, want in enumerate(nodes):
        hi = pf.decode_point_cloud_reference_math(*pf.compress_point_cloud(want["high"])[:2])
        lo = pf.decode_point_cloud_reference_math(*pf.compress_point_cloud(want["low"])[:2])
        r = oracle_m[want["submap"]].match(hi, lo, np.asarray(want["histogram"], np.float32), 0.45,
                                           np.asarray(want["global_pose"], np.float32), ob.IDENTITY_POSE7,
                                           np.asarray(want["gravity_alignment"], np.float64), False)
        assert bool(found[i]) == bool(r.found), i
        if r.found:
            assert np.array_equal(transform[i].astype(np.float32), np.array(r.pose, np.float32))
    assert found.sum() >= 5
