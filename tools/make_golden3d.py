#!/usr/bin/env python
"""Generates tests/golden/csm3d_golden.{npz,json} from the CPU oracle (3D path).

As for 2D the reference ships no golden vectors; what it does state are the inputs
and tolerances of its 3D tests.  Those fixtures are rebuilt here (axes cloud of
fast_correlative_scan_matcher_3d_test.cc:43-56 inserted through the restated
RangeDataInserter3D; the analytic expectation t / hypot(t, 1 - t) of
rotational_scan_matcher_test.cc:45-66 is stored as "reference_rotational") and the
oracle's exact outputs are recorded so that
 * `-m "not gpu"` tests detect any drift of the oracle, and
 * `-m gpu` tests compare the CUDA path with committed expectations.

Run from the repo root:  python tools/make_golden3d.py
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import golden_util as gu  # noqa: E402
import oracle_binding as ob  # noqa: E402
import synth3d  # noqa: E402


def result_doc(r):
    return {"found": int(r.found), "num_scans": int(r.num_scans), "scan_index": int(r.scan_index),
            "offset": [int(v) for v in r.offset], "score_bits": gu.score_bits(r.score),
            "low_bits": gu.score_bits(r.low_resolution_score), "rot_bits": gu.score_bits(r.rotational_score),
            "pose_bits": [gu.score_bits(v) for v in r.pose]}


def main():
    arrays, doc = {}, {"stacks": [], "matches": [], "discrete": []}
    grids = {}

    def add_grid(key, res, xyz, val):
        arrays[key + "_xyz"], arrays[key + "_val"] = xyz, val
        grids[key] = (res, ob.OracleGrid3D(res, xyz, val))

    # -- grids -----------------------------------------------------------------------
    rng = np.random.default_rng(23847)
    xyz = np.unique(rng.integers(-50, 50, size=(1000, 3)), axis=0).astype(np.int32)
    add_grid("random", 2.0, xyz, synth3d.probability_to_value(rng.uniform(0.1, 0.9, len(xyz))))
    w = synth3d.World3D(7, size_xy=6.0, samples_per_m2=150.0)
    add_grid("world_hi", 0.1, *w.voxels(w.res))
    add_grid("world_lo", 0.45, *w.voxels(w.low_res))
    axes = np.array([[d if a == 0 else 0, d if a == 1 else 0, d if a == 2 else 0]
                     for a in range(3) for d in (4.0, 4.5, 5.0, 5.5)], np.float32)
    arrays["axes"] = axes
    fixture_poses = [([0.31, -0.22, 0.4], 0.12), ([-0.55, 0.6, -0.1], -0.17), ([0.0, 0.0, 0.0], 0.0)]
    for k, (t, theta) in enumerate(fixture_poses):
        g = ob.OracleGrid3D(0.05)
        world = synth3d.quat_rotate(synth3d.yaw_quat(theta), axes.astype(np.float64)) + np.array(t)
        g.insert(np.asarray(t, np.float32), world.astype(np.float32))
        add_grid("fixture%d" % k, 0.05, *g.cells())
    add_grid("far_voxel", 0.1, np.array([[500, 500, 500]], np.int32), synth3d.probability_to_value([0.9]))

    # -- stacks: canonical (non-zero bounding box) form of every level -----------------
    for key, depth, frd in [("random", 4, 4), ("random", 6, 3), ("random", 5, 1), ("world_hi", 8, 3),
                            ("fixture0", 6, 6)]:
        res, g = grids[key]
        m = ob.OracleFCSM3D(g, g, np.zeros(4, np.float32), ob.options3d(depth, frd))
        levels = []
        for i in range(depth):
            lo, dense = gu.canonical_level(*m.level(i))
            levels.append({"lo": lo.tolist(), "shape": list(dense.shape), "sha1": gu.sha(dense)})
        doc["stacks"].append({"grid": key, "depth": depth, "frd": frd, "grid_size": g.grid_size,
                              "levels": levels})

    # -- matches -----------------------------------------------------------------------
    def add_match(name, hi_key, lo_key, hist_key, sub_hist_key, grid, low_grid, opt, min_score,
                  node_pose=ob.IDENTITY_POSE7, submap_pose=ob.IDENTITY_POSE7, gravity=(1., 0., 0., 0.),
                  full=False):
        m = ob.OracleFCSM3D(grids[grid][1], grids[low_grid][1], arrays[sub_hist_key], opt)
        r = m.match(arrays[hi_key], arrays[lo_key], arrays[hist_key], min_score,
                    np.asarray(node_pose, np.float32), np.asarray(submap_pose, np.float32), gravity, full)
        doc["matches"].append({
            "name": name, "grid": grid, "low_grid": low_grid, "hi": hi_key, "lo": lo_key, "hist": hist_key,
            "submap_hist": sub_hist_key, "options": [float(v) for v in opt], "min_score": min_score,
            "node_pose": [float(np.float32(v)) for v in node_pose],
            "submap_pose": [float(np.float32(v)) for v in submap_pose],
            "gravity": [float(v) for v in gravity], "full": int(full), "expect": result_doc(r)})
        return r

    arrays["zeros10"] = np.zeros(10, np.float32)
    arrays["far_low"] = np.array([[42.0, 42.0, 42.0]], np.float32)
    fixture_opt = ob.options3d(6, 6, 0.1, 0.15, 0.8, 0.8, 0.3)
    for k in range(len(fixture_poses)):
        r = add_match("fixture%d" % k, "axes", "axes", "zeros10", "zeros10", "fixture%d" % k, "fixture%d" % k,
                      fixture_opt, 0.1)
        assert r.found
        r = add_match("fixture%d_far" % k, "axes", "far_low", "zeros10", "zeros10", "fixture%d" % k,
                      "fixture%d" % k, fixture_opt, 0.1)
        assert not r.found
    r = add_match("fixture0_full", "axes", "axes", "zeros10", "zeros10", "fixture0", "fixture0", fixture_opt,
                  0.1, full=True)
    assert r.found
    arrays["world_hist"] = w.submap_histogram()
    nrng = np.random.default_rng(77)
    for k in range(5):
        pos = np.append(nrng.uniform(-1.5, 1.5, 2), nrng.uniform(-0.2, 0.5))
        yaw = nrng.uniform(-1, 1)
        gyaw = nrng.uniform(-0.2, 0.2)
        hi, lo, hist, grav, true = w.node(pos, yaw, max_range=4.0, n_high=150 + 11 * k, n_low=40 + k,
                                          gravity_yaw=gyaw, tilt=(0.01 * k, -0.02))
        arrays["node%d_hi" % k], arrays["node%d_lo" % k], arrays["node%d_hist" % k] = hi, lo, hist
        init = true.copy()
        init[:3] += nrng.uniform(-0.5, 0.5, 3) * [1, 1, 0.3]
        init[3:] = synth3d.quat_mul(synth3d.yaw_quat(nrng.uniform(-0.1, 0.1)), true[3:])
        if k == 4:
            init[:3] += [30.0, 0.0, 0.0]
        submap = synth3d.pose7([0.5, 0.25, -0.125], synth3d.yaw_quat(0.2)) if k % 2 else ob.IDENTITY_POSE7
        if k % 2:  # express the node pose in a moved global frame
            init = synth3d.pose7(synth3d.quat_rotate(submap[3:], init[:3][None])[0] + submap[:3],
                                 synth3d.quat_mul(submap[3:], init[3:]))
        opt = ob.options3d(6 + k % 2, 3 - k % 2, 0.0 if k == 0 else 0.3, 0.4, 1.2, 0.4, np.radians(10))
        add_match("world_node%d" % k, "node%d_hi" % k, "node%d_lo" % k, "node%d_hist" % k, "world_hist",
                  "world_hi", "world_lo", opt, 0.45, init, submap, grav)
        if k == 0:
            m = ob.OracleFCSM3D(grids["world_hi"][1], grids["world_lo"][1], arrays["world_hist"], opt)
            poses, rot, cells = m.discrete_scans(hi, hist, init.astype(np.float32),
                                                 np.asarray(submap, np.float32), grav)
            doc["discrete"].append({"match": len(doc["matches"]) - 1, "num_scans": int(len(rot)),
                                    "poses_sha1": gu.sha(poses), "rot_sha1": gu.sha(rot),
                                    "cells_sha1": gu.sha(cells), "cells_first": cells[0, :3].tolist()})
    hi, lo, hist, grav, true = w.node([0.6, -0.3, 0.1], 0.4, max_range=4.0, n_high=120, n_low=30)
    arrays["full_hi"], arrays["full_lo"], arrays["full_hist"] = hi, lo, hist
    add_match("world_full", "full_hi", "full_lo", "full_hist", "world_hist", "world_hi", "world_lo",
              ob.options3d(7, 3, 0.5, 0.4, 1.2, 0.4, np.radians(10)), 0.45,
              synth3d.pose7([0, 0, 0], true[3:]), full=True)
    trng = np.random.default_rng(5)
    arrays["tie_hi"] = trng.uniform(-2, 2, (40, 3)).astype(np.float32)
    arrays["tie_lo"] = trng.uniform(-2, 2, (9, 3)).astype(np.float32)
    arrays["ones12"] = np.ones(12, np.float32)
    r = add_match("tie_order", "tie_hi", "tie_lo", "ones12", "ones12", "far_voxel", "far_voxel",
                  ob.options3d(4, 2, 0.0, 0.05, 1.0, 0.4, np.radians(20)), 0.05)
    assert r.found

    # -- the reference's analytic rotational expectation ----------------------------------
    doc["reference_rotational"] = {"buckets": 10, "submap_unit": 3, "scan_unit": 2,
                                   "t": [0.1 * k for k in range(10)], "tolerance": 1e-6}

    out = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out, exist_ok=True)
    np.savez_compressed(os.path.join(out, "csm3d_golden.npz"), **arrays)
    with open(os.path.join(out, "csm3d_golden.json"), "w") as f:
        json.dump(doc, f, indent=1, sort_keys=True)
    found = sum(m["expect"]["found"] for m in doc["matches"])
    print("wrote", len(doc["stacks"]), "stacks,", len(doc["matches"]), "matches (%d found)," % found,
          len(doc["discrete"]), "discrete-scan cases")


if __name__ == "__main__":
    main()
