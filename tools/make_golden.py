#!/usr/bin/env python
"""Generates tests/golden/csm2d_golden.npz + csm2d_golden.json from the CPU oracle.

The reference ships no golden vectors for this path; these fixtures
are produced by oracle/ (the restatement pinned by the reference's own known-answer
tests, oracle/oracle_tests.cc) on fixed seeds, inputs included, so that
 * `-m "not gpu"` tests detect any drift of the oracle itself, and
 * `-m gpu` tests compare the CUDA path with committed expectations without
   depending on the generator.
The one vector that IS the reference's own is the DiscretizeScans fixture of
cartographer/mapping/internal/2d/scan_matching/correlative_scan_matcher_test.cc:72-96
(exact expected cell indices), stored under "reference_discretize".

Run from the repo root:  python tools/make_golden.py
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_binding as ob  # noqa: E402
import synth  # noqa: E402


def sha(a):
    return hashlib.sha1(np.ascontiguousarray(a).tobytes()).hexdigest()


def score_bits(x):
    return int(np.float32(x).view(np.uint32))


def main():
    arrays, doc = {}, {"stacks": [], "discretize": [], "matches": [], "rt": []}

    # -- stacks: random grids incl. width > grid and non-square shapes ---------------
    for k, (nx, ny, depth) in enumerate([(97, 61, 6), (5, 3, 7), (128, 128, 5), (33, 150, 4)]):
        rng = np.random.default_rng(100 + k)
        cells = rng.integers(1, 32768, size=(ny, nx)).astype(np.uint16)
        cells[rng.random((ny, nx)) < 0.3] = 0
        cells[rng.random((ny, nx)) < 0.02] |= 0x8000  # update marker bit is ignored
        arrays["stack%d" % k] = cells
        m = ob.OracleFCSM2D(ob.OracleGrid(cells, 0.05, 1.0, 1.0), 1.0, 0.1, depth)
        doc["stacks"].append({"cells": "stack%d" % k, "depth": depth,
                              "levels": [{"shape": list(m.level(i).shape), "sha1": sha(m.level(i))}
                                         for i in range(depth)]})

    # -- the reference's own DiscretizeScans fixture ---------------------------------
    pts = np.array([[0.025, 0.175, 0], [-0.025, 0.175, 0], [-0.075, 0.175, 0], [-0.125, 0.175, 0],
                    [-0.125, 0.125, 0], [-0.125, 0.075, 0], [-0.125, 0.025, 0]], dtype=np.float32)
    doc["reference_discretize"] = {
        "points": pts.tolist(), "resolution": 0.05, "max": [0.05, 0.25], "cells": [6, 6],
        "expected_xy": [[1, 0], [1, 1], [1, 2], [1, 3], [2, 3], [3, 3], [4, 3]]}

    # -- worlds shared by discretize / match / rt cases --------------------------------
    worlds = [synth.World(160, 128, seed=500), synth.World(96, 96, seed=501),
              synth.World(200, 200, seed=502)]
    for k, w in enumerate(worlds):
        arrays["world%d" % k] = w.cells
    clouds = []
    for k in range(10):
        w = worlds[k % 3]
        pose, c = synth.make_scan(w, 600 + k, max_range=5.0)
        arrays["cloud%d" % k] = c
        clouds.append((k % 3, pose, c))

    # -- discretize + ShrinkToFit with rotation and translation ------------------------
    for k in (0, 1, 2, 3):
        wi, pose, c = clouds[k]
        w = worlds[wi]
        init = (pose[0] + 0.31, pose[1] - 0.17, pose[2] + 0.23)
        nap, step, ns, nlin = ob.search_parameters(2.0, 0.4, c, w.res)
        xy, bounds = ob.discretize(w.res, w.max_x, w.max_y, w.nx, w.ny, c, init, nlin, nap, step)
        doc["discretize"].append({"world": wi, "cloud": k, "init": list(init), "linear": 2.0,
                                  "angular": 0.4, "num_scans": ns, "xy_sha1": sha(xy),
                                  "bounds_sha1": sha(bounds), "xy_first": xy[0, :4].tolist(),
                                  "bounds_first": bounds[0].tolist()})

    # -- FastCorrelativeScanMatcher2D: full-submap and local searches -------------------
    def add_match(wi, cells_key, res, max_x, max_y, cells, cloud_key, c, depth, lin, ang, min_score,
                  init):
        m = ob.OracleFCSM2D(ob.OracleGrid(cells, res, max_x, max_y), lin, ang, depth)
        r = m.match(c, min_score, initial=init)
        doc["matches"].append({
            "cells": cells_key, "res": res, "max": [max_x, max_y], "cloud": cloud_key,
            "depth": depth, "linear": lin, "angular": ang, "min_score": min_score,
            "init": None if init is None else list(init), "found": int(r.found),
            "winner": [r.scan_index, r.x_index_offset, r.y_index_offset],
            "score_bits": score_bits(r.score), "pose": [r.pose_x, r.pose_y, r.pose_theta]})

    for k, (wi, pose, c) in enumerate(clouds):
        w = worlds[wi]
        depth = [5, 4, 6][wi]
        add_match(wi, "world%d" % wi, w.res, w.max_x, w.max_y, w.cells, "cloud%d" % k, c, depth,
                  7.0, float(np.deg2rad(30.0)), 0.5 if k % 2 else 0.2, None)
        init = (pose[0] + 0.4, pose[1] - 0.3, pose[2] + 0.15)
        add_match(wi, "world%d" % wi, w.res, w.max_x, w.max_y, w.cells, "cloud%d" % k, c, depth,
                  2.0, 0.5, 0.3, init)
    # tie stress: empty grid (every candidate ties), single point; symmetric corridor
    empty = np.zeros((110, 100), dtype=np.uint16)
    arrays["empty"] = empty
    one = np.array([[0.1, 0.2, 0.3]], dtype=np.float32)
    arrays["one_point"] = one
    for depth in (1, 3, 7):
        add_match(-1, "empty", 1.0, 2.0, 3.0, empty, "one_point", one, depth, 7.0,
                  float(np.deg2rad(30.0)), 0.0, None)
        add_match(-1, "empty", 1.0, 2.0, 3.0, empty, "one_point", one, depth, 7.0,
                  float(np.deg2rad(30.0)), 0.0, (4.0, 5.0, 0.0))
    corridor = np.zeros((200, 200), dtype=np.uint16)
    wall = synth.probability_to_value(np.float32(0.8))
    corridor[60:62, :] = wall
    corridor[140:142, :] = wall
    arrays["corridor"] = corridor
    ys = np.linspace(-3, 3, 40, dtype=np.float32)
    cc = np.concatenate([np.stack([np.full(40, 2.0, np.float32), ys, np.zeros(40, np.float32)], 1),
                         np.stack([np.full(40, -2.0, np.float32), ys, np.zeros(40, np.float32)], 1)])
    arrays["corridor_cloud"] = cc
    add_match(-1, "corridor", 0.05, 5.0, 5.0, corridor, "corridor_cloud", cc, 6, 7.0,
              float(np.deg2rad(30.0)), 0.3, None)

    # -- RealTimeCorrelativeScanMatcher2D ------------------------------------------------
    for k in (0, 1, 2, 4, 5):
        wi, pose, c = clouds[k]
        w = worlds[wi]
        opts = {"linear_search_window": 0.1 + 0.05 * (k % 2), "angular_search_window": 0.35,
                "translation_delta_cost_weight": 0.1 * k, "rotation_delta_cost_weight": 0.1}
        init = (pose[0] + 0.06, pose[1] - 0.04, pose[2] + 0.2)
        r = ob.rt_match(ob.OracleGrid(w.cells, w.res, w.max_x, w.max_y), opts, init, c)
        doc["rt"].append({"world": wi, "cloud": k, "options": opts, "init": list(init),
                          "winner": [r.scan_index, r.x_index_offset, r.y_index_offset],
                          "score_bits": score_bits(r.score),
                          "pose": [r.pose_x, r.pose_y, r.pose_theta]})
    doc["worlds"] = [{"res": w.res, "max": [w.max_x, w.max_y]} for w in worlds]

    out = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out, exist_ok=True)
    np.savez_compressed(os.path.join(out, "csm2d_golden.npz"), **arrays)
    with open(os.path.join(out, "csm2d_golden.json"), "w") as f:
        json.dump(doc, f, indent=1, sort_keys=True)
    print("wrote", len(doc["stacks"]), "stacks,", len(doc["discretize"]), "discretize,",
          len(doc["matches"]), "matches,", len(doc["rt"]), "rt cases")


if __name__ == "__main__":
    main()
