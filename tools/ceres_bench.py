#!/usr/bin/env python
"""CeresScanMatcher2D refinement throughput: poses refined per second
through the host API (host buffers in, poses out) at batch sizes 1, 100, 1000 on a 1000x1000
submap with ~280-point scans and the pose_graph.lua options, next to the CPU restatement on
one thread (bounded sample, results compared at 1e-6).  One JSON line."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from __graft_entry__ import _load_package  # noqa: E402
import oracle_binding as ob  # noqa: E402
import synth  # noqa: E402

OPTS = {"occupied_space_weight": 20.0, "translation_weight": 10.0, "rotation_weight": 1.0,
        "use_nonmonotonic_steps": True, "max_num_iterations": 10}


def main():
    pkg = _load_package()
    ctx = pkg.Context(0)
    w = synth.World(1000, 1000, seed=5000)
    g = pkg.Grid(w.cells, w.res, w.max_x, w.max_y)
    rng = np.random.default_rng(6)
    clouds, inits = [], []
    for i in range(1000):
        truth, pts = synth.make_scan(w, 5000 + (i % 100))
        clouds.append(np.ascontiguousarray(pts))
        inits.append((truth[0] + rng.uniform(-0.05, 0.05), truth[1] + rng.uniform(-0.05, 0.05),
                      truth[2] + rng.uniform(-0.02, 0.02)))
    targets = [(p[0], p[1]) for p in inits]
    out = {"workload": "CeresScanMatcher2D refinement, 1000x1000 grid at 5 cm, ~%d points, pose_graph.lua "
                       "options (20 / 10 / 1, non-monotonic, 10 iterations)" % int(np.mean([len(c) for c in clouds]))}
    for b in (1, 100, 1000):
        reps = {1: 50, 100: 10, 1000: 3}[b]
        pkg.ceres_match_batch(ctx, [g] * b, clouds[:b], targets[:b], inits[:b], **OPTS)
        t0 = time.perf_counter()
        for _ in range(reps):
            poses, summ = pkg.ceres_match_batch(ctx, [g] * b, clouds[:b], targets[:b], inits[:b], **OPTS)
        dt = (time.perf_counter() - t0) / reps
        out["batch_%d" % b] = {"poses_per_s": b / dt, "ms_per_call": 1e3 * dt,
                               "mean_iterations": float(np.mean(summ[:, 2]))}
    og = ob.OracleGrid(w.cells, w.res, w.max_x, w.max_y)
    n = 64
    t0 = time.perf_counter()
    worst = 0.0
    for i in range(n):
        want, _ = ob.ceres_match(og, OPTS, targets[i], inits[i], clouds[i])
        worst = max(worst, float(np.max(np.abs(want - poses[i]))))
    out["cpu_port_1_thread"] = {"poses_per_s": n / (time.perf_counter() - t0), "sample": n}
    out["max_abs_pose_difference_vs_cpu_port"] = worst
    assert worst < 1e-6
    print(json.dumps(out))


if __name__ == "__main__":
    main()
