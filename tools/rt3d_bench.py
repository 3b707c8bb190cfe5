#!/usr/bin/env python
"""RealTimeCorrelativeScanMatcher3D throughput: scans matched per second
through the host API on a synthetic 3D room (0.1 m HybridGrid voxels), trajectory_builder_3d.lua
window (0.15 m, 1 degree, weights 0.1), ~1500-point clouds, at batch sizes 1 and 16, with the CPU
restatement on one thread beside it (results compared bit for bit).  One JSON line."""
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
import synth3d  # noqa: E402


def main():
    pkg = _load_package()
    ctx = pkg.Context(0)
    w = synth3d.World3D(81, size_xy=20.0)
    xyz, val = w.voxels(w.res)
    opt = pkg.rt_options(0.15, np.radians(1.0), 1e-1, 1e-1)
    rng = np.random.default_rng(82)
    clouds, inits = [], []
    for k in range(16):
        hi, lo, hist, grav, true = w.node(np.append(rng.uniform(-5, 5, 2), 0.2), rng.uniform(-1, 1),
                                          n_high=1500, n_low=100, max_range=15.0, tilt=(0.01, -0.02))
        init = true.copy()
        init[:3] += rng.uniform(-0.12, 0.12, 3)
        init[3:] = synth3d.quat_mul(synth3d.yaw_quat(rng.uniform(-0.01, 0.01)), true[3:])
        clouds.append(hi)
        inits.append(init.astype(np.float32).astype(np.float64))
    out = {"workload": "RT-CSM 3D: %d voxels at 0.1 m, ~%d points, 0.15 m / 1 deg window" %
                       (len(val), int(np.mean([len(c) for c in clouds])))}
    for b in (1, 16):
        reps = {1: 10, 16: 3}[b]
        pkg.rt_match3d_batch(ctx, 0.1, xyz, val, opt, np.array(inits[:b]), clouds[:b])
        t0 = time.perf_counter()
        for _ in range(reps):
            scores, poses, best, num = pkg.rt_match3d_batch(ctx, 0.1, xyz, val, opt, np.array(inits[:b]), clouds[:b])
        dt = (time.perf_counter() - t0) / reps
        out["batch_%d" % b] = {"scans_per_s": b / dt, "ms_per_call": 1e3 * dt,
                               "candidates_per_scan": int(num[0]),
                               "note": "includes the upload of the voxel list and the dense-box build every call"}
    dg = pkg.DeviceGrid3D(ctx, 0.1, xyz, val)
    for b in (1, 16):
        reps = {1: 20, 16: 5}[b]
        dg.match(opt, np.array(inits[:b]), clouds[:b])
        t0 = time.perf_counter()
        for _ in range(reps):
            s2, p2, b2, n2 = dg.match(opt, np.array(inits[:b]), clouds[:b])
        dt = (time.perf_counter() - t0) / reps
        out["resident_batch_%d" % b] = {"scans_per_s": b / dt, "ms_per_call": 1e3 * dt,
                                        "note": "grid resident on the device (DeviceHybridGrid)"}
    assert np.array_equal(b2, best) and np.array_equal(s2, scores)
    og = ob.OracleGrid3D(0.1, xyz, val)
    t0 = time.perf_counter()
    n = 2
    for k in range(n):
        score_o, pose_o, best_o, num_o = ob.rt_match3d(og, opt, inits[k].astype(np.float32), clouds[k])
        assert best[k] == best_o and np.float32(scores[k]).tobytes() == np.float32(score_o).tobytes()
    out["cpu_port_1_thread"] = {"scans_per_s": n / (time.perf_counter() - t0), "sample": n}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
