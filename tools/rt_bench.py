#!/usr/bin/env python
"""C2: RealTimeCorrelativeScanMatcher2D throughput (scans/s) at batch sizes 1, 64, 1024 and
single-call latency, through the host API (host buffers in, poses out), with the CPU
oracle timed on a bounded sample beside it.  One JSON line."""
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


def main():
    pkg = _load_package()
    ctx = pkg.Context(0)
    w = synth.World(400, 400, seed=2000)
    g = pkg.Grid(w.cells, w.res, w.max_x, w.max_y)
    rng = np.random.default_rng(2)
    clouds, inits = [], []
    for i in range(1024):
        pose, pts = synth.make_scan(w, 2000 + (i % 128))
        pts = np.ascontiguousarray(pts[::max(1, len(pts) // 250)][:250])
        clouds.append(pts)
        inits.append((pose[0] + rng.uniform(-0.08, 0.08), pose[1] + rng.uniform(-0.08, 0.08),
                      pose[2] + rng.uniform(-0.26, 0.26)))
    out = {"workload": "C2 RT-CSM 2D: 400x400 live grid, ~250 points, +-0.1 m, +-20 deg, weights 0.1"}
    for b in (1, 64, 1024):
        reps = {1: 200, 64: 20, 1024: 3}[b]
        pkg.rt_match_batch(ctx, [g] * b, clouds[:b], inits[:b])
        t0 = time.perf_counter()
        for _ in range(reps):
            res = pkg.rt_match_batch(ctx, [g] * b, clouds[:b], inits[:b])
        dt = (time.perf_counter() - t0) / reps
        out["batch_%d" % b] = {"scans_per_s": b / dt, "ms_per_call": 1e3 * dt}
    og = ob.OracleGrid(w.cells, w.res, w.max_x, w.max_y)
    opts = {"linear_search_window": 0.1, "angular_search_window": float(np.deg2rad(20.0)),
            "translation_delta_cost_weight": 0.1, "rotation_delta_cost_weight": 0.1}
    t0 = time.perf_counter()
    n = 64
    for i in range(n):
        r = ob.rt_match(og, opts, inits[i], clouds[i])
        assert (r.scan_index, r.x_index_offset, r.y_index_offset) == \
            (res[i].scan_index, res[i].x_index_offset, res[i].y_index_offset)
    out["cpu_port_1_thread"] = {"scans_per_s": n / (time.perf_counter() - t0), "sample": n}
    # The same scans against a TSDF2D of the same world (RT2D.cc:38-59).
    import tsdf_util
    tv, wv = tsdf_util.from_world(w, 0.3, 10.0, seed=3)
    tg = pkg.TsdfGrid(tv, wv, w.res, w.max_x, w.max_y, 0.3, 10.0)
    otg = ob.OracleTsdf(tv, wv, w.res, w.max_x, w.max_y, 0.3, 10.0)
    pkg.rt_match_batch(ctx, [tg] * 64, clouds[:64], inits[:64])
    t0 = time.perf_counter()
    for _ in range(20):
        tres = pkg.rt_match_batch(ctx, [tg] * 64, clouds[:64], inits[:64])
    out["tsdf_batch_64"] = {"scans_per_s": 64 * 20 / (time.perf_counter() - t0)}
    t0 = time.perf_counter()
    for i in range(32):
        r = ob.rt_match_tsdf(otg, opts, inits[i], clouds[i])
        assert (r.scan_index, r.x_index_offset, r.y_index_offset) == \
            (tres[i].scan_index, tres[i].x_index_offset, tres[i].y_index_offset)
    out["tsdf_cpu_port_1_thread"] = {"scans_per_s": 32 / (time.perf_counter() - t0), "sample": 32}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
