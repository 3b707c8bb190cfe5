#!/usr/bin/env python
"""ProbabilityGridRangeDataInserter2D throughput: scans inserted per second
into a 1000x1000 grid at 5 cm (~360 returns per scan, 10 m range, free space inserted), through the
host API: the reference-signature Insert per scan (grid up and down every call) and the
device-resident path (grid uploaded once, all scans inserted, downloaded once), next to the CPU
restatement on one thread; final cells compared bit for bit.  One JSON line."""
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
    w = synth.World(1000, 1000, seed=6000)
    origins, returns = [], []
    for i in range(200):
        pose, pts = synth.make_scan(w, 6000 + i)
        c, s = np.cos(pose[2]), np.sin(pose[2])
        returns.append(np.stack([pose[0] + c * pts[:, 0] - s * pts[:, 1],
                                 pose[1] + s * pts[:, 0] + c * pts[:, 1], np.zeros(len(pts))],
                                axis=1).astype(np.float32))
        origins.append((np.float32(pose[0]), np.float32(pose[1]), 0.0))
    empty = np.zeros((1000, 1000), np.uint16)
    out = {"workload": "ProbabilityGridRangeDataInserter2D, 1000x1000 grid at 5 cm, %d scans of ~%d returns, "
                       "insert_free_space" % (len(origins), int(np.mean([len(r) for r in returns])))}
    for name, resident in (("per_scan_round_trip", False), ("device_resident", True)):
        g = pkg.Grid(empty, w.res, w.max_x, w.max_y)
        pkg.grid_insert(ctx, g, origins[:2], returns[:2], resident=resident)
        g = pkg.Grid(empty, w.res, w.max_x, w.max_y)
        t0 = time.perf_counter()
        cells, _, _ = pkg.grid_insert(ctx, g, origins, returns, resident=resident)
        out[name] = {"scans_per_s": len(origins) / (time.perf_counter() - t0)}
    og = ob.OracleGrid(empty, w.res, w.max_x, w.max_y)
    t0 = time.perf_counter()
    for o, r in zip(origins, returns):
        ob.grid_insert(og, o[:2], r)
    out["cpu_port_1_thread"] = {"scans_per_s": len(origins) / (time.perf_counter() - t0)}
    assert np.array_equal(cells, og.cells)
    out["cells_identical"] = True
    print(json.dumps(out))


if __name__ == "__main__":
    main()
