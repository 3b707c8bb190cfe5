#!/usr/bin/env python
"""TSDFRangeDataInserter2D throughput: scans inserted per second into a TSDF2D
at 5 cm (~360 returns per scan, 10 m range, trajectory_builder_2d.lua defaults), through the host API:
the reference-signature Insert per scan (both cell arrays up and down every call) and the
device-resident path, next to the CPU on one thread (the restatement and, where oracle/_ref was
built, the reference's own source); final cells compared bit for bit.  One JSON line."""
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
    for i in range(100):
        pose, pts = synth.make_scan(w, 6000 + i)
        c, s = np.cos(pose[2]), np.sin(pose[2])
        returns.append(np.stack([pose[0] + c * pts[:, 0] - s * pts[:, 1], pose[1] + s * pts[:, 0] + c * pts[:, 1],
                                 np.zeros(len(pts))], axis=1).astype(np.float32))
        origins.append(np.array([pose[0], pose[1]], np.float32))
    zero = np.zeros((1000, 1000), np.uint16)
    o9 = ob.tsdf_inserter_options()
    out = {"workload": "TSDFRangeDataInserter2D, 1000x1000 TSDF2D at 5 cm, %d scans of ~%d returns, default options"
                       % (len(origins), int(np.mean([len(r) for r in returns])))}

    def run(count, resident, reps=1):
        g = pkg.TsdfGrid(zero, zero, w.res, w.max_x, w.max_y, 0.3, 10.0)
        t0 = time.perf_counter()
        res = pkg.tsdf_insert(ctx, g, (origins * reps)[:count], (returns * reps)[:count], o9, resident=resident)
        return time.perf_counter() - t0, res

    run(2, False), run(2, True)
    t, _ = run(20, False)
    out["per_scan_round_trip"] = {"scans_per_s": 20 / t, "scans": 20}
    t_short, got = run(100, True)
    t_long, _ = run(1000, True, reps=10)
    out["device_resident"] = {"scans_per_s": 100 / t_short, "scans": 100,
                              "marginal_scans_per_s": 900 / (t_long - t_short)}
    og = ob.OracleTsdf(zero, zero, w.res, w.max_x, w.max_y, 0.3, 10.0)
    t0 = time.perf_counter()
    for o, r in zip(origins, returns):
        ob.tsdf_insert(og, o, r)
    out["cpu_port_1_thread"] = {"scans_per_s": len(origins) / (time.perf_counter() - t0),
                                "note": "includes reading both cell arrays back per scan (4 MB)"}
    assert got[0].tobytes() == og.tsd_cells.tobytes() and got[1].tobytes() == og.weight_cells.tobytes()
    out["cells_identical"] = True
    out["cells_updated"] = int((og.weight_cells > 0).sum())
    if os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libref.so")):
        import ref_binding as rb
        rg = rb.RefTsdf(zero, zero, w.res, w.max_x, w.max_y, 0.3, 10.0)
        t0 = time.perf_counter()
        for o, r in zip(origins, returns):
            rb.tsdf_insert(rg, o, r)
        out["cpu_reference_1_thread"] = {"scans_per_s": len(origins) / (time.perf_counter() - t0),
                                         "note": "includes reading both cell arrays back per scan through ToProto"}
        assert rg.tsd_cells.tobytes() == og.tsd_cells.tobytes()
    print(json.dumps(out))
    ctx.close()


if __name__ == "__main__":
    main()
