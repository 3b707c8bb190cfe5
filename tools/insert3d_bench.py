#!/usr/bin/env python
"""RangeDataInserter3D throughput (SURVEY 8(f) rank 3, 3D half): scans inserted per second into a
HybridGrid at 0.1 m (a 40 x 40 x 6 m room, 1500 returns per scan, 20 m range, 2 free-space
voxels), through the host API: the reference-signature Insert per scan (voxels up and down every
call) and the device-resident path (one DeviceGrid, every scan inserted, exported once), next to the
CPU on one thread -- the oracle restatement and, where oracle/_ref was built, the reference's own
range_data_inserter_3d.cc (which in this fork also walks the whole grid twice per insertion for a
debugging printout, hgrid_info).  Final voxels compared bit for bit.  One JSON line."""
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
    w = synth3d.World3D(9, size_xy=40.0, height=6.0, n_boxes=40, samples_per_m2=300.0)
    rng = np.random.default_rng(3)
    origins, returns = [], []
    for _ in range(100):
        o = np.array([rng.uniform(-15, 15), rng.uniform(-15, 15), rng.uniform(0.3, 1.5)], np.float32)
        near = np.nonzero(np.linalg.norm(w.points - o, axis=1) < 20.0)[0]
        pick = rng.choice(near, min(1500, len(near)), replace=False)
        origins.append(o)
        returns.append(np.ascontiguousarray(w.points[pick].astype(np.float32)))
    kw = dict(hit_probability=0.55, miss_probability=0.49, num_free_space_voxels=2)
    none_xyz, none_val = np.zeros((0, 3), np.int32), np.zeros(0, np.uint16)
    out = {"workload": "RangeDataInserter3D, HybridGrid at 0.1 m, %d scans of %d returns, 2 free-space voxels"
                       % (len(origins), int(np.mean([len(r) for r in returns])))}
    got = None
    for name, resident, count in (("per_scan_round_trip", False, 20), ("device_resident", True, len(origins))):
        pkg.grid3d_insert(ctx, 0.1, none_xyz, none_val, origins[:2], returns[:2], resident=resident, **kw)
        t0 = time.perf_counter()
        res = pkg.grid3d_insert(ctx, 0.1, none_xyz, none_val, origins[:count], returns[:count], resident=resident, **kw)
        out[name] = {"scans_per_s": count / (time.perf_counter() - t0), "scans": count}
        if resident:
            got = res
    # The resident call above also pays for the box allocation, the export of 345 k voxels and their
    # way back into a host grid once; the cost of one more scan on a live grid is the slope:
    t0 = time.perf_counter()
    pkg.grid3d_insert(ctx, 0.1, none_xyz, none_val, origins * 10, returns * 10, resident=True, **kw)
    t_long = time.perf_counter() - t0
    t_short = out["device_resident"]["scans"] / out["device_resident"]["scans_per_s"]
    out["device_resident"]["marginal_scans_per_s"] = 9 * len(origins) / (t_long - t_short)
    g = ob.OracleGrid3D(0.1)
    t0 = time.perf_counter()
    for o, r in zip(origins, returns):
        g.insert(o, r, **kw)
    out["cpu_port_1_thread"] = {"scans_per_s": len(origins) / (time.perf_counter() - t0)}
    xyz, val = g.cells()
    assert got[0].tobytes() == xyz.tobytes() and got[1].tobytes() == val.tobytes() and got[2] == g.grid_size
    out["voxels_identical"] = True
    out["voxels"] = int(len(val))
    if os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libref.so")):
        import ref_binding as rb
        rg = rb.RefGrid3D(0.1)
        devnull, saved = os.open(os.devnull, os.O_WRONLY), os.dup(1)
        os.dup2(devnull, 1)  # the fork's hgrid_info printout
        t0 = time.perf_counter()
        try:
            for o, r in zip(origins[:20], returns[:20]):
                rg.insert(o, r, **kw)
        finally:
            dt = time.perf_counter() - t0
            os.dup2(saved, 1)
            os.close(devnull)
        out["cpu_reference_1_thread"] = {"scans_per_s": 20 / dt, "scans": 20,
                                         "note": "includes the fork's two full-grid walks per insertion (hgrid_info)"}
    print(json.dumps(out))
    ctx.close()


if __name__ == "__main__":
    main()
