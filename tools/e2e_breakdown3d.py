#!/usr/bin/env python
"""Where the end-to-end time of a C5 batch goes (host prep + upload, kernels, fetch, free)."""
import sys, os, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
sys.path.insert(0, ROOT)
import bench3d


def main():
    import torch
    from __graft_entry__ import _load_package
    pkg = _load_package()
    a = bench3d.parse()
    w, (hx, hv, lx, lv), nodes = bench3d.make_workload(a)
    ctx = pkg.Context(0)
    m = pkg.Matcher3D(ctx, 0.1, hx, hv, 0.45, lx, lv, w.submap_histogram(), bench3d.OPT)
    n = len(nodes)
    objs = [pkg.Node3D(hi, lo, h_, g) for hi, lo, h_, g, _, _, _ in nodes]
    inits = np.array([x[4] for x in nodes])
    full = np.array([1 if x[5] else 0 for x in nodes], np.int32)
    mins = [bench3d.MIN_SCORE] * n
    acc = np.zeros(4)
    reps = 10
    for r in range(reps + 3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        b = pkg.PreparedBatch3D(ctx, [m] * n, objs, mins, inits, None, full)
        t1 = time.perf_counter()
        b.run()
        pkg.abi().csm_ctx_synchronize(ctx.raw)
        t2 = time.perf_counter()
        b.fetch()
        t3 = time.perf_counter()
        b.close()
        t4 = time.perf_counter()
        if r >= 3:
            acc += [t1 - t0, t2 - t1, t3 - t2, t4 - t3]
    print("ms per %d-search batch: prepare+upload %.2f, run %.2f, fetch %.2f, destroy %.2f" %
          ((n,) + tuple(acc / reps * 1e3)))


if __name__ == "__main__":
    main()
