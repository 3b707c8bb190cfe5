#!/usr/bin/env python
"""C4: PrecomputationGridStack2D build time on the device (cells resident in HBM).
Prints one JSON line per grid size: algorithmic GB/s = (2*W*H + sum of level bytes) / time."""
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from __graft_entry__ import _load_package  # noqa: E402
import synth  # noqa: E402


def main():
    import torch
    pkg = _load_package()
    ctx = pkg.Context(0)
    abi = pkg.abi()
    stream = torch.cuda.ExternalStream(abi.csm_ctx_stream(C.c_void_p(ctx.raw)))
    depth = 7
    peak = 6549.1
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = json.load(open(p))["hbm_gbs"]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for size in (400, 1000, 2000):
        w = synth.World(size, size, seed=4000 + size)
        g = pkg.Grid(w.cells, w.res, w.max_x, w.max_y)
        m = pkg.Matcher(ctx, g, 7.0, 0.5, depth)
        cells = torch.from_numpy(w.cells.view(np.int16)).cuda()
        ptr = C.c_void_p(cells.data_ptr())
        for _ in range(10):
            pkg.host().csmh_matcher_rebuild_dev(ctx.h, m.h, ptr)
        torch.cuda.synchronize()
        reps, warm_ms, cold_ms = 100, 0.0, 0.0
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(reps):
            pkg.host().csmh_matcher_rebuild_dev(ctx.h, m.h, ptr)
        e1.record(stream)
        e1.synchronize()
        warm_ms = e0.elapsed_time(e1) / reps
        for _ in range(10):  # cold: L2 flushed before every build
            flush.fill_(1)
            torch.cuda.synchronize()
            e0.record(stream)
            pkg.host().csmh_matcher_rebuild_dev(ctx.h, m.h, ptr)
            e1.record(stream)
            e1.synchronize()
            cold_ms += e0.elapsed_time(e1) / 10
        alg = 2 * size * size + sum((size + 2 ** i - 1) ** 2 for i in range(depth))
        quad = sum(4 * ((size + 2 * 2 ** i - 1 + 2 * 2 ** i - 1) // (2 * 2 ** i) * (2 * 2 ** i)) ** 2
                   for i in range(depth))
        print(json.dumps({
            "workload": "C4 PrecomputationGridStack2D build %dx%d depth %d" % (size, size, depth),
            "mode": os.environ.get("CSM_STACK_BUILD", "rows"),
            "ms_back_to_back": warm_ms, "ms_l2_flushed": cold_ms,
            "algorithmic_bytes": alg, "quad_layout_bytes_written": quad,
            "algorithmic_gbs": alg / (cold_ms * 1e6), "frac_of_measured_hbm_peak": alg / (cold_ms * 1e6) / peak,
            "all_bytes_gbs": (alg + quad) / (cold_ms * 1e6)}))


if __name__ == "__main__":
    main()
