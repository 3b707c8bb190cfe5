#!/usr/bin/env python
"""Prints the branch-and-bound tile occupancy histogram of the bench workload
(diagnostic; needs a GPU)."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
from __graft_entry__ import _load_package  # noqa: E402


def main():
    sys.argv = [sys.argv[0]] + sys.argv[1:]
    args = bench.parse()
    pkg = _load_package()
    worlds, clouds, target = bench.make_workload(args, 0)
    ctx = pkg.Context(0)
    grids = [pkg.Grid(w.cells, w.res, w.max_x, w.max_y) for w in worlds]
    matchers = [pkg.Matcher(ctx, g, bench.LINEAR, bench.ANGULAR, bench.DEPTH) for g in grids]
    batch = pkg.PreparedBatch(ctx, [matchers[t] for t in target], clouds,
                              [bench.MIN_SCORE] * len(clouds))
    abi = pkg.abi()
    batch.run()
    batch.fetch()
    abi.csm_ctx_set_profiling(C.c_void_p(ctx.raw), 2)
    batch.run()
    res = batch.fetch()
    hist = (C.c_double * (16 * 33))()
    abi.csm_ctx_get_tile_histogram(C.c_void_p(ctx.raw), hist)
    h = np.array(list(hist)).reshape(16, 33)
    print("searches", len(clouds), "found", sum(1 for r in res if r.found))
    for d in range(16):
        if h[d].sum() == 0:
            continue
        tiles = h[d].sum()
        alive = (h[d] * np.arange(33)).sum()
        bins = [h[d][1], h[d][2], h[d][3:5].sum(), h[d][5:9].sum(), h[d][9:17].sum(), h[d][17:].sum()]
        print("depth %d: tiles %9d alive %10d mean %.2f | 1:%d 2:%d 3-4:%d 5-8:%d 9-16:%d 17-32:%d" %
              (d, tiles, alive, alive / tiles, *bins))
    ph = (C.c_double * 5)()
    sc = (C.c_double * 5)()
    runs = C.c_int()
    abi.csm_ctx_get_profile(C.c_void_p(ctx.raw), ph, sc, C.byref(runs))
    print("phase ms", list(ph), "scored", list(sc))
    import hashlib
    sig = hashlib.sha1(repr([(r.found, r.scan_index, r.x_index_offset, r.y_index_offset, r.score)
                             for r in res]).encode()).hexdigest()[:12]
    print("result signature", sig)
    per = [(r.found, r.candidates_scored) for r in res]
    pos = [c for f, c in per if f]
    neg = [c for f, c in per if not f]
    print("scored/search found: mean %.0f  not found: mean %.0f max %.0f" %
          (np.mean(pos) if pos else 0, np.mean(neg) if neg else 0, max(neg) if neg else 0))


if __name__ == "__main__":
    main()
