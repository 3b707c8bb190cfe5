#!/usr/bin/env python
"""Wall-clock breakdown of one end-to-end bench step (diagnostic; needs a GPU)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
from __graft_entry__ import _load_package  # noqa: E402


def main():
    args = bench.parse()
    pkg = _load_package()
    worlds, clouds, target = bench.make_workload(args, 0)
    ctx = pkg.Context(0)
    grids = [pkg.Grid(w.cells, w.res, w.max_x, w.max_y) for w in worlds]
    n = len(clouds)
    for it in range(4):
        t0 = time.perf_counter()
        ms = [pkg.Matcher(ctx, g, bench.LINEAR, bench.ANGULAR, bench.DEPTH) for g in grids]
        t1 = time.perf_counter()
        b = pkg.PreparedBatch(ctx, [ms[t] for t in target], clouds, [bench.MIN_SCORE] * n)
        t2 = time.perf_counter()
        b.run()
        t3 = time.perf_counter()
        b.fetch()
        t4 = time.perf_counter()
        b.close()
        del ms
        t5 = time.perf_counter()
        print("iter %d: matchers %.2f ms, prepare+upload %.2f ms, run(enqueue) %.2f ms, fetch(wait) %.2f ms, "
              "destroy %.2f ms, total %.2f ms" % (it, 1e3 * (t1 - t0), 1e3 * (t2 - t1), 1e3 * (t3 - t2),
                                                  1e3 * (t4 - t3), 1e3 * (t5 - t4), 1e3 * (t5 - t0)))


if __name__ == "__main__":
    main()
