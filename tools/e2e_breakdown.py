#!/usr/bin/env python
"""Wall-clock breakdown of one end-to-end bench step (diagnostic; needs a GPU)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402


def main():
    args = bench.parse()
    job = bench.Job(args)
    from __graft_entry__ import _load_package
    pkg = _load_package()
    ctx = pkg.Context(0)
    grids = [pkg.Grid(job.cells[s], *job.limits[s]) for s in range(job.num_submaps)]
    n = len(job)
    target = [int(s) for s in job.submap]
    for it in range(4):
        t0 = time.perf_counter()
        ms = [pkg.Matcher(ctx, g, bench.LINEAR, bench.ANGULAR, bench.DEPTH) for g in grids]
        t1 = time.perf_counter()
        b = pkg.PreparedBatch(ctx, [ms[t] for t in target], job.clouds, [bench.MIN_SCORE] * n)
        t2 = time.perf_counter()
        b.run()
        t3 = time.perf_counter()
        b.fetch()
        t4 = time.perf_counter()
        b.close()
        del ms
        t5 = time.perf_counter()
        res = pkg.match_batch(ctx, [pkg.Matcher(ctx, g, bench.LINEAR, bench.ANGULAR, bench.DEPTH) for g in grids][0:0] or
                              [m for m in [pkg.Matcher(ctx, g, bench.LINEAR, bench.ANGULAR, bench.DEPTH) for g in grids]
                               for _ in range(args.nodes)], job.clouds, [bench.MIN_SCORE] * n)
        t6 = time.perf_counter()
        print("iter %d: matchers %.1f ms, prepare+upload %.1f ms, run(enqueue) %.1f ms, fetch(wait) %.1f ms, "
              "destroy %.1f ms, total %.1f ms | one-call match_batch incl. matchers %.1f ms" %
              (it, 1e3 * (t1 - t0), 1e3 * (t2 - t1), 1e3 * (t3 - t2), 1e3 * (t4 - t3), 1e3 * (t5 - t4),
               1e3 * (t5 - t0), 1e3 * (t6 - t5)), flush=True)


if __name__ == "__main__":
    main()
