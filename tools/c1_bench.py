#!/usr/bin/env python
"""C1 (BASELINE.json configs[0]): 2D MatchFullSubmap on 400x400 submaps
(20 m at 5 cm), ~360-beam scans, branch_and_bound_depth 7; 256 (grid, scan) pairs, 30 % true
positives; min_score 0.6 and the stress variant 0.1.  This is the reference's own
CPU-runnable case: the thread-pooled oracle restatement is timed at T = 1, 4 and all
cores on a bounded sample, next to the GPU's single-search latency (one search per call
through the public host API, host buffers in, result out) and its batched throughput.
Prints one JSON line per min_score; results of the CPU sample must equal the GPU's."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from __graft_entry__ import _load_package  # noqa: E402
import oracle_binding as ob  # noqa: E402  (CPU baseline leg: the checker / timed port)
import synth  # noqa: E402

GRID, DEPTH, LINEAR, ANGULAR = 400, 7, 7.0, float(np.deg2rad(30.0))
N_SUB, N_NODES = 16, 16


def workload():
    worlds = [synth.World(GRID, GRID, seed=1000 + s) for s in range(N_SUB)]
    clouds, target = [], []
    for s in range(N_SUB):
        for k in range(N_NODES):
            src = worlds[s] if (k % 10) < 3 else worlds[(s + 1 + k) % N_SUB]
            _, pts = synth.make_scan(src, 1000 * s + k)
            if len(pts) < 8:
                _, pts = synth.make_scan(src, 1000 * s + k + 500)
            clouds.append(pts)
            target.append(s)
    return worlds, clouds, target


def cpu_run(worlds, clouds, target, idx, threads, min_score):
    used = sorted(set(target[i] for i in idx))
    remap = {s: k for k, s in enumerate(used)}
    grids = [ob.OracleGrid(worlds[s].cells, worlds[s].res, worlds[s].max_x, worlds[s].max_y) for s in used]
    n = len(idx)
    secs, out, _ = ob.constraint_batch(
        threads, grids, np.zeros((len(grids), 3)), [remap[target[i]] for i in idx], [1] * n,
        np.zeros((n, 3)), [clouds[i] for i in idx], sampling_ratio=1.0, global_min_score=min_score,
        min_score=min_score, linear_window=LINEAR, angular_window=ANGULAR, depth=DEPTH)
    return secs, out


def main():
    import torch
    pkg = _load_package()
    ctx = pkg.Context(0)
    worlds, clouds, target = workload()
    grids = [pkg.Grid(w.cells, w.res, w.max_x, w.max_y) for w in worlds]
    matchers = [pkg.Matcher(ctx, g, LINEAR, ANGULAR, DEPTH) for g in grids]
    cores = os.cpu_count() or 1
    for min_score in (0.6, 0.1):
        ms = [min_score] * len(clouds)
        for _ in range(3):
            res = pkg.match_batch(ctx, [matchers[t] for t in target], clouds, ms)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        reps = 5
        for _ in range(reps):
            res = pkg.match_batch(ctx, [matchers[t] for t in target], clouds, ms)
        batched_s = (time.perf_counter() - t0) / reps
        lat = []
        for i in range(0, len(clouds), 4):  # 64 single-search calls
            t0 = time.perf_counter()
            one = pkg.match_batch(ctx, [matchers[target[i]]], [clouds[i]], [min_score])[0]
            lat.append(time.perf_counter() - t0)
            assert (one.found, one.scan_index, one.x_index_offset, one.y_index_offset) == \
                (res[i].found, res[i].scan_index, res[i].x_index_offset, res[i].y_index_offset)
        line = {
            "workload": "C1 MatchFullSubmap 400x400 depth 7, min_score %.1f, %d searches on %d submaps"
                        % (min_score, len(clouds), N_SUB),
            "found": int(sum(r.found for r in res)),
            "candidates_scored_per_search": float(np.mean([r.candidates_scored for r in res])),
            "gpu_batched_searches_per_s": len(clouds) / batched_s,
            "gpu_single_search_latency_ms": {"median": 1e3 * float(np.median(lat)),
                                             "p90": 1e3 * float(np.percentile(lat, 90))},
            "cpu_port": {},
        }
        for threads, n in ((1, 8), (4, 24), (cores, max(32, 2 * cores))):
            idx = list(range(0, len(clouds), max(1, len(clouds) // n)))[:n]
            secs, out = cpu_run(worlds, clouds, target, idx, threads, min_score)
            for k, i in enumerate(idx):  # the timed sample doubles as a parity check
                a, b = out[k], res[i]
                assert a.found == b.found and (not a.found or (
                    a.scan_index, a.x_index_offset, a.y_index_offset, a.score) == (
                    b.scan_index, b.x_index_offset, b.y_index_offset, b.score)), "parity broke at %d" % i
            line["cpu_port"]["threads_%d" % threads] = {"searches_per_s": len(idx) / secs,
                                                        "sample": len(idx), "seconds": secs}
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
