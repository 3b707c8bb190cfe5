#!/usr/bin/env python
"""C5 of BASELINE.json: FastCorrelativeScanMatcher3D on B200 -- a measurement tool for
the 3D row of SURVEY.md 8 (bench.py stays on the headline 2D metric).

Workload: synthetic room (--size m square, 6 m high, boxes) voxelised at 0.1 m / 0.45 m,
node clouds of ~1500 / ~300 points within 20 m, 120-bucket histograms, options =
configuration_files/pose_graph.lua 3D defaults (depth 8, full_resolution_depth 3,
min_rotational_score 0.77, min_low_resolution_score 0.55, 5 m / 1 m / 15 deg).
One step = --matches local Match searches + --full MatchFullSubmap searches.

Prints one JSON line: searches/s device-resident (CUDA events on the context's
stream, L2 flushed between steps) and end to end from host buffers, per-phase times,
lookups/s against the measured L2 gather roofline, and the oracle restatement timed on
the host cores on a bounded sample (results compared with the GPU's).
"""
import argparse
import ctypes as C
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)
import synth3d  # noqa: E402

OPT = [8, 3, 0.77, 0.55, 5.0, 1.0, float(np.radians(15.0))]
MIN_SCORE = 0.55


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--size", type=float, default=40.0)
    p.add_argument("--matches", type=int, default=200)
    p.add_argument("--full", type=int, default=4)
    p.add_argument("--steps", type=int, default=5)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--points", type=int, default=1500)
    p.add_argument("--low-points", type=int, default=300)
    p.add_argument("--cpu-searches", type=int, default=0)
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--seed", type=int, default=9)
    return p.parse_args()


def make_workload(a):
    w = synth3d.World3D(a.seed, size_xy=a.size, height=6.0, n_boxes=int(a.size), samples_per_m2=300.0)
    hi_xyz, hi_val = w.voxels(w.res)
    lo_xyz, lo_val = w.voxels(w.low_res)
    rng = np.random.default_rng(a.seed + 1)
    nodes = []
    h = a.size / 2 - 2.0
    for k in range(a.matches + a.full):
        full = k >= a.matches
        pos = np.append(rng.uniform(-h, h, 2), rng.uniform(-0.3, 1.0))
        yaw = rng.uniform(-np.pi, np.pi)
        hi, lo, hist, grav, true = w.node(pos, yaw, max_range=20.0, n_high=a.points, n_low=a.low_points,
                                          gravity_yaw=rng.uniform(-0.3, 0.3))
        init = true.copy()
        if full:
            init[:3] = 0.0
        else:
            init[:3] += rng.uniform(-3.0, 3.0, 3) * [1, 1, 0.2]
            init[3:] = synth3d.quat_mul(synth3d.yaw_quat(rng.uniform(-0.15, 0.15)), true[3:])
        nodes.append((hi, lo, hist, grav, init, full, true))
    return w, (hi_xyz, hi_val, lo_xyz, lo_val), nodes


def main():
    a = parse()
    import torch
    from __graft_entry__ import _load_package
    pkg = _load_package()
    if not torch.cuda.is_available():
        raise RuntimeError("bench3d.py needs a CUDA device: the scan matcher has no CPU fallback")
    torch.cuda.set_device(0)
    w, (hi_xyz, hi_val, lo_xyz, lo_val), nodes = make_workload(a)
    ctx = pkg.Context(0)
    raw = C.c_void_p(ctx.raw)
    abi = pkg.abi()
    hist = w.submap_histogram()
    t0 = time.time()
    matcher = pkg.Matcher3D(ctx, 0.1, hi_xyz, hi_val, 0.45, lo_xyz, lo_val, hist, OPT)
    build_s = time.time() - t0
    n = len(nodes)
    node_objs = [pkg.Node3D(hi, lo, h_, g) for hi, lo, h_, g, _, _, _ in nodes]
    inits = np.array([x[4] for x in nodes])
    full = np.array([1 if x[5] else 0 for x in nodes], np.int32)
    mins = [MIN_SCORE] * n
    batch = pkg.PreparedBatch3D(ctx, [matcher] * n, node_objs, mins, inits, None, full)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    stream = torch.cuda.ExternalStream(abi.csm_ctx_stream(raw))

    def timed(fn):
        flush.fill_(1)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        out = fn()
        e1.record(stream)
        e1.synchronize()
        return out, e0.elapsed_time(e1)

    def resident_step():
        batch.run()
        return batch.fetch()

    def e2e_step():
        return pkg.match3d_batch(ctx, [matcher] * n, node_objs, mins, inits, None, full)

    for _ in range(a.warmup):
        timed(resident_step)
    abi.csm_ctx_set_profiling(raw, 1)
    launches0 = abi.csm_ctx_launch_count(raw)
    ms_res = 0.0
    for _ in range(a.steps):
        res, ms = timed(resident_step)
        ms_res += ms
    launches = abi.csm_ctx_launch_count(raw) - launches0
    phase_ms, phase_scored, runs = (C.c_double * 5)(), (C.c_double * 5)(), C.c_int()
    abi.csm_ctx_get_profile(raw, phase_ms, phase_scored, C.byref(runs))
    abi.csm_ctx_set_profiling(raw, 0)
    for _ in range(min(a.warmup, 3)):
        timed(e2e_step)
    ms_e2e = 0.0
    for _ in range(a.steps):
        out, ms = timed(e2e_step)
        ms_e2e += ms
    for x, y in zip(out, res):
        assert (x.found, x.scan_index, tuple(x.offset)) == (y.found, y.scan_index, tuple(y.offset))
    found = sum(1 for r in res if r.found)
    err = [float(np.linalg.norm(np.array(r.pose[:3]) - nd[6][:3])) for r, nd in zip(res, nodes) if r.found]
    scored = float(sum(r.candidates_scored for r in res))
    lookups = float(sum(r.candidates_scored * len(nd[0]) for r, nd in zip(res, nodes)))
    names = ["k3_discretize", "k3_score_top", "k3_dive", "k3_expand+k3_filter", "phase2"]
    search_ms = sum(phase_ms) / a.steps
    # L2 gather roofline on a table of the full-resolution level's size with a scan of
    # this workload's shape (2D projection of the first cloud's cells).
    stack_bytes = int(sum(np.prod(matcher.level(i)[1].shape) for i in range(OPT[0])))
    cells2d = np.ascontiguousarray(np.round(nodes[0][0][:, :2] / 0.1).astype(np.int32))
    cells2d -= cells2d.min(0)
    pitch = int(cells2d[:, 0].max() + 64)
    g_l2 = C.c_double()
    rc = abi.csm_measure_gather_rate(raw, C.c_size_t(stack_bytes), pitch, cells2d.ctypes.data_as(C.c_void_p),
                                     len(cells2d), 1 << 19, 3, C.byref(g_l2))
    assert rc == 0, abi.csm_last_error()
    line = {
        "metric": "FastCorrelativeScanMatcher3D searches/sec (C5)", "unit": "searches/s",
        "value": n * a.steps / (ms_res * 1e-3), "n_gpus": 1, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": ms_res / a.steps, "higher_is_better": True, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "C5: %gx%gx6 m room @0.1/0.45 m, %d Match + %d MatchFullSubmap per step, "
                               "%d/%d points, depth 8, full_resolution_depth 3, 5 m/1 m/15 deg; L2 flushed "
                               "between steps" % (a.size, a.size, a.matches, a.full, a.points, a.low_points),
                   "voxels": int(len(hi_val)), "low_voxels": int(len(lo_val)), "stack_bytes": stack_bytes},
        "found": found, "median_position_error_m": float(np.median(err)) if err else None,
        "host_tie_replays": sum(1 for r in res if r.tie_resolution),
        "mean_scans_per_search": float(np.mean([r.num_scans for r in res])),
        "candidate_scores_per_s": scored * a.steps / (ms_res * 1e-3),
        "candidates_scored_per_search": scored / n,
        "e2e": {"value": n * a.steps / (ms_e2e * 1e-3), "unit": "searches/s", "ms_per_step": ms_e2e / a.steps,
                "h2d_bytes_per_step": int(sum(nd[0].nbytes + nd[1].nbytes for nd in nodes)),
                "d2h_bytes_per_step": n * 48},
        "gpu_launches": int(launches),
        "phase_ms_per_step": {nm: phase_ms[i] / a.steps for i, nm in enumerate(names)},
        "gather_roofline": {"g_l2_lookups_per_s": g_l2.value,
                            "achieved_lookups_per_s": lookups / (search_ms * 1e-3),
                            "frac": lookups / (search_ms * 1e-3) / g_l2.value},
        "stack_build_seconds_incl_upload": build_s,
    }
    if not a.no_cpu_baseline:
        import oracle_binding as ob
        threads = os.cpu_count() or 1
        k = a.cpu_searches or min(n, 2 * threads)
        og, olg = ob.OracleGrid3D(0.1, hi_xyz, hi_val), ob.OracleGrid3D(0.45, lo_xyz, lo_val)
        t0 = time.time()
        om = ob.OracleFCSM3D(og, olg, hist, np.array(OPT))
        cpu_build = time.time() - t0

        def one(i):
            hi, lo, h_, g, init, fl, _ = nodes[i]
            return om.match(hi, lo, h_, MIN_SCORE, init.astype(np.float32), ob.IDENTITY_POSE7, g, bool(fl))

        t0 = time.time()
        with ThreadPoolExecutor(threads) as ex:
            cpu = list(ex.map(one, range(k)))
        secs = time.time() - t0
        for i, r in enumerate(cpu):  # full-size parity check
            g = res[i]
            assert bool(r.found) == bool(g.found) and (not r.found or (
                r.scan_index, tuple(r.offset), np.float32(r.score)) == (
                g.scan_index, tuple(g.offset), np.float32(g.score))), "parity broke at %d" % i
        line["cpu_baseline"] = {"value": k / secs, "unit": "searches/s", "cores": threads, "kind": "port",
                                "sample": "first %d searches of the step on a thread pool of the oracle "
                                          "restatement, %.1f s (+ %.1f s stack build); results identical to "
                                          "the GPU's" % (k, secs, cpu_build)}
    print(json.dumps(line))


if __name__ == "__main__":
    main()
