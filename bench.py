#!/usr/bin/env python
"""bench.py -- MatchFullSubmap constraint searches / s on B200 (BASELINE.json metric).

Workload (config C3 of BASELINE.json / SURVEY.md 8(d)): global loop closure,
independent (node, submap) MatchFullSubmap searches on 1000x1000-cell (50 m) synthetic
submaps, ~360-point scans, branch_and_bound_depth 7, min_score 0.6, 30 % of the scans
drawn inside the target submap.  The 10k-search job is sharded over the ranks with no
data-path collective (weak scaling: every rank runs the same number of searches per
step against its own submaps).  One step = every rank builds the precomputation stacks
of its `--submaps` submaps and runs `--submaps x --nodes` searches against them.

  value : searches/s over all ranks with inputs (cells, points, tables) resident in
          HBM when the timed region starts; CUDA-event time on the context's stream,
          max over ranks; L2 flushed (untimed) between steps.
  e2e   : the same through the public host API from HOST buffers (grid upload, stack
          build, search, results back on the host) every step.
  --impl reference : the reference algorithm (oracle restatement, thread-pooled like
          ConstraintBuilder2D) on the host cores, bounded sample per step.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)

MIN_SCORE = 0.6
DEPTH = 7
LINEAR = 7.0
ANGULAR = float(np.deg2rad(30.0))


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=3)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--grid", type=int, default=1000, help="submap cells per side")
    p.add_argument("--submaps", type=int, default=4, help="submaps per rank per step")
    p.add_argument("--nodes", type=int, default=25, help="searches per submap per step")
    p.add_argument("--cpu-searches", type=int, default=0, help="CPU sample size (0 = auto)")
    p.add_argument("--no-cpu-baseline", action="store_true")
    return p.parse_args()


def make_workload(args, rank):
    """Deterministic synthetic submaps and scans (seed = 1000 * config + index).
    Weak scaling: every rank runs a shard of the same size AND the same content (the
    generator ignores the rank), so that value(N) / value(1) measures the machine and not
    the luck of which scans landed on which rank; ranks never exchange data."""
    rank = 0
    import synth
    worlds, clouds, target = [], [], []
    n_sub = args.submaps
    for s in range(n_sub):
        worlds.append(synth.World(args.grid, args.grid, seed=3000 + rank * 1000 + s))
    decoy = synth.World(args.grid, args.grid, seed=3000 + rank * 1000 + 999)
    for s in range(n_sub):
        for k in range(args.nodes):
            inside = (k % 10) < 3  # 30 % true positives
            src = worlds[s] if inside else (worlds[(s + 1 + k) % n_sub] if n_sub > 1 and k % 2 else decoy)
            _, pts = synth.make_scan(src, 100000 * rank + 1000 * s + k)
            if len(pts) < 8:
                _, pts = synth.make_scan(src, 100000 * rank + 1000 * s + k + 500)
            clouds.append(pts)
            target.append(s)
    return worlds, clouds, target


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.samples, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "200"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append([x.strip() for x in line.split(",")])

    def mark(self):
        """Index of the next sample: call at the start of the timed region."""
        return len(self.samples)

    def stop(self, first=0):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        # Samples of the timed region; a region shorter than nvidia-smi's period falls
        # back to the last samples of the identical warm-up steps just before it.
        chosen = self.samples[first:] or self.samples[-3:]
        for s in chosen:
            try:
                sm.append(float(s[0]))
                mx.append(float(s[1]))
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, s[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": float(max(mx)) if mx else None, "reasons": sorted(reasons)}


def cpu_reference_run(args, worlds, clouds, target, n_searches, threads):
    """Times the thread-pooled reference restatement on `n_searches` searches of the
    workload (stack construction included, as one pool task per submap)."""
    import oracle_binding as ob
    idx = list(range(min(n_searches, len(clouds))))
    used = sorted(set(target[i] for i in idx))
    remap = {s: k for k, s in enumerate(used)}
    grids = [ob.OracleGrid(worlds[s].cells, worlds[s].res, worlds[s].max_x, worlds[s].max_y)
             for s in used]
    n = len(idx)
    secs, out, stats = ob.constraint_batch(
        threads, grids, np.zeros((len(grids), 3)), [remap[target[i]] for i in idx], [1] * n,
        np.zeros((n, 3)), [clouds[i] for i in idx], sampling_ratio=1.0,
        global_min_score=MIN_SCORE, min_score=MIN_SCORE, linear_window=LINEAR,
        angular_window=ANGULAR, depth=DEPTH)
    return secs, out, stats, n




def workload_config(args, world, n_search, n_pts):
    return {"workload": "C3 2D global loop closure: MatchFullSubmap on %dx%d submaps, depth %d, "
                        "min_score %.2f, 30%% true positives" % (args.grid, args.grid, DEPTH, MIN_SCORE),
            "submaps_per_rank_per_step": args.submaps, "searches_per_rank_per_step": n_search,
            "mean_points_per_scan": n_pts,
            "l2": "flushed between steps (256 MiB write, untimed); stacks + work lists of a step "
                  "exceed L2",
            "parallelism": "independent searches sharded over %d ranks by submap (identical synthetic "
                           "shard per rank), no collective" % world}


def run_reference(args):
    """The reference algorithm on the host cores: thread-pooled oracle restatement driven
    like ConstraintBuilder2D (one pool task per stack build, one per search).  The
    reference's own sources cannot be built here (Eigen, glog, abseil, protobuf, Lua,
    Ceres absent), so kind = "port".  Each step is a bounded sample of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    worlds, clouds, target = make_workload(args, 0)
    n = args.cpu_searches or max(4, min(len(clouds), 2 * threads))
    times, total_scored = [], 0.0
    steps, warmup = args.steps, args.warmup
    it = 0
    while it < warmup + steps:
        secs, out, stats, n_done = cpu_reference_run(args, worlds, clouds, target, n, threads)
        if it >= warmup:
            times.append(secs)
            total_scored += stats[0] + stats[1]
        if it == 0 and secs * (warmup + steps) > 200:
            warmup = min(warmup, 1)  # keep the whole arm within a few minutes
            steps = max(1, min(steps, int(200 / secs) - warmup))
        it += 1
    total = sum(times)
    value = n_done * len(times) / total
    n_pts = float(np.mean([len(c) for c in clouds]))
    line = {
        "impl": "reference", "metric": "MatchFullSubmap constraint searches/sec",
        "value": value, "unit": "searches/s", "n_gpus": args.gpus, "steps": len(times),
        "warmup": warmup, "ms_per_step": 1000.0 * total / len(times),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
        "data": "synthetic",
        "config": workload_config(args, max(1, args.gpus), len(clouds), n_pts),
        "candidate_scores_per_s": total_scored / total,
        "cpu_baseline": {"value": value, "unit": "searches/s", "cores": threads, "kind": "port",
                         "sample": "%d searches per step over %d submaps (stack builds included), "
                                   "thread-pooled oracle restatement of ConstraintBuilder2D; the "
                                   "reference sources are unbuildable here" %
                                   (n_done, len(set(target[:n_done])))},
        "e2e": {"value": value, "unit": "searches/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
    }
    emit(line)


def num_scans_of(cloud, res):
    """SearchParameters::num_scans for a full-submap search (CSM2D.cc:27-55), for byte counts."""
    r = max(3.0 * res, float(np.sqrt((cloud[:, 0].astype(np.float64) ** 2 +
                                      cloud[:, 1].astype(np.float64) ** 2).max())))
    step = (1.0 - 1e-3) * np.arccos(1.0 - res * res / (2.0 * r * r))
    return 2 * int(np.ceil(np.pi / step)) + 1


def run_b200(args):
    import torch
    import torch.distributed as dist
    from __graft_entry__ import _load_package
    pkg = _load_package()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the scan matcher has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    worlds, clouds, target = make_workload(args, rank)
    n_search = len(clouds)
    ctx = pkg.Context(local)
    raw = C.c_void_p(ctx.raw)
    grids = [pkg.Grid(w.cells, w.res, w.max_x, w.max_y) for w in worlds]
    matchers = [pkg.Matcher(ctx, g, LINEAR, ANGULAR, DEPTH) for g in grids]
    cells_dev = [torch.from_numpy(w.cells.view(np.int16)).cuda() for w in worlds]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    batch = pkg.PreparedBatch(ctx, [matchers[t] for t in target], clouds, [MIN_SCORE] * n_search)
    abi = pkg.abi()
    stream = torch.cuda.ExternalStream(abi.csm_ctx_stream(raw))

    def timed(fn):
        """One step: L2 flush (untimed), then `fn` bracketed by CUDA events on the
        context's stream.  Returns (result, milliseconds)."""
        flush.fill_(1)  # 256 MiB write > 126 MB L2
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        out = fn()
        e1.record(stream)
        e1.synchronize()
        return out, e0.elapsed_time(e1)

    build_ms = []

    def resident_step():
        # inputs already in HBM: cells of every submap, points / tables of every search
        b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        b0.record(stream)
        for m, c in zip(matchers, cells_dev):
            rc = pkg.host().csmh_matcher_rebuild_dev(ctx.h, m.h, C.c_void_p(c.data_ptr()))
            assert rc == 0, abi.csm_last_error()
        b1.record(stream)
        batch.run()
        res = batch.fetch()
        build_ms.append((b0, b1))
        return res

    def e2e_step():
        # host buffers in, host results out: grid H2D + stack build + upload + search + D2H
        ms = [pkg.Matcher(ctx, g, LINEAR, ANGULAR, DEPTH) for g in grids]
        return pkg.match_batch(ctx, [ms[t] for t in target], clouds, [MIN_SCORE] * n_search)

    # ---- device-resident arm ---------------------------------------------------------
    sampler = ClockSampler(local)
    sampler.start()  # before the warm-up: nvidia-smi needs a moment to deliver samples
    for _ in range(args.warmup):
        timed(resident_step)
    build_ms.clear()
    abi.csm_ctx_set_profiling(raw, 1)  # per-phase events; resets the accumulators
    barrier()
    first_sample = sampler.mark()
    launches0 = abi.csm_ctx_launch_count(raw)
    ms_resident = 0.0
    for _ in range(args.steps):
        res, ms = timed(resident_step)
        ms_resident += ms
    launches = abi.csm_ctx_launch_count(raw) - launches0
    barrier()
    clocks = sampler.stop(first_sample)
    phase_ms = (C.c_double * 5)()
    phase_scored = (C.c_double * 5)()
    runs = C.c_int()
    abi.csm_ctx_get_profile(raw, phase_ms, phase_scored, C.byref(runs))
    abi.csm_ctx_set_profiling(raw, 0)
    build = sum(a.elapsed_time(b) for a, b in build_ms) / max(1, len(build_ms))
    scored_per_step = sum(r.candidates_scored for r in res)
    lookups_per_step = float(sum(r.candidates_scored * len(c) for r, c in zip(res, clouds)))
    found = sum(1 for r in res if r.found)
    ties = sum(1 for r in res if r.tie_resolution)

    # ---- end-to-end arm ----------------------------------------------------------------
    for _ in range(min(args.warmup, 3)):
        timed(e2e_step)
    barrier()
    ms_e2e = 0.0
    for _ in range(args.steps):
        out, ms = timed(e2e_step)
        ms_e2e += ms
    barrier()
    for a, b in zip(out, res):
        assert (a.found, a.scan_index, a.x_index_offset, a.y_index_offset) == \
               (b.found, b.scan_index, b.x_index_offset, b.y_index_offset)

    times = torch.tensor([ms_resident, ms_e2e], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    ms_resident, ms_e2e = float(times[0]), float(times[1])
    total_searches = world * n_search * args.steps
    value = total_searches / (ms_resident * 1e-3)
    e2e_value = total_searches / (ms_e2e * 1e-3)

    n_pts = float(np.mean([len(c) for c in clouds]))
    res0 = worlds[0].res
    h2d = sum(w.cells.nbytes + 65536 for w in worlds) + sum(
        c.nbytes + 8 * num_scans_of(c, res0) + 160 for c in clouds)
    d2h = n_search * 40
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "measured"
    else:
        peak, peak_src = 6650.0, "fallback"
    traffic_path = os.path.join(ROOT, "profiles", "traffic.json")
    traffic = json.load(open(traffic_path)) if os.path.exists(traffic_path) else {}

    # Dominant kernel: k_expand, the level-synchronous tile expansion (6 launches per
    # batch).  Algorithmic bytes = 1 B per (candidate, point) lookup.
    steps = max(1, runs.value)  # profiled batch runs (a step may split into several)
    names = ["k_discretize", "k_score_top", "k_dive", "k_expand", "phase2"]
    dom = int(np.argmax(list(phase_ms)))
    launches_of_dom = {"k_expand": DEPTH - 1}.get(names[dom], 1) * steps
    dom_bytes = phase_scored[dom] * n_pts
    dom_s = phase_ms[dom] / 1000.0
    achieved = dom_bytes / dom_s / 1e9 if dom_s > 0 else 0.0

    # Gather roofline G_L2: naive 1-byte gathers, warp per candidate,
    # lane per point, L2-resident table of the stack's size, this workload's scan shape.
    xy, _ = pkg.discretize(ctx, res0, worlds[0].max_x, worlds[0].max_y, args.grid, args.grid,
                           clouds[0], (0.0, 0.0, 0.3), 1.0, 0.0, shrink=False)
    pts_xy = np.ascontiguousarray(xy[0], dtype=np.int32)
    stack_bytes = sum((args.grid + 2 ** i - 1) ** 2 for i in range(DEPTH))
    g_l2 = C.c_double()
    rc = abi.csm_measure_gather_rate(raw, C.c_size_t(stack_bytes), args.grid + 63,
                                     pts_xy.ctypes.data_as(C.c_void_p), len(pts_xy), 1 << 21, 3,
                                     C.byref(g_l2))
    assert rc == 0, abi.csm_last_error()
    search_ms = sum(phase_ms) / steps
    line = {
        "metric": "MatchFullSubmap constraint searches/sec", "value": value, "unit": "searches/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_resident / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(args, world, n_search, n_pts),
        "candidate_scores_per_s": world * scored_per_step * args.steps / (ms_resident * 1e-3),
        "candidates_scored_per_search": scored_per_step / n_search,
        "found": found, "host_tie_replays": ties,
        "e2e": {"value": e2e_value, "unit": "searches/s", "h2d_bytes_per_step": int(h2d),
                "d2h_bytes_per_step": int(d2h), "ms_per_step": ms_e2e / args.steps},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "hbm", "kernel": {"k_expand": "k_expand_nodes"}.get(names[dom], names[dom]),
                     "achieved": achieved, "peak": peak,
                     "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic.get(names[dom]),
                     "peak_source": peak_src,
                     "launch_ms": phase_ms[dom] / launches_of_dom,
                     "algorithmic_bytes_per_launch": dom_bytes / launches_of_dom,
                     "note": "algorithmic bytes = 1 B per (candidate, point) lookup; the stack is "
                             "L2-resident (97-99 % L2 hits, traffic = DRAM bytes of the largest "
                             "launch), so the HBM peak is a loose bound: the unit that saturates "
                             "is the L1 data pipe, 85-88 % of its peak under ncu "
                             "(profiles/r01_v8_summary.md, DESIGN.md section 3)"},
        "gather_roofline": {"g_l2_lookups_per_s": g_l2.value,
                            "achieved_lookups_per_s": lookups_per_step / (search_ms * 1e-3),
                            "frac": lookups_per_step / (search_ms * 1e-3) / g_l2.value,
                            "searches_per_s_at_g_l2": g_l2.value / (lookups_per_step / n_search),
                            "how": "csm_measure_gather_rate: %d-byte table, %d-point scan, 2^21 "
                                   "candidates x 3" % (stack_bytes, len(pts_xy)),
                            "note": "G_L2 is the rate of the reference's naive pattern (one byte per "
                                    "candidate and point, SURVEY 8(d)); the quad-phase layout serves 4 "
                                    "lookups per 4-byte load from neighbouring words, so frac > 1 is a "
                                    "speed-up over that pattern, not a fraction of a ceiling"},
        "phase_ms_per_step": {n: phase_ms[i] / args.steps for i, n in enumerate(names)},
        "phase_candidates_per_step": {n: phase_scored[i] / args.steps for i, n in enumerate(names)},
        "stack_build": {"ms_per_step": build, "submaps": len(matchers),
                        "algorithmic_gbs": (len(matchers) * (2 * args.grid ** 2 + stack_bytes)) /
                        (build * 1e6) if build > 0 else None},
    }
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        n = args.cpu_searches or max(4, min(n_search, 4 * threads))
        secs, cpu_out, stats, n_done = cpu_reference_run(args, worlds, clouds, target, n, threads)
        for i in range(n_done):  # the oracle doubles as a full-size parity check
            a, b = cpu_out[i], res[i]
            assert a.found == b.found and (not a.found or (
                a.scan_index, a.x_index_offset, a.y_index_offset, a.score) == (
                b.scan_index, b.x_index_offset, b.y_index_offset, b.score)), "parity broke at %d" % i
        line["cpu_baseline"] = {
            "value": n_done / secs, "unit": "searches/s", "cores": threads, "kind": "port",
            "sample": "first %d searches of the step, thread-pooled oracle restatement "
                      "(ConstraintBuilder2D shape, stack builds included), %.1f s; results "
                      "identical to the GPU's" % (n_done, secs),
            "candidate_scores_per_s": (stats[0] + stats[1]) / secs}
    if rank == 0:
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def emit(line):
    """The one JSON line goes to the real stdout; everything else a library prints there
    (NCCL's version banner, torchrun notes) was redirected to stderr at start-up."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = 1

if __name__ == "__main__":
    a = parse()
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)  # fd 1 now points at stderr for this process and anything it loads
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
