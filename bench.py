#!/usr/bin/env python
"""bench.py -- MatchFullSubmap constraint searches / s on B200 (BASELINE.json metric).

Workload = config C3 of BASELINE.json / SURVEY.md 8(d): ONE global loop-closure job of
`--submaps` x `--nodes` (default 100 x 100 = 10 000) independent (node, submap)
MatchFullSubmap searches on 1000x1000-cell (50 m) synthetic submaps, ~360-beam scans,
branch_and_bound_depth 7, min_score 0.6, 30 % of a submap's scans drawn inside it.

One step = the whole job.  With N ranks (one process per GPU) the job is PARTITIONED:
`sharding.partition_searches` assigns the searches to ranks by submap affinity (greedy LPT
on an estimated cost, a hot submap's node list split and the submap replicated), every rank
builds the stacks of its submaps and runs its shard through the host API, rank 0 gathers the
results in submission order.  No data-path collective; strong scaling (the job is fixed).

  value : searches/s of the whole job, inputs (cells, points, tables) resident in HBM
          when the timed region starts; CUDA events on each context's stream, MAX over
          ranks; L2 flushed (untimed) before every step.
  e2e   : the same from HOST buffers through the public host API every step: grid H2D,
          stack builds, job upload, searches, results D2H and the gather of every rank's
          results on rank 0.
  result_sha1 : hash of the gathered result array in submission order -- identical at
          every N (1-GPU and N-GPU runs return the same array).
  --impl reference : the reference algorithm (oracle restatement, thread-pooled like
          ConstraintBuilder2D) on the host cores, bounded sample per step.
"""
import argparse
import ctypes as C
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
sys.path.insert(0, ROOT)

MIN_SCORE = 0.6
DEPTH = 7
LINEAR = 7.0
ANGULAR = float(np.deg2rad(30.0))


def parse(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=3)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--grid", type=int, default=1000, help="submap cells per side")
    p.add_argument("--submaps", type=int, default=100, help="submaps of the job")
    p.add_argument("--nodes", type=int, default=100, help="searches per submap")
    p.add_argument("--cpu-searches", type=int, default=0, help="CPU sample size (0 = auto)")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-extras", action="store_true",
                   help="skip the C1 / C2 / C4 / C5 sub-records (they run at N = 1 only)")
    return p.parse_args(argv)


# ---------------------------------------------------------------------------------------
# The C3 job.  Deterministic (seed = 1000 * config + index).  Every submap s has
# 3 * ceil(nodes / 10) scans ray-cast inside it; node k of submap s is one of them when
# k % 10 < 3 (30 % true positives) and otherwise a scan drawn inside ANOTHER submap (a
# negative that must return "no match").
# ---------------------------------------------------------------------------------------
def _gen_submap(task):
    import synth
    s, grid, n_pos = task
    w = synth.World(grid, grid, seed=3000 + s)
    scans = []
    for k in range(n_pos):
        _, pts = synth.make_scan(w, 1000 * s + k)
        if len(pts) < 8:
            _, pts = synth.make_scan(w, 1000 * s + k + 500)
        scans.append(pts)
    return s, w.cells, (w.res, w.max_x, w.max_y), scans


class Job:
    """cells[s], limits[s]; per search (submission order, submap-major): submap[i], cloud[i]."""

    def __init__(self, args, limit_submaps=None):
        n_sub, n_nodes, grid = args.submaps, args.nodes, args.grid
        n_pos = 3 * ((n_nodes + 9) // 10)
        tasks = [(s, grid, n_pos) for s in range(n_sub)]
        world = int(os.environ.get("WORLD_SIZE", "1"))
        procs = max(1, min(len(tasks), (os.cpu_count() or 1) // max(1, world)))
        if procs > 1:
            import multiprocessing as mp
            with mp.get_context("fork").Pool(procs) as pool:  # before any CUDA initialisation
                made = pool.map(_gen_submap, tasks, chunksize=1)
        else:
            made = [_gen_submap(t) for t in tasks]
        self.cells = [m[1] for m in made]
        self.limits = [m[2] for m in made]
        pos = [m[3] for m in made]
        self.submap, self.clouds, self.inside = [], [], []
        for s in range(n_sub):
            for k in range(n_nodes):
                inside = (k % 10) < 3
                if inside or n_sub == 1:
                    src, which = s, ((k // 10) * 3 + k % 10) % n_pos
                else:
                    src, which = (s + 1 + k % (n_sub - 1)) % n_sub, (7 * k + s) % n_pos
                self.submap.append(s)
                self.clouds.append(pos[src][which])
                self.inside.append(bool(inside))
        self.submap = np.asarray(self.submap, dtype=np.int64)
        self.num_submaps, self.grid, self.res = n_sub, grid, self.limits[0][0]
        self.points = np.array([len(c) for c in self.clouds], dtype=np.int64)
        self.scans = np.array([num_scans_of(c, self.res) for c in self.clouds], dtype=np.int64)

    def __len__(self):
        return len(self.clouds)


def num_scans_of(cloud, res):
    """SearchParameters::num_scans for a full-submap search (CSM2D.cc:27-55)."""
    r = max(3.0 * res, float(np.sqrt((cloud[:, 0].astype(np.float64) ** 2 +
                                      cloud[:, 1].astype(np.float64) ** 2).max())))
    step = (1.0 - 1e-3) * np.arccos(1.0 - res * res / (2.0 * r * r))
    return 2 * int(np.ceil(np.pi / step)) + 1


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


def cpu_reference_kind():
    """"reference": oracle/_ref/libref.so, the reference's own scan-matcher sources compiled unmodified
    (oracle/Makefile `ref`; prebuilt, it travels to the GPU box).  "port": the oracle restatement."""
    if os.environ.get("CSM_CPU_BASELINE") == "port":
        return "port"
    return "reference" if os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libref.so")) else "port"


def cpu_reference_run(job, idx, threads):
    """Times the constraint searches `idx` of the job on the host threads: the reference's own
    FastCorrelativeScanMatcher2D where oracle/_ref was built, else the oracle restatement; stack
    construction included, as one pool task per submap (ConstraintBuilder2D's shape)."""
    if cpu_reference_kind() == "reference":
        import ref_binding as ob
        make_grid = ob.RefGrid
    else:
        import oracle_binding as ob
        make_grid = ob.OracleGrid
    used = sorted(set(int(job.submap[i]) for i in idx))
    remap = {s: k for k, s in enumerate(used)}
    grids = [make_grid(job.cells[s], *job.limits[s]) for s in used]
    n = len(idx)
    secs, out, stats = ob.constraint_batch(
        threads, grids, np.zeros((len(grids), 3)), [remap[int(job.submap[i])] for i in idx], [1] * n,
        np.zeros((n, 3)), [job.clouds[i] for i in idx], sampling_ratio=1.0,
        global_min_score=MIN_SCORE, min_score=MIN_SCORE, linear_window=LINEAR,
        angular_window=ANGULAR, depth=DEPTH)
    return secs, out, stats


def cpu_baseline_text(kind):
    return ("the reference's own FastCorrelativeScanMatcher2D (oracle/_ref: its sources compiled unmodified "
            "against stand-in headers for Eigen/glog/abseil/protobuf), driven like ConstraintBuilder2D"
            if kind == "reference" else
            "thread-pooled oracle restatement of ConstraintBuilder2D (oracle/_ref not built on this box)")


def cpu_sample(job, n):
    """`n` searches spread over the job: whole groups of 8 nodes from evenly spaced submaps, so
    that the sample keeps the job's 30 % positives and its stack-build share."""
    per = 8
    subs = max(1, (n + per - 1) // per)
    chosen = np.unique(np.linspace(0, job.num_submaps - 1, min(subs, job.num_submaps)).astype(int))
    idx = []
    for s in chosen:
        idx.extend(np.nonzero(job.submap == s)[0][:max(per, (n + len(chosen) - 1) // len(chosen))].tolist())
    return idx[:max(n, 1)]


def workload_config(args, world, job, n_pts):
    return {"workload": "C3 2D global loop closure: one job of %d (node, submap) MatchFullSubmap searches "
                        "on %d submaps of %dx%d cells (50 m), depth %d, min_score %.2f, 30%% true positives"
                        % (len(job), args.submaps, args.grid, args.grid, DEPTH, MIN_SCORE),
            "searches_per_step": len(job), "submaps": args.submaps,
            "mean_points_per_scan": n_pts,
            "l2": "flushed before every step (256 MiB write, untimed); the %d stacks of a step (36 MB "
                  "each with their quad words) exceed L2" % args.submaps,
            "parallelism": "one job partitioned over %d ranks by submap affinity (LPT on estimated cost, "
                           "hot submaps replicated), results gathered on rank 0 in submission order; no "
                           "data-path collective" % world}


def run_reference(args):
    """The reference's CPU implementation of the path on the host cores: its own matcher sources
    (oracle/_ref, kind "reference") where they were built, else the oracle restatement (kind
    "port"); driven like ConstraintBuilder2D (one task per stack build, one per search) on all
    host threads.  Each step is a bounded sample of the job."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    # the sample touches few submaps: generate only those
    n = args.cpu_searches or max(8, 4 * threads)
    job = Job(args)
    idx = cpu_sample(job, n)
    times, total_scored = [], 0.0
    steps, warmup = args.steps, args.warmup
    it = 0
    while it < warmup + steps:
        secs, out, stats = cpu_reference_run(job, idx, threads)
        if it >= warmup:
            times.append(secs)
            total_scored += stats[0] + stats[1]
        if it == 0 and secs * (warmup + steps) > 200:
            warmup = min(warmup, 1)  # keep the whole arm within a few minutes
            steps = max(1, min(steps, int(200 / secs) - warmup))
        it += 1
    total = sum(times)
    value = len(idx) * len(times) / total
    n_pts = float(job.points.mean())
    line = {
        "impl": "reference", "metric": "MatchFullSubmap constraint searches/sec",
        "value": value, "unit": "searches/s", "n_gpus": args.gpus, "steps": len(times),
        "warmup": warmup, "ms_per_step": 1000.0 * total / len(times),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u8",
        "data": "synthetic",
        "config": workload_config(args, max(1, args.gpus), job, n_pts),
        "candidate_scores_per_s": total_scored / total if total_scored else None,
        "cpu_baseline": {"value": value, "unit": "searches/s", "cores": threads, "kind": cpu_reference_kind(),
                         "sample": "%d searches of the job per step over %d submaps (stack builds included), "
                                   "%s, on %d host threads" %
                                   (len(idx), len(set(int(job.submap[i]) for i in idx)),
                                    cpu_baseline_text(cpu_reference_kind()), threads)},
        "e2e": {"value": value, "unit": "searches/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
    }
    emit(line)


def _load_sharding():
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        "csm_sharding", os.path.join(ROOT, "cartographer-1_b200", "sharding.py"))
    sh = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(sh)
    return sh


def time_grid_broadcast(torch, dist, world, local, sizes=(2000000, 8000000), reps=10):
    """A new submap's uint16 grid reaching every replica: NCCL broadcast over NVLink from the
    owner's HBM against the alternative of each replica copying it from pinned host memory
    ("measure both").  Device-timed, max over ranks, microseconds."""
    out = {}
    for nbytes in sizes:
        dev = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
        host = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
        res = {}
        for name in (("nccl_broadcast", "h2d_copy") if world > 1 else ("h2d_copy",)):
            for it in range(3 + reps):
                if it == 3:
                    if world > 1:
                        dist.barrier()
                    torch.cuda.synchronize()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                if name == "nccl_broadcast":
                    dist.broadcast(dev, src=0)
                else:
                    dev.copy_(host, non_blocking=True)
            e1.record()
            e1.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) / reps], dtype=torch.float64, device="cuda")
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            res[name + "_us"] = 1e3 * float(t[0])
        res["gbs"] = {k[:-3]: nbytes / (v * 1e-6) / 1e9 for k, v in res.items()}
        out["%d_bytes" % nbytes] = res
    return out


def run_b200(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    t_gen = time.time()
    job = Job(args)  # forks worker processes: must precede CUDA initialisation
    t_gen = time.time() - t_gen
    import torch
    import torch.distributed as dist
    from __graft_entry__ import _load_package
    pkg = _load_package()
    sh = _load_sharding()
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the scan matcher has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- partition of the job (identical on every rank: deterministic inputs) ----------
    n_total = len(job)
    cost = sh.search_costs(job.points, job.scans)
    owner = sh.partition_searches(job.submap, cost, job.num_submaps, world)
    shard_sizes = np.bincount(owner, minlength=world)
    idx = np.nonzero(owner == rank)[0]
    my_submaps = sorted(set(int(s) for s in job.submap[idx]))
    slot = {s: k for k, s in enumerate(my_submaps)}
    replicated = sum(1 for r in sh.replicas(job.submap, owner, job.num_submaps) if len(r) > 1)
    clouds = [job.clouds[i] for i in idx]
    target = [slot[int(job.submap[i])] for i in idx]
    n_search = len(idx)

    ctx = pkg.Context(local)
    raw = C.c_void_p(ctx.raw)
    # the submaps' grids: host buffers, page-locked (the step copies them from pinned host memory)
    grids = [pkg.Grid(job.cells[s], *job.limits[s]).pin() for s in my_submaps]
    matchers = [pkg.Matcher(ctx, g, LINEAR, ANGULAR, DEPTH) for g in grids]
    cells_dev = [torch.from_numpy(job.cells[s].view(np.int16)).cuda() for s in my_submaps]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    batch = pkg.PreparedBatch(ctx, [matchers[t] for t in target], clouds, [MIN_SCORE] * n_search)
    abi = pkg.abi()
    stream = torch.cuda.ExternalStream(abi.csm_ctx_stream(raw))

    def timed(fn, end_on_current_stream=False):
        """One step: L2 flush (untimed), then `fn` bracketed by CUDA events; the first on the
        context's stream, the second there too or (e2e: the gather runs on torch's stream
        after the results reached the host) on the current stream.  (result, ms)."""
        flush.fill_(1)  # 256 MiB write > 126 MB L2
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        out = fn()
        if end_on_current_stream:
            e1.record()
        else:
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

    # the searches' point clouds as the host buffers the C entry point takes (xyz + offsets)
    host_clouds = pkg.PackedClouds(clouds)
    e2e_host = {"matchers_h2d_and_stack_builds": 0.0, "match_batch": 0.0, "pack_and_gather": 0.0,
                "matcher_teardown": 0.0}

    def e2e_step():
        # host buffers in, gathered host results out: grid H2D + stack builds + upload +
        # searches + D2H + gather of every rank's results on rank 0 (one collective)
        t0 = time.perf_counter()
        ms = pkg.Matcher.create_many(ctx, grids, LINEAR, ANGULAR, DEPTH)  # FastCorrelativeScanMatcher2D::CreateMany
        t1 = time.perf_counter()
        res = pkg.match_batch(ctx, [ms[t] for t in target], host_clouds, [MIN_SCORE] * n_search)
        t2 = time.perf_counter()
        packed = sh.pack_results(idx, res)
        if world > 1:
            packed = sh.gather_packed(packed, shard_sizes, n_total, device=torch.device("cuda", local))
        t3 = time.perf_counter()
        del ms
        e2e_host["matchers_h2d_and_stack_builds"] += t1 - t0  # one pipelined call, synchronised at its end
        e2e_host["match_batch"] += t2 - t1
        e2e_host["pack_and_gather"] += t3 - t2
        e2e_host["matcher_teardown"] += time.perf_counter() - t3
        return packed

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
    import roofline2d
    timed_profile = roofline2d.depth_profile(abi, raw)
    abi.csm_ctx_set_profiling(raw, 0)
    build = sum(a.elapsed_time(b) for a, b in build_ms) / max(1, len(build_ms))
    scored_per_step = float(sum(r.candidates_scored for r in res))
    lookups_per_step = float(sum(r.candidates_scored * len(c) for r, c in zip(res, clouds)))
    found = sum(1 for r in res if r.found)
    ties = sum(1 for r in res if r.tie_resolution)

    # ---- end-to-end arm ----------------------------------------------------------------
    for _ in range(min(args.warmup, 2)):
        timed(e2e_step, True)
    barrier()
    ms_e2e = 0.0
    for k in e2e_host:
        e2e_host[k] = 0.0
    for _ in range(args.steps):
        gathered, ms = timed(e2e_step, True)
        ms_e2e += ms
    barrier()
    mine = sh.pack_results(idx, res)
    assert np.array_equal(gathered[idx] if world > 1 else gathered, mine), \
        "the end-to-end results differ from the resident ones"
    result_sha1 = hashlib.sha1(np.ascontiguousarray(gathered[:, 1:]).tobytes()).hexdigest()

    # ---- whole-job totals over ranks -----------------------------------------------------
    tot = torch.tensor([scored_per_step, lookups_per_step, float(found), float(ties), float(launches),
                        float(replicated if rank == 0 else 0)], dtype=torch.float64, device="cuda")
    mx = torch.tensor([ms_resident, ms_e2e, build] + [phase_ms[k] for k in range(5)],
                      dtype=torch.float64, device="cuda")
    per_rank = torch.zeros(world, dtype=torch.float64, device="cuda")
    per_rank[rank] = ms_resident / args.steps
    if world > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        dist.all_reduce(per_rank, op=dist.ReduceOp.SUM)
        mx_all = mx.clone()
        dist.all_reduce(mx_all, op=dist.ReduceOp.MAX)
        ms_resident, ms_e2e = float(mx_all[0]), float(mx_all[1])
    scored_job, lookups_job, found_job, ties_job, launches_job = (float(tot[k]) for k in range(5))
    value = n_total * args.steps / (ms_resident * 1e-3)
    e2e_value = n_total * args.steps / (ms_e2e * 1e-3)
    broadcast = time_grid_broadcast(torch, dist, world, local)

    n_pts = float(job.points.mean())
    res0 = job.res
    h2d = sum(job.cells[s].nbytes + 65536 for s in range(job.num_submaps)) + sum(
        c.nbytes + 8 * int(k) + 160 for c, k in zip(job.clouds, job.scans))
    d2h = n_total * 40
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "measured"
    else:
        peak, peak_src = 6650.0, "fallback"

    line = {
        "metric": "MatchFullSubmap constraint searches/sec", "value": value, "unit": "searches/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_resident / args.steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(args, world, job, n_pts),
        "candidate_scores_per_s": scored_job * args.steps / (ms_resident * 1e-3),
        "candidates_scored_per_search": scored_job / n_total,
        "found": int(found_job), "host_tie_replays": int(ties_job),
        "result_sha1": result_sha1,
        "e2e": {"value": e2e_value, "unit": "searches/s", "h2d_bytes_per_step": int(h2d),
                "d2h_bytes_per_step": int(d2h), "ms_per_step": ms_e2e / args.steps,
                "host_wall_ms_per_step": {k: 1e3 * v / args.steps for k, v in e2e_host.items()},
                "includes": "grid H2D, stack builds, job upload, searches, results D2H, gather on rank 0"},
        "gpu_launches": int(launches_job),
        "clocks": clocks,
        "partition": {"searches_per_rank": [int(x) for x in shard_sizes],
                      "replicated_submaps": int(tot[5]),
                      "resident_ms_per_step_per_rank": [float(x) for x in per_rank.cpu()],
                      "idle_fraction_of_slowest": [float(1.0 - x / max(1e-9, float(per_rank.max())))
                                                   for x in per_rank.cpu()],
                      "estimated_cost_imbalance": float(np.bincount(owner, weights=cost, minlength=world).max() /
                                                        (cost.sum() / world))},
        "grid_broadcast_us": broadcast,
        "workload_generation_s": t_gen,
    }
    names = ["k_discretize", "k_score_top", "k_dive", "k_expand", "phase2"]
    line["phase_ms_per_step"] = {n: phase_ms[i] / args.steps for i, n in enumerate(names)}
    line["phase_candidates_per_step"] = {n: phase_scored[i] / args.steps for i, n in enumerate(names)}
    if rank == 0:  # rank 0's shard: per-depth times of its own launches against the measured ceiling
        xy, _ = pkg.discretize(ctx, res0, job.limits[my_submaps[0]][1], job.limits[my_submaps[0]][2],
                               args.grid, args.grid, clouds[0], (0.0, 0.0, 0.3), 1.0, 0.0, shrink=False)
        line.update(roofline2d.measure(
            pkg, abi, raw, timed_profile, max(1, runs.value) // max(1, args.steps), list(phase_ms),
            list(phase_scored), args.steps, float(np.mean([len(c) for c in clouds])), args.grid, DEPTH,
            peak, peak_src, resident_step, np.ascontiguousarray(xy[0], dtype=np.int32)))
    line["stack_build"] = {"ms_per_step": build, "submaps": len(matchers),
                           "algorithmic_gbs": (len(matchers) * (2 * args.grid ** 2 + sum(
                               (args.grid + 2 ** i - 1) ** 2 for i in range(DEPTH)))) / (build * 1e6)
                           if build > 0 else None}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        n = args.cpu_searches or max(8, min(n_total, 4 * threads))
        sample = cpu_sample(job, n)
        secs, cpu_out, stats = cpu_reference_run(job, sample, threads)
        kind = cpu_reference_kind()
        for k, i in enumerate(sample):  # the CPU leg doubles as a full-size parity check
            a, b = cpu_out[k], res[i]
            assert a.found == b.found and (not a.found or a.score == b.score), "parity broke at %d" % i
            if a.found and kind == "port":  # the reference's Match() does not return the winner's indices
                assert (a.scan_index, a.x_index_offset, a.y_index_offset) == (
                    b.scan_index, b.x_index_offset, b.y_index_offset), "parity broke at %d" % i
        line["cpu_baseline"] = {
            "value": len(sample) / secs, "unit": "searches/s", "cores": threads, "kind": kind,
            "sample": "%d searches of the job over %d submaps (stack builds included), %s, on %d host "
                      "threads, %.1f s; found flags and scores identical to the GPU's" %
                      (len(sample), len(set(int(job.submap[i]) for i in sample)), cpu_baseline_text(kind),
                       threads, secs),
            "candidate_scores_per_s": (stats[0] + stats[1]) / secs if kind == "port" else None}
    if rank == 0 and world == 1 and not args.no_extras:
        import extra_configs
        line["extra"] = {"configs": extra_configs.measure_all(pkg, ctx, torch, peak)}
    if rank == 0:
        emit(line)
    if world > 1:
        dist.barrier()
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
