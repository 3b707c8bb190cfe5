"""The other configs of BASELINE.json as sub-records of bench.py's line (extra.configs.C1 / C2 /
C4 / C5), each a bounded run of a few seconds on rank 0 at N = 1: value on the GPU, what bounds
it, and the CPU restatement timed beside it on the box's host cores (its results compared
with the GPU's, so every record doubles as a parity check at that config's size).

  C1  2D MatchFullSubmap, 400 x 400 @ 5 cm, 360-beam scans, depth 7 (the reference's own
      CPU-runnable case): batched throughput and single-search latency
  C2  RealTimeCorrelativeScanMatcher2D, +-0.1 m / +-20 deg, ~250 points: scans/s at batch
      1 / 64 / 1024 through the host API
  C4  PrecomputationGridStack2D build, 2000 x 2000, depth 7: GB/s against the HBM peak
  C5  FastCorrelativeScanMatcher3D, 40 m room at 0.1 / 0.45 m, depth 8: searches/s
"""
import ctypes as C
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def _event_ms(torch, stream, fn, flush=None):
    if flush is not None:
        flush.fill_(1)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    out = fn()
    e1.record(stream)
    e1.synchronize()
    return out, e0.elapsed_time(e1)


def _cpu_kind():
    """"reference": oracle/_ref/libref.so (the reference's own sources) exists;
    "port": only the oracle restatement does."""
    return "reference" if os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libref.so")) and \
        os.environ.get("CSM_CPU_BASELINE") != "port" else "port"


_CPU_TEXT = {"reference": "the reference's own sources (oracle/_ref)", "port": "oracle restatement"}


def c1(pkg, ctx, torch):
    import synth
    kind = _cpu_kind()
    if kind == "reference":
        import ref_binding as ob
        make_grid = ob.RefGrid
    else:
        import oracle_binding as ob
        make_grid = ob.OracleGrid
    grid, depth, lin, ang = 400, 7, 7.0, float(np.deg2rad(30.0))
    n_sub, n_nodes = 8, 16
    worlds = [synth.World(grid, grid, seed=1000 + s) for s in range(n_sub)]
    clouds, target = [], []
    for s in range(n_sub):
        for k in range(n_nodes):
            src = worlds[s] if (k % 10) < 3 else worlds[(s + 1 + k) % n_sub]
            _, pts = synth.make_scan(src, 1000 * s + k)
            if len(pts) < 8:
                _, pts = synth.make_scan(src, 1000 * s + k + 500)
            clouds.append(pts)
            target.append(s)
    grids = [pkg.Grid(w.cells, w.res, w.max_x, w.max_y) for w in worlds]
    matchers = [pkg.Matcher(ctx, g, lin, ang, depth) for g in grids]
    ms = [0.6] * len(clouds)
    for _ in range(2):
        res = pkg.match_batch(ctx, [matchers[t] for t in target], clouds, ms)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    reps = 5
    for _ in range(reps):
        res = pkg.match_batch(ctx, [matchers[t] for t in target], clouds, ms)
    batched_s = (time.perf_counter() - t0) / reps
    lat = []
    for i in range(0, len(clouds), 4):
        t0 = time.perf_counter()
        one = pkg.match_batch(ctx, [matchers[target[i]]], [clouds[i]], [0.6])[0]
        lat.append(time.perf_counter() - t0)
        assert (one.found, one.scan_index, one.x_index_offset, one.y_index_offset) == \
            (res[i].found, res[i].scan_index, res[i].x_index_offset, res[i].y_index_offset)
    cores = os.cpu_count() or 1
    idx = list(range(0, len(clouds), max(1, len(clouds) // (2 * cores))))[:2 * cores]
    used = sorted(set(target[i] for i in idx))
    remap = {s: k for k, s in enumerate(used)}
    ogrids = [make_grid(worlds[s].cells, worlds[s].res, worlds[s].max_x, worlds[s].max_y) for s in used]
    secs, out, _ = ob.constraint_batch(
        cores, ogrids, np.zeros((len(ogrids), 3)), [remap[target[i]] for i in idx], [1] * len(idx),
        np.zeros((len(idx), 3)), [clouds[i] for i in idx], sampling_ratio=1.0, global_min_score=0.6,
        min_score=0.6, linear_window=lin, angular_window=ang, depth=depth)
    for k, i in enumerate(idx):
        a, b = out[k], res[i]
        assert a.found == b.found and (not a.found or a.score == b.score), "C1 parity broke at %d" % i
        if a.found and kind == "port":  # the reference's Match() does not return the winner's indices
            assert (a.scan_index, a.x_index_offset, a.y_index_offset) == (
                b.scan_index, b.x_index_offset, b.y_index_offset), "C1 parity broke at %d" % i
    return {
        "workload": "C1 MatchFullSubmap 400x400 @5 cm, depth 7, min_score 0.6, %d searches on %d submaps"
                    % (len(clouds), n_sub),
        "value": len(clouds) / batched_s, "unit": "searches/s",
        "how": "host API, host buffers in / results out (wall clock around %d batched calls)" % reps,
        "single_search_latency_ms": {"median": 1e3 * float(np.median(lat)), "p90": 1e3 * float(np.percentile(lat, 90))},
        "found": int(sum(r.found for r in res)),
        "candidates_scored_per_search": float(np.mean([r.candidates_scored for r in res])),
        "bound": "the sweep's gather ceiling (see roofline) plus per-call latency at this size",
        "cpu_baseline": {"value": len(idx) / secs, "unit": "searches/s", "cores": cores, "kind": kind,
                         "sample": "%d of the searches, thread-pooled, %s, %.1f s; found flags and scores "
                                   "identical to the GPU's" % (len(idx), _CPU_TEXT[kind], secs)}}


def c2(pkg, ctx, torch):
    import synth
    kind = _cpu_kind()
    if kind == "reference":
        import ref_binding as ob
        make_grid = ob.RefGrid
    else:
        import oracle_binding as ob
        make_grid = ob.OracleGrid
    w = synth.World(400, 400, seed=2000)
    g = pkg.Grid(w.cells, w.res, w.max_x, w.max_y)
    rng = np.random.default_rng(2)
    base = []
    for i in range(64):
        pose, pts = synth.make_scan(w, 2000 + i)
        base.append((pose, np.ascontiguousarray(pts[::max(1, len(pts) // 250)][:250])))
    clouds, inits = [], []
    for i in range(1024):
        pose, pts = base[i % 64]
        clouds.append(pts)
        inits.append((pose[0] + rng.uniform(-0.08, 0.08), pose[1] + rng.uniform(-0.08, 0.08),
                      pose[2] + rng.uniform(-0.26, 0.26)))
    out = {"workload": "C2 RealTimeCorrelativeScanMatcher2D: 400x400 live grid, ~250 points, +-0.1 m, +-20 deg, "
                       "weights 0.1", "unit": "scans/s",
           "how": "host API, host buffers in / poses out (wall clock)"}
    res = None
    for b, reps in ((1, 100), (64, 10), (1024, 5)):
        host_clouds = pkg.PackedClouds(clouds[:b])  # the host buffers of the call: xyz + offsets
        grids_b, inits_b = [g] * b, np.ascontiguousarray(inits[:b], dtype=np.float64)
        pkg.rt_match_batch(ctx, grids_b, host_clouds, inits_b)
        t0 = time.perf_counter()
        for _ in range(reps):
            res = pkg.rt_match_batch(ctx, grids_b, host_clouds, inits_b)
        dt = (time.perf_counter() - t0) / reps
        out["batch_%d" % b] = {"scans_per_s": b / dt, "ms_per_call": 1e3 * dt}
    out["value"] = max(out["batch_%d" % b]["scans_per_s"] for b in (1, 64, 1024))
    og = make_grid(w.cells, w.res, w.max_x, w.max_y)
    opts = {"linear_search_window": 0.1, "angular_search_window": float(np.deg2rad(20.0)),
            "translation_delta_cost_weight": 0.1, "rotation_delta_cost_weight": 0.1}
    n = 48

    def one(i):
        return ob.rt_match(og, opts, inits[i], clouds[i])

    cores = os.cpu_count() or 1
    t0 = time.perf_counter()
    with ThreadPoolExecutor(cores) as ex:
        cpu = list(ex.map(one, range(n)))
    secs = time.perf_counter() - t0
    for i, r in enumerate(cpu):
        if kind == "port":  # the reference's Match() does not return the winner's indices
            assert (r.scan_index, r.x_index_offset, r.y_index_offset) == \
                (res[i].scan_index, res[i].x_index_offset, res[i].y_index_offset), "C2 parity broke at %d" % i
        assert np.float32(r.score).tobytes() == np.float32(res[i].score).tobytes(), "C2 parity broke at %d" % i
        assert (r.pose_x, r.pose_y, r.pose_theta) == (res[i].pose_x, res[i].pose_y, res[i].pose_theta), \
            "C2 pose parity broke at %d" % i
    out["bound"] = "host-side preparation (penalty tables, packing) at large batches; k_rt_score itself is L1-latency bound"
    out["cpu_baseline"] = {"value": n / secs, "unit": "scans/s", "cores": cores, "kind": kind,
                           "sample": "%d scans, one Match call per pool thread, %s, %.2f s; score bits and poses "
                                     "identical to the GPU's" % (n, _CPU_TEXT[kind], secs)}
    return out


def c4(pkg, ctx, torch, hbm_peak):
    import oracle_binding as ob
    import synth
    abi = pkg.abi()
    stream = torch.cuda.ExternalStream(abi.csm_ctx_stream(C.c_void_p(ctx.raw)))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    depth = 7
    out = {"unit": "GB/s", "sizes": {}}
    # C4 is the reference's PrecomputationGridStack2D build: the copy of the cells that builds
    # otherwise leave for the refinement stage is switched off for this record.
    abi.csm_ctx_set_keep_cells(C.c_void_p(ctx.raw), 0)
    try:
        return _c4_sizes(pkg, ctx, torch, hbm_peak, abi, stream, flush, depth, out)
    finally:
        abi.csm_ctx_set_keep_cells(C.c_void_p(ctx.raw), 1)


def _c4_sizes(pkg, ctx, torch, hbm_peak, abi, stream, flush, depth, out):
    import oracle_binding as ob
    import synth
    for size in (400, 1000, 2000):
        w = synth.World(size, size, seed=4000 + size)
        g = pkg.Grid(w.cells, w.res, w.max_x, w.max_y)
        m = pkg.Matcher(ctx, g, 7.0, 0.5, depth)
        cells = torch.from_numpy(w.cells.view(np.int16)).cuda()
        ptr = C.c_void_p(cells.data_ptr())

        def build():
            return pkg.host().csmh_matcher_rebuild_dev(ctx.h, m.h, ptr)

        for _ in range(5):
            build()
        reps = 20
        _, warm = _event_ms(torch, stream, lambda: [build() for _ in range(reps)])
        cold = 0.0
        for _ in range(10):
            cold += _event_ms(torch, stream, build, flush)[1] / 10
        alg = 2 * size * size + sum((size + 2 ** i - 1) ** 2 for i in range(depth))
        out["sizes"]["%dx%d" % (size, size)] = {
            "ms_l2_flushed": cold, "ms_back_to_back": warm / reps, "algorithmic_bytes": alg,
            "algorithmic_gbs": alg / (cold * 1e6), "frac_of_hbm_peak": alg / (cold * 1e6) / hbm_peak}
        if size == 2000:
            out["workload"] = "C4 PrecomputationGridStack2D build 2000x2000, depth 7, cells resident in HBM"
            out["value"] = alg / (cold * 1e6)
            out["ms"] = cold
            out["roofline"] = {"bound": "hbm", "achieved": alg / (cold * 1e6), "peak": hbm_peak, "unit": "GB/s",
                               "frac": alg / (cold * 1e6) / hbm_peak,
                               "note": "algorithmic bytes = 2*W*H + sum of level bytes (SURVEY.md 8(d)); the build "
                                       "also writes the matcher's quad-phase words"}
            kind = _cpu_kind()
            if kind == "reference":
                import ref_binding as rb
                rg = rb.RefGrid(w.cells, w.res, w.max_x, w.max_y)
                secs = rb.stack_build_seconds(rg, depth)
                om = rb.RefFCSM2D(rg, 7.0, 0.5, depth)
            else:
                og = ob.OracleGrid(w.cells, w.res, w.max_x, w.max_y)
                t0 = time.perf_counter()
                om = ob.OracleFCSM2D(og, 7.0, 0.5, depth)
                secs = time.perf_counter() - t0
            assert np.array_equal(om.level(depth - 1), m.level(depth - 1)), "C4 parity broke"
            out["cpu_baseline"] = {"value": alg / secs / 1e9, "unit": "GB/s", "cores": 1, "kind": kind,
                                   "sample": "one 2000x2000 depth-7 PrecomputationGridStack2D, %s "
                                             "(SlidingWindowMaximum), one thread, %.2f s; coarsest level "
                                             "identical" % (_CPU_TEXT[kind], secs)}
    return out


def c5(pkg, ctx, torch):
    import fullsize_cases as fc
    import oracle_binding as ob
    import synth3d
    abi = pkg.abi()
    stream = torch.cuda.ExternalStream(abi.csm_ctx_stream(C.c_void_p(ctx.raw)))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    w = synth3d.World3D(9, size_xy=40.0, height=6.0, n_boxes=40, samples_per_m2=300.0)
    hi_xyz, hi_val = w.voxels(w.res)
    lo_xyz, lo_val = w.voxels(w.low_res)
    hist = w.submap_histogram()
    rng = np.random.default_rng(11)
    n_match, n_full = 100, 2
    nodes = []
    for k in range(n_match + n_full):
        full = k >= n_match
        pos = np.append(rng.uniform(-18, 18, 2), rng.uniform(-0.3, 1.0))
        hi, lo, h_, grav, true = w.node(pos, rng.uniform(-np.pi, np.pi), max_range=20.0, n_high=1500, n_low=300,
                                        gravity_yaw=rng.uniform(-0.3, 0.3))
        init = true.copy()
        if full:
            init[:3] = 0.0
        else:
            init[:3] += rng.uniform(-3.0, 3.0, 3) * [1, 1, 0.2]
            init[3:] = synth3d.quat_mul(synth3d.yaw_quat(rng.uniform(-0.15, 0.15)), true[3:])
        nodes.append((hi, lo, h_, grav, init.astype(np.float32), full))
    matcher = pkg.Matcher3D(ctx, 0.1, hi_xyz, hi_val, 0.45, lo_xyz, lo_val, hist, fc.C5_OPT)
    n = len(nodes)
    objs = [pkg.Node3D(hi, lo, h_, g) for hi, lo, h_, g, _, _ in nodes]
    inits = np.array([nd[4].astype(np.float64) for nd in nodes])
    full = np.array([1 if nd[5] else 0 for nd in nodes], np.int32)
    mins = [fc.C5_MIN_SCORE] * n
    batch = pkg.PreparedBatch3D(ctx, [matcher] * n, objs, mins, inits, None, full)

    def resident():
        batch.run()
        return batch.fetch()

    for _ in range(2):
        resident()
    steps, ms_res, ms_e2e = 3, 0.0, 0.0
    for _ in range(steps):
        res, ms = _event_ms(torch, stream, resident, flush)
        ms_res += ms
    for _ in range(steps):
        _, ms = _event_ms(torch, stream,
                          lambda: pkg.match3d_batch(ctx, [matcher] * n, objs, mins, inits, None, full), flush)
        ms_e2e += ms
    # CPU leg: the reference's own FastCorrelativeScanMatcher3D on its own HybridGrid where oracle/_ref
    # was built (kind "reference"), else the oracle restatement, whose HybridGrid is a hash map (kind "port").
    kind = "reference" if os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libref.so")) and \
        os.environ.get("CSM_CPU_BASELINE") != "port" else "port"
    if kind == "reference":
        import ref_binding as rb
        om = rb.RefFCSM3D(rb.RefGrid3D(0.1, hi_xyz, hi_val), rb.RefGrid3D(0.45, lo_xyz, lo_val), hist,
                          np.array(fc.C5_OPT))
    else:
        om = ob.OracleFCSM3D(ob.OracleGrid3D(0.1, hi_xyz, hi_val), ob.OracleGrid3D(0.45, lo_xyz, lo_val), hist,
                             np.array(fc.C5_OPT))
    cores = os.cpu_count() or 1
    k = min(n_match, 2 * cores if kind == "reference" else cores)

    def one(i):
        hi, lo, h_, g, init, _ = nodes[i]
        return om.match(hi, lo, h_, fc.C5_MIN_SCORE, init, ob.IDENTITY_POSE7, g, False)

    t0 = time.perf_counter()
    with ThreadPoolExecutor(cores) as ex:
        cpu = list(ex.map(one, range(k)))
    secs = time.perf_counter() - t0
    for i, r in enumerate(cpu):
        g = res[i]
        assert bool(r.found) == bool(g.found) and (not r.found or np.float32(r.score).tobytes() ==
                                                   np.float32(g.score).tobytes()), "C5 parity broke at %d" % i
        if r.found and kind == "port":  # the reference's Result does not carry the winner's indices
            assert (r.scan_index, tuple(r.offset)) == (g.scan_index, tuple(g.offset)), "C5 parity broke at %d" % i
        if r.found and kind == "reference":
            assert np.array(list(r.pose), np.float32).tobytes() == np.array(list(g.pose), np.float32).tobytes(), \
                "C5 pose parity broke at %d" % i
    scored = float(sum(r.candidates_scored for r in res))
    return {
        "workload": "C5 FastCorrelativeScanMatcher3D: 40x40x6 m room @0.1/0.45 m, %d Match + %d MatchFullSubmap "
                    "per step, 1500/300 points, depth 8, full_resolution_depth 3, 5 m / 1 m / 15 deg" % (n_match, n_full),
        "value": n * steps / (ms_res * 1e-3), "unit": "searches/s", "ms_per_step": ms_res / steps,
        "e2e": {"value": n * steps / (ms_e2e * 1e-3), "unit": "searches/s"},
        "found": int(sum(1 for r in res if r.found)),
        "candidate_scores_per_s": scored * steps / (ms_res * 1e-3),
        "bound": "L1/L2 sector throughput of the 8-child pack gather (k3_expand), profiles/r01_3d_summary.md",
        "cpu_baseline": {"value": k / secs, "unit": "searches/s", "cores": cores, "kind": kind,
                         "sample": "%d Match searches on %d host threads, %s, %.1f s; found flags, scores%s identical "
                                   "to the GPU's" % (k, cores,
                                                     "the reference's own FastCorrelativeScanMatcher3D on its own "
                                                     "HybridGrid (oracle/_ref)" if kind == "reference" else
                                                     "the oracle restatement (its HybridGrid is a hash map, slower per "
                                                     "lookup than the reference's brick tree)", secs,
                                                     " and poses" if kind == "reference" else " and winners")}}


def measure_all(pkg, ctx, torch, hbm_peak):
    out = {}
    for name, fn in (("C1", lambda: c1(pkg, ctx, torch)), ("C2", lambda: c2(pkg, ctx, torch)),
                     ("C4", lambda: c4(pkg, ctx, torch, hbm_peak)), ("C5", lambda: c5(pkg, ctx, torch))):
        t0 = time.perf_counter()
        try:
            out[name] = fn()
        except Exception as e:  # a failing sub-record must not take the headline line down
            out[name] = {"error": repr(e)}
        out[name]["seconds"] = time.perf_counter() - t0
    return out
