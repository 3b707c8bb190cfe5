"""Roofline of the dominant kernel of the 2D path, k_expand_nodes (the level-synchronous
branch-and-bound sweep), measured inside bench.py's own run.

The sweep is a gather from L2-resident quad planes: DRAM traffic is a fraction of a percent of
the algorithmic bytes, so the HBM peak is not its ceiling.  What binds is how fast the L1 / L2
path serves warp-wide 4-byte loads whose lanes are spread over S 32-byte sectors (S follows
from how sparse the alive nodes of a work item are along a plane row).  The ceiling is
therefore MEASURED: csm_measure_gather_ceiling runs a micro-kernel with the sweep's access
shape and nothing else, at several S; the run's own S per depth comes from k_item_stats
(csm_ctx_set_profiling(ctx, 2)), the run's time per depth from CUDA events around every
k_expand_nodes launch.

  per depth d:  requests_d = sum over work items of scan points (one warp load each)
                peak_d     = ceiling(S_d) requests/s,  ideal_ms_d = requests_d / peak_d
  frac = sum_d ideal_ms_d / sum_d measured_ms_d;  in GB/s: achieved = 1 B per (candidate,
  point) lookup / measured time, peak = achieved / frac.
"""
import ctypes as C
import json
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEPTH_FIELDS = 8
CEILING_SECTORS = (4, 6, 8, 10, 12, 14, 16, 20, 24, 32)


def depth_profile(abi, raw):
    out = (C.c_double * (16 * DEPTH_FIELDS))()
    rc = abi.csm_ctx_get_depth_profile(raw, out)
    assert rc == 0, abi.csm_last_error()
    return np.array(out[:]).reshape(16, DEPTH_FIELDS)


def measure_ceiling(abi, raw, table_bytes, num_sms_items=75776, points=256, repeats=3):
    table = {}
    for s in CEILING_SECTORS:
        v = C.c_double()
        rc = abi.csm_measure_gather_ceiling(raw, C.c_size_t(table_bytes), s, num_sms_items, points, repeats,
                                            C.byref(v))
        assert rc == 0, abi.csm_last_error()
        table[s] = v.value
    return table


def interpolate(table, s):
    xs = np.array(sorted(table))
    ys = np.array([table[x] for x in xs])
    return float(np.interp(min(max(s, xs[0]), xs[-1]), xs, ys))


def measure(pkg, abi, raw, timed_profile, timed_runs_steps, phase_ms, phase_scored, steps, n_pts,
            grid, depth, hbm_peak, hbm_peak_src, one_step, discretized_scan):
    """timed_profile: csm_ctx_get_depth_profile() of the timed steps (profiling 1);
    one_step(): runs one more (untimed) step, here with profiling 2 for the work-item statistics."""
    abi.csm_ctx_set_profiling(raw, 2)
    one_step()
    stats = depth_profile(abi, raw)
    abi.csm_ctx_set_profiling(raw, 0)
    quad_bytes = int(sum(4 * (grid + 4 * (1 << l)) ** 2 for l in range(depth)))
    ceiling = measure_ceiling(abi, raw, min(quad_bytes, 96 << 20))
    per_depth, ideal_ms, measured_ms, requests_total = [], 0.0, 0.0, 0.0
    for d in range(15, 0, -1):
        ms = timed_profile[d, 0] / steps
        items, requests, lane_loads, sectors, lines, nodes = stats[d, 2:8]
        if requests <= 0 or ms <= 0:
            continue
        s_d = sectors / requests
        peak_d = interpolate(ceiling, s_d)
        ideal = 1e3 * requests / peak_d
        per_depth.append({
            "depth": d, "expand_ms": ms, "list_bookkeeping_ms": timed_profile[d, 1] / steps,
            "work_items": items, "nodes": nodes, "lanes_alive_per_item": nodes / items,
            "requests": requests, "sectors_per_request": s_d, "lines_per_request": lines / requests,
            "requests_per_s": requests / (ms * 1e-3), "ceiling_requests_per_s": peak_d,
            "frac": ideal / ms})
        ideal_ms += ideal
        measured_ms += ms
        requests_total += requests
    frac = ideal_ms / measured_ms if measured_ms > 0 else None
    lookups = phase_scored[3] / steps * n_pts  # 1 B per (candidate, point) lookup, SURVEY.md 8(d)
    achieved = lookups / (measured_ms * 1e-3) / 1e9 if measured_ms > 0 else 0.0
    traffic_path = os.path.join(ROOT, "profiles", "r02_traffic.json")
    traffic = json.load(open(traffic_path)) if os.path.exists(traffic_path) else {}
    launches = len(per_depth) * max(1, timed_runs_steps)
    out = {}
    out["roofline"] = {
        "bound": "l1tex/L2 gather (not HBM): warp loads per second of a micro-kernel with the sweep's access "
                 "shape at the run's own sectors per request",
        "kernel": "k_expand_nodes", "achieved": achieved, "peak": achieved / frac if frac else None,
        "unit": "GB/s", "frac": frac,
        "traffic": traffic.get("k_expand_nodes", {}).get("dram_bytes_per_launch"),
        "traffic_source": traffic.get("source"),
        "peak_source": "csm_measure_gather_ceiling in this run (requests/s at 4..32 sectors per request, "
                       "interpolated at every depth's measured sectors per request)",
        "launch_ms": measured_ms / max(1, launches),
        "launches_per_step": launches,
        "algorithmic_bytes_per_launch": lookups / max(1, launches),
        "requests_per_step": requests_total,
        "lookups_per_request": lookups / requests_total if requests_total else None,
        "ceiling_requests_per_s_by_sectors": {str(k): v for k, v in ceiling.items()},
        "per_depth": per_depth,
        "hbm": {"peak": hbm_peak, "peak_source": hbm_peak_src, "frac_of_hbm_peak": achieved / hbm_peak,
                "note": "for reference only: the planes are L2-resident"},
    }
    # The naive pattern of the reference's loop (one byte per candidate and point, warp per
    # candidate): kept as a speed-up figure, it is not a ceiling.
    g_l2 = C.c_double()
    stack_bytes = sum((grid + 2 ** i - 1) ** 2 for i in range(depth))
    rc = abi.csm_measure_gather_rate(raw, C.c_size_t(stack_bytes), grid + 63,
                                     discretized_scan.ctypes.data_as(C.c_void_p), len(discretized_scan),
                                     1 << 21, 3, C.byref(g_l2))
    assert rc == 0, abi.csm_last_error()
    search_ms = sum(phase_ms) / steps
    all_lookups = sum(phase_scored) / steps * n_pts
    out["naive_pattern_speedup"] = {
        "g_l2_lookups_per_s": g_l2.value,
        "achieved_lookups_per_s": all_lookups / (search_ms * 1e-3),
        "speedup": all_lookups / (search_ms * 1e-3) / g_l2.value,
        "note": "G_L2 of SURVEY.md 8(d): one byte per (candidate, point), one warp per candidate, on an "
                "L2-resident table of the stack's size; the quad-phase layout serves 4 lookups per 4-byte "
                "load from neighbouring words, so this is a speed-up over the naive pattern, not a fraction "
                "of a ceiling"}
    return out
