"""Pins the oracle (oracle/csm_oracle.h, a restatement) and the committed golden vectors to the
REFERENCE'S OWN CODE: oracle/_ref/libref.so is the reference's 2D scan-matcher, grid, value-table
and point-cloud sources compiled unmodified (oracle/Makefile `ref`, oracle/ref_capi.cc).  CPU only.

Everything is compared bit for bit: precomputation levels byte by byte, discrete scans and
shrunk bounds as int32, scores by their fp32 bit pattern, poses as the doubles they are.
"""
import numpy as np
import pytest

import golden_util as gu
import oracle_binding as ob
import ref_binding as rb
import synth
import tsdf_util

pytestmark = pytest.mark.skipif(not rb.available(), reason="oracle/_ref/libref.so was not built (no /root/reference)")


def _same_match(a, b):
    assert a.found == b.found
    if a.found:  # the reference leaves score and pose untouched when nothing reaches min_score
        assert gu.score_bits(a.score) == gu.score_bits(b.score)
        assert [a.pose_x, a.pose_y, a.pose_theta] == [b.pose_x, b.pose_y, b.pose_theta]


def test_value_tables_are_identical():
    t = np.empty(65536, dtype=np.float32)
    ob.lib().oracle_correspondence_cost_table(t)
    assert t.tobytes() == rb.correspondence_cost_table().tobytes()
    for p in (0.55, 0.49, 0.7, 0.3):
        assert ob.apply_odds_table(p).tobytes() == rb.apply_odds_table(p).tobytes()
    for p in np.linspace(0.0, 1.0, 1001, dtype=np.float32):
        assert ob.lib().oracle_probability_to_cost_value(float(p)) == rb.lib().oracle_probability_to_cost_value(
            float(p))


def test_reference_reproduces_golden_stacks():
    doc, arr = gu.load()
    for case in doc["stacks"]:
        m = rb.RefFCSM2D(rb.RefGrid(arr[case["cells"]], 0.05, 1.0, 1.0), 1.0, 0.1, case["depth"])
        for i, lv in enumerate(case["levels"]):
            got = m.level(i)
            assert list(got.shape) == lv["shape"] and gu.sha(got) == lv["sha1"]


def test_reference_reproduces_golden_discretize():
    doc, arr = gu.load()
    ref = doc["reference_discretize"]
    pts = np.array(ref["points"], dtype=np.float32)
    xy, _ = rb.discretize(ref["resolution"], ref["max"][0], ref["max"][1], ref["cells"][0], ref["cells"][1], pts,
                          (0.0, 0.0, 0.0), 0, 0, 0.0, shrink=False)
    assert xy[0].tolist() == ref["expected_xy"]
    for case in doc["discretize"]:
        w = doc["worlds"][case["world"]]
        cells = arr["world%d" % case["world"]]
        c = arr["cloud%d" % case["cloud"]]
        nap, step, ns, nlin = rb.search_parameters(case["linear"], case["angular"], c, w["res"])
        assert (nap, step, ns, nlin) == ob.search_parameters(case["linear"], case["angular"], c, w["res"])
        assert ns == case["num_scans"]
        xy, bounds = rb.discretize(w["res"], w["max"][0], w["max"][1], cells.shape[1], cells.shape[0], c,
                                   case["init"], nlin, nap, step)
        assert gu.sha(xy) == case["xy_sha1"] and gu.sha(bounds) == case["bounds_sha1"]


def test_reference_reproduces_golden_matches():
    doc, arr = gu.load()
    for case in doc["matches"]:
        g = rb.RefGrid(arr[case["cells"]], case["res"], case["max"][0], case["max"][1])
        m = rb.RefFCSM2D(g, case["linear"], case["angular"], case["depth"])
        r = m.match(arr[case["cloud"]], case["min_score"],
                    initial=None if case["init"] is None else tuple(case["init"]))
        assert r.found == case["found"]
        if r.found:
            assert gu.score_bits(r.score) == case["score_bits"]
            assert [r.pose_x, r.pose_y, r.pose_theta] == case["pose"]


def test_reference_reproduces_golden_rt():
    doc, arr = gu.load()
    for case in doc["rt"]:
        w = doc["worlds"][case["world"]]
        g = rb.RefGrid(arr["world%d" % case["world"]], w["res"], w["max"][0], w["max"][1])
        r = rb.rt_match(g, case["options"], tuple(case["init"]), arr["cloud%d" % case["cloud"]])
        assert gu.score_bits(r.score) == case["score_bits"]
        assert [r.pose_x, r.pose_y, r.pose_theta] == case["pose"]


@pytest.mark.parametrize("seed", [11, 12, 13])
def test_oracle_equals_reference_on_random_worlds(seed):
    """Fresh worlds and scans, not in the golden file: stacks, local matches around a perturbed
    pose (found and not found), MatchFullSubmap, the real-time matcher."""
    rng = np.random.default_rng(seed)
    w = synth.World(int(rng.integers(150, 260)), int(rng.integers(150, 260)), seed)
    depth = int(rng.integers(3, 8))
    og = ob.OracleGrid(w.cells, w.res, w.max_x, w.max_y)
    rg = rb.RefGrid(w.cells, w.res, w.max_x, w.max_y)
    assert rg.cells().tobytes() == np.ascontiguousarray(w.cells).tobytes()
    om = ob.OracleFCSM2D(og, 1.5, 0.4, depth)
    rm = rb.RefFCSM2D(rg, 1.5, 0.4, depth)
    for i in range(depth):
        assert om.level(i).tobytes() == rm.level(i).tobytes()
    rt_opts = {"linear_search_window": 0.3, "angular_search_window": 0.15,
               "translation_delta_cost_weight": 1e-1, "rotation_delta_cost_weight": 1e-1}
    for k in range(4):
        pose, pts = synth.make_scan(w, 100 * seed + k)
        init = (pose[0] + rng.uniform(-0.6, 0.6), pose[1] + rng.uniform(-0.6, 0.6), pose[2] + rng.uniform(-0.2, 0.2))
        for min_score in (0.3, 0.98):
            _same_match(om.match(pts, min_score, initial=init), rm.match(pts, min_score, initial=init))
        near = (pose[0] + rng.uniform(-0.1, 0.1), pose[1] + rng.uniform(-0.1, 0.1), pose[2] + rng.uniform(-0.05, 0.05))
        _same_match(ob.rt_match(og, rt_opts, near, pts), rb.rt_match(rg, rt_opts, near, pts))
    pose, pts = synth.make_scan(w, 100 * seed + 50)
    _same_match(om.match(pts, 0.4), rm.match(pts, 0.4))


def test_oracle_equals_reference_on_a_tsdf_grid():
    w = synth.World(160, 140, 21)
    tv, wv = tsdf_util.from_world(w, truncation=0.3, max_weight=10.0, seed=3)
    ot = ob.OracleTsdf(tv, wv, w.res, w.max_x, w.max_y, 0.3, 10.0)
    rt = rb.RefTsdf(tv, wv, w.res, w.max_x, w.max_y, 0.3, 10.0)
    om = ob.OracleFCSM2D(ot, 1.0, 0.3, 5)
    rm = rb.RefFCSM2D(rt, 1.0, 0.3, 5)
    for i in range(5):
        assert om.level(i).tobytes() == rm.level(i).tobytes()
    rng = np.random.default_rng(5)
    for k in range(3):
        pose, pts = synth.make_scan(w, 900 + k)
        init = (pose[0] + rng.uniform(-0.4, 0.4), pose[1] + rng.uniform(-0.4, 0.4), pose[2] + rng.uniform(-0.1, 0.1))
        _same_match(om.match(pts, 0.2, initial=init), rm.match(pts, 0.2, initial=init))


def test_constraint_batch_of_the_oracle_equals_the_reference_driver():
    """Three submaps with their own poses, local and global searches mixed, 4 threads each side."""
    rng = np.random.default_rng(77)
    worlds = [synth.World(180, 160, 40 + k) for k in range(3)]
    poses = np.array([[0.3, -0.2, 0.1], [-1.0, 0.5, -0.4], [2.0, 1.0, 0.7]])
    sub, full, rel, clouds = [], [], [], []
    for k in range(12):
        s = k % 3
        pose, pts = synth.make_scan(worlds[(s + (k % 4 == 3)) % 3], 700 + k)  # every 4th from another world
        sub.append(s)
        full.append(int(k % 2))
        rel.append([pose[0] + rng.uniform(-0.3, 0.3), pose[1] + rng.uniform(-0.3, 0.3), pose[2] + rng.uniform(-0.1, 0.1)])
        clouds.append(pts)
    rel[5][0] = 40.0  # beyond max_constraint_distance: skipped by both
    kw = dict(min_score=0.5, global_min_score=0.55, linear_window=2.0, angular_window=0.35, depth=5)
    og = [ob.OracleGrid(w.cells, w.res, w.max_x, w.max_y) for w in worlds]
    rg = [rb.RefGrid(w.cells, w.res, w.max_x, w.max_y) for w in worlds]
    _, o_out, _ = ob.constraint_batch(4, og, poses, sub, full, rel, clouds, **kw)
    _, r_out, _ = rb.constraint_batch(4, rg, poses, sub, full, rel, clouds, **kw)
    assert sum(o.found for o in o_out) >= 6
    assert o_out[5].found == 0 and r_out[5].found == 0
    for o, r in zip(o_out, r_out):
        _same_match(o, r)


def test_reference_reproduces_fullsize_golden():
    """BASELINE.json's full sizes (tests/golden/fullsize_golden.json, written by the oracle):
    every level of the C4 stack (2000 x 2000, depth 7) and the C3 MatchFullSubmap winners
    (1000 x 1000, 7 m / 30 deg window ignored by the full-submap search, depth 7)."""
    import json
    import os
    from concurrent.futures import ThreadPoolExecutor

    import fullsize_cases as fc
    doc = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "fullsize_golden.json")))
    w = fc.c4_world()
    assert gu.sha(w.cells) == doc["c4"]["cells_sha1"]
    m = rb.RefFCSM2D(rb.RefGrid(w.cells, w.res, w.max_x, w.max_y), 7.0, 0.5, 7)
    assert [gu.sha(m.level(i)) for i in range(7)] == doc["c4"]["level_sha1"]
    del m
    w, clouds = fc.c3_case()
    assert gu.sha(w.cells) == doc["c3"]["cells_sha1"]
    m = rb.RefFCSM2D(rb.RefGrid(w.cells, w.res, w.max_x, w.max_y), fc.LINEAR, fc.ANGULAR, 7)
    with ThreadPoolExecutor(len(clouds)) as ex:
        res = list(ex.map(lambda c: m.match(c, fc.C3_MIN_SCORE), clouds))
    for r, want in zip(res, doc["c3"]["results"]):
        assert r.found == want["found"]
        if r.found:
            assert gu.score_bits(r.score) == want["score_bits"]
            assert [r.pose_x, r.pose_y, r.pose_theta] == want["pose"]


def test_range_data_inserter_of_the_oracle_equals_the_reference():
    """ProbabilityGridRangeDataInserter2D: three successive scans into a small grid
    that has to grow, hits and misses, free space on and off; cells and limits after every insert."""
    rng = np.random.default_rng(5)
    for free_space in (True, False):
        og = ob.OracleGrid(np.zeros((20, 24), np.uint16), 0.05, 0.6, 0.5)
        rg = rb.RefGrid(np.zeros((20, 24), np.uint16), 0.05, 0.6, 0.5)
        for k in range(3):
            origin = rng.uniform(-0.5, 0.5, 2).astype(np.float32)
            ang = rng.uniform(-np.pi, np.pi, 120)
            rad = rng.uniform(0.3, 6.0, 120)
            pts = np.stack([origin[0] + rad * np.cos(ang), origin[1] + rad * np.sin(ang), np.zeros(120)], 1)
            returns, misses = pts[:100].astype(np.float32), pts[100:].astype(np.float32)
            ob.grid_insert(og, origin, returns, misses, 0.55 + 0.05 * k, 0.49 - 0.03 * k, free_space)
            cells, max_x, max_y = rg.insert(origin, returns, misses, 0.55 + 0.05 * k, 0.49 - 0.03 * k, free_space)
            assert (og.nx, og.ny, og.max_x, og.max_y) == (rg.nx, rg.ny, max_x, max_y)
            assert og.cells.tobytes() == cells.tobytes()
        assert (og.cells > 0).sum() > (1000 if free_space else 250)


@pytest.mark.parametrize("opts", [
    {},                                                                # trajectory_builder_2d.lua defaults
    {"update_free_space": True},
    {"project_sdf_distance_to_scan_normal": False, "update_weight_angle_scan_normal_to_ray_kernel_bandwidth": 0.0},
    {"update_weight_range_exponent": 2, "update_weight_distance_cell_to_hit_kernel_bandwidth": 0.0},
    {"truncation_distance": 0.15, "maximum_weight": 3.0, "num_normal_samples": 8, "sample_radius": 0.2},
])
def test_tsdf_range_data_inserter_of_the_oracle_equals_the_reference(opts):
    """TSDFRangeDataInserter2D with its normal estimation: five scans of a synthetic
    office into a small TSDF2D that has to grow; TSD cells, weight cells and limits after every scan."""
    w = synth.World(200, 180, seed=33)
    trunc = opts.get("truncation_distance", 0.3)
    maxw = opts.get("maximum_weight", 10.0)
    zero = np.zeros((40, 40), np.uint16)
    og = ob.OracleTsdf(zero, zero, 0.05, 1.0, 1.0, trunc, maxw)
    rg = rb.RefTsdf(zero, zero, 0.05, 1.0, 1.0, trunc, maxw)
    for k in range(5):
        pose, pts = synth.make_scan(w, 400 + k, max_range=6.0)
        c, s = np.cos(pose[2]), np.sin(pose[2])
        world = np.stack([pose[0] + c * pts[:, 0] - s * pts[:, 1], pose[1] + s * pts[:, 0] + c * pts[:, 1],
                          np.zeros(len(pts))], axis=1).astype(np.float32)
        ob.tsdf_insert(og, pose[:2], world, opts)
        rb.tsdf_insert(rg, pose[:2], world, opts)
        assert (og.nx, og.ny, og.max_x, og.max_y) == (rg.nx, rg.ny, rg.max_x, rg.max_y)
        assert og.tsd_cells.tobytes() == rg.tsd_cells.tobytes()
        assert og.weight_cells.tobytes() == rg.weight_cells.tobytes()
    assert (og.weight_cells > 0).sum() > 1500 and og.nx > 40
