"""Parity at BASELINE.json's full sizes.  Against the oracle: every level of the C4 stack
(2000 x 2000, depth 7), C3-size MatchFullSubmap searches (1000 x 1000), C5-size 3D searches
(40 m room at 0.1 / 0.45 m, depth 8) -- live where the restatement finishes in seconds, and
against tests/golden/fullsize_golden.json (tools/make_golden_fullsize.py) everywhere.  Plus
size-independent properties: the C4 stack obeys its defining identities, C3-size searches
recover the scan's pose, do not depend on the order or the company of the searches in a
batch, and are reproducible bit for bit."""
import json
import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest

import fullsize_cases as fc
import golden_util as gu
import oracle_binding as ob
import synth

pytestmark = pytest.mark.gpu

GOLDEN = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden",
                                     "fullsize_golden.json")))


def test_c4_stack_levels_match_the_oracle(csm, gpu_ctx):
    """Every byte of every level of the 2000 x 2000, depth-7 stack against the oracle's
    PrecomputationGrid2D (literal SlidingWindowMaximum, FCSM2D.cc:41-161), and against the
    committed SHA-1s."""
    w = fc.c4_world()
    assert gu.sha(w.cells) == GOLDEN["c4"]["cells_sha1"], "the input generator drifted"
    om = ob.OracleFCSM2D(ob.OracleGrid(w.cells, w.res, w.max_x, w.max_y), 7.0, 0.5, 7)
    m = csm.Matcher(gpu_ctx, csm.Grid(w.cells, w.res, w.max_x, w.max_y), 7.0, 0.5, 7)
    for lvl in range(7):
        got = m.level(lvl)
        assert list(got.shape) == GOLDEN["c4"]["level_shape"][lvl]
        assert np.array_equal(got, om.level(lvl)), "level %d differs from the oracle" % lvl
        assert gu.sha(got) == GOLDEN["c4"]["level_sha1"][lvl], "level %d differs from the golden file" % lvl


def test_c3_full_submap_searches_match_the_oracle(csm, gpu_ctx):
    """Four MatchFullSubmap searches on a 1000 x 1000 submap (two scans from inside, two from
    another world), depth 7, min_score 0.6: winner, score bits and pose against the oracle run
    live (one thread per search) and against the golden file (FCSM2D.cc:211-262)."""
    w, clouds = fc.c3_case()
    assert gu.sha(w.cells) == GOLDEN["c3"]["cells_sha1"], "the input generator drifted"
    om = ob.OracleFCSM2D(ob.OracleGrid(w.cells, w.res, w.max_x, w.max_y), fc.LINEAR, fc.ANGULAR, 7)
    m = csm.Matcher(gpu_ctx, csm.Grid(w.cells, w.res, w.max_x, w.max_y), fc.LINEAR, fc.ANGULAR, 7)
    res = csm.match_batch(gpu_ctx, [m] * len(clouds), clouds, [fc.C3_MIN_SCORE] * len(clouds))
    with ThreadPoolExecutor(len(clouds)) as ex:
        want = list(ex.map(lambda c: om.match(c, fc.C3_MIN_SCORE), clouds))
    assert sum(1 for r in want if r.found) == 2 and sum(1 for r in want if not r.found) == 2
    for r_g, r_o, gold in zip(res, want, GOLDEN["c3"]["results"]):
        assert r_g.found == r_o.found == gold["found"]
        if not r_o.found:
            continue
        assert (r_g.scan_index, r_g.x_index_offset, r_g.y_index_offset) == \
               (r_o.scan_index, r_o.x_index_offset, r_o.y_index_offset) == \
               (gold["scan_index"], gold["x_index_offset"], gold["y_index_offset"])
        assert gu.score_bits(r_g.score) == gu.score_bits(r_o.score) == gold["score_bits"]
        assert (r_g.pose_x, r_g.pose_y, r_g.pose_theta) == (r_o.pose_x, r_o.pose_y, r_o.pose_theta)
        assert [r_g.pose_x, r_g.pose_y, r_g.pose_theta] == gold["pose"]


def test_c5_searches_match_the_oracle(csm, gpu_ctx):
    """C5 size (40 x 40 x 6 m room at 0.1 / 0.45 m, depth 8, full_resolution_depth 3,
    pose_graph.lua defaults): two Match searches against the oracle run live and all three
    searches (the third a MatchFullSubmap, ~90 s for the restatement) against the golden file."""
    w3, (hi_xyz, hi_val, lo_xyz, lo_val), nodes = fc.c5_case()
    assert gu.sha(hi_xyz) + gu.sha(hi_val) == GOLDEN["c5"]["voxels_sha1"], "the input generator drifted"
    hist = w3.submap_histogram()
    gm = csm.Matcher3D(gpu_ctx, 0.1, hi_xyz, hi_val, 0.45, lo_xyz, lo_val, hist, fc.C5_OPT)
    objs = [csm.Node3D(hi, lo, h, g) for hi, lo, h, g, _, _ in nodes]
    inits = np.array([nd[4].astype(np.float64) for nd in nodes])
    full = np.array([1 if nd[5] else 0 for nd in nodes], np.int32)
    res = csm.match3d_batch(gpu_ctx, [gm] * len(nodes), objs, [fc.C5_MIN_SCORE] * len(nodes), inits, None, full)
    for r_g, gold in zip(res, GOLDEN["c5"]["results"]):
        assert bool(r_g.found) == bool(gold["found"])
        if not gold["found"]:
            continue
        assert (r_g.scan_index, list(r_g.offset)) == (gold["scan_index"], gold["offset"])
        assert gu.score_bits(r_g.score) == gold["score_bits"]
        assert gu.score_bits(r_g.low_resolution_score) == gold["low_resolution_score_bits"]
        assert gu.score_bits(r_g.rotational_score) == gold["rotational_score_bits"]
        assert np.array_equal(np.array(r_g.pose, np.float64).astype(np.float32), np.array(gold["pose"], np.float32))
    assert sum(g["found"] for g in GOLDEN["c5"]["results"]) >= 2
    om = ob.OracleFCSM3D(ob.OracleGrid3D(0.1, hi_xyz, hi_val), ob.OracleGrid3D(0.45, lo_xyz, lo_val), hist,
                         np.array(fc.C5_OPT))

    def one(nd):
        hi, lo, h, g, init, _ = nd
        return om.match(hi, lo, h, fc.C5_MIN_SCORE, init, ob.IDENTITY_POSE7, g, False)

    with ThreadPoolExecutor(2) as ex:
        live = list(ex.map(one, nodes[:2]))
    for r_g, r_o in zip(res, live):
        assert bool(r_g.found) == bool(r_o.found) and r_g.num_scans == r_o.num_scans
        if r_o.found:
            assert (r_g.scan_index, tuple(r_g.offset)) == (r_o.scan_index, tuple(r_o.offset))
            for name in ("score", "low_resolution_score", "rotational_score"):
                assert gu.score_bits(getattr(r_g, name)) == gu.score_bits(getattr(r_o, name)), name


def test_c4_stack_identities(csm, gpu_ctx):
    w = synth.World(2000, 2000, seed=4000)
    g = csm.Grid(w.cells, w.res, w.max_x, w.max_y)
    m = csm.Matcher(gpu_ctx, g, 7.0, 0.5, 7)
    lut = g.quantization_table()
    prev = m.level(0)
    assert np.array_equal(prev, lut[w.cells])               # level 0 = quantised cells
    for lvl in range(1, 7):
        s = 1 << (lvl - 1)
        cur = m.level(lvl)
        assert cur.shape == (2000 + 2 * s - 1, 2000 + 2 * s - 1)
        big = np.zeros((prev.shape[0] + 2 * s, prev.shape[1] + 2 * s), np.uint8)  # prev, s zeros around
        big[s:s + prev.shape[0], s:s + prev.shape[1]] = prev
        h, wd = cur.shape
        # level i+1 at (X, Y) = max of level i at {X - s, X} x {Y - s, Y} (window-end coordinates)
        want = np.maximum(np.maximum(big[s:s + h, s:s + wd], big[0:h, s:s + wd]),
                          np.maximum(big[s:s + h, 0:wd], big[0:h, 0:wd]))
        assert np.array_equal(cur, want), "level %d" % lvl
        prev = cur
    # a checksum of checksums over the quad-phase words: byte 3 of every word is the level itself
    q = m.quad(6)
    two_s, jn, n_i = q.shape[0], q.shape[2], q.shape[3]
    lvl6 = (q >> 24).astype(np.uint8).transpose(2, 0, 3, 1).reshape(jn * two_s, n_i * two_s)
    assert np.array_equal(lvl6[:prev.shape[0], :prev.shape[1]], prev)


def test_c3_round_trip_order_independence_and_reproducibility(csm, gpu_ctx):
    w = synth.World(1000, 1000, seed=3000)
    decoy = synth.World(1000, 1000, seed=3999)
    g = csm.Grid(w.cells, w.res, w.max_x, w.max_y)
    m = csm.Matcher(gpu_ctx, g, 7.0, float(np.deg2rad(30.0)), 7)
    clouds, truth = [], []
    for k in range(12):
        src = w if k % 3 == 0 else decoy
        pose, pts = synth.make_scan(src, 3000 + k)
        clouds.append(pts)
        truth.append(pose if src is w else None)
    res = csm.match_batch(gpu_ctx, [m] * 12, clouds, [0.6] * 12)
    key = lambda r: (r.found, r.scan_index, r.x_index_offset, r.y_index_offset,
                     np.float32(r.score).tobytes(), r.pose_x, r.pose_y, r.pose_theta)
    found_true = 0
    for r, t in zip(res, truth):
        if t is None:
            continue
        if r.found:
            found_true += 1
            assert abs(r.pose_x - t[0]) < 0.1 and abs(r.pose_y - t[1]) < 0.1
            d = (r.pose_theta - t[2] + np.pi) % (2 * np.pi) - np.pi
            assert abs(d) < np.deg2rad(1.0)
            assert r.score > 0.6
    assert found_true >= 2  # a true positive can score below 0.6 (unknown cells, noise): it is then a miss for the reference too
    again = csm.match_batch(gpu_ctx, [m] * 12, clouds, [0.6] * 12)
    assert [key(a) for a in again] == [key(r) for r in res]          # reproducible
    perm = [7, 0, 11, 3, 5, 1, 9, 2, 10, 4, 8, 6]
    shuffled = csm.match_batch(gpu_ctx, [m] * 12, [clouds[i] for i in perm], [0.6] * 12)
    for slot, i in enumerate(perm):
        assert key(shuffled[slot]) == key(res[i])                    # order / company independent
    alone = csm.match_batch(gpu_ctx, [m], [clouds[0]], [0.6])[0]
    assert key(alone) == key(res[0])
