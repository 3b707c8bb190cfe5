"""The 3D oracle (oracle/csm_oracle3d.h) and the committed 3D golden vectors against the REFERENCE'S
OWN 3D sources in oracle/_ref/libref.so: HybridGrid, PrecomputationGridStack3D,
FastCorrelativeScanMatcher3D (with RotationalScanMatcher and the low-resolution matcher) and
RealTimeCorrelativeScanMatcher3D, compiled unmodified (oracle/ref_capi3d.cc).  CPU only.

Byte-exact: HybridGrid contents and grid_size, every precomputation level.  Bit-exact under the
Eigen stand-in's scalar formulas (shared with the oracle): scores,
rotational / low-resolution scores and poses.  The reference's Result has no winner indices; a
different winner shows as a different pose.
"""
import numpy as np
import pytest

import golden_util as gu
import oracle_binding as ob
import ref_binding as rb

pytestmark = pytest.mark.skipif(not rb.available(), reason="oracle/_ref/libref.so was not built (no /root/reference)")

RES = {"random": 2.0, "world_hi": 0.1, "world_lo": 0.45, "fixture0": 0.05, "fixture1": 0.05,
       "fixture2": 0.05, "far_voxel": 0.1}


def _grid(arr, key):
    return rb.RefGrid3D(RES[key], arr[key + "_xyz"], arr[key + "_val"])


def test_hybrid_grid_round_trip_and_size():
    _, arr = gu.load3d()
    for key in RES:
        o = ob.OracleGrid3D(RES[key], arr[key + "_xyz"], arr[key + "_val"])
        r = _grid(arr, key)
        assert o.grid_size == r.grid_size
        (oxyz, oval), (rxyz, rval) = o.cells(), r.cells()
        assert oxyz.tobytes() == rxyz.tobytes() and oval.tobytes() == rval.tobytes()


def test_reference_reproduces_golden_stacks3d():
    doc, arr = gu.load3d()
    for case in doc["stacks"]:
        g = _grid(arr, case["grid"])
        assert g.grid_size == case["grid_size"]
        m = rb.RefFCSM3D(g, g, np.zeros(4, np.float32), ob.options3d(case["depth"], case["frd"]))
        for i, lv in enumerate(case["levels"]):
            lo, dense = gu.canonical_level(*m.level(i))
            assert lo.tolist() == lv["lo"] and list(dense.shape) == lv["shape"] and gu.sha(dense) == lv["sha1"]


def test_reference_reproduces_golden_matches3d():
    doc, arr = gu.load3d()
    found = 0
    for case in doc["matches"]:
        m = rb.RefFCSM3D(_grid(arr, case["grid"]), _grid(arr, case["low_grid"]), arr[case["submap_hist"]],
                         np.array(case["options"]))
        r = m.match(arr[case["hi"]], arr[case["lo"]], arr[case["hist"]], case["min_score"],
                    np.array(case["node_pose"], np.float32), np.array(case["submap_pose"], np.float32),
                    case["gravity"], bool(case["full"]))
        e = case["expect"]
        assert int(r.found) == e["found"], case["name"]
        if e["found"]:
            found += 1
            assert gu.score_bits(r.score) == e["score_bits"], case["name"]
            assert gu.score_bits(r.low_resolution_score) == e["low_bits"], case["name"]
            assert gu.score_bits(r.rotational_score) == e["rot_bits"], case["name"]
            assert [gu.score_bits(v) for v in r.pose] == e["pose_bits"], case["name"]
    assert found >= 4


def test_rotational_matcher_equals_reference():
    rng = np.random.default_rng(3)
    for n in (4, 30, 120):
        a = rng.uniform(0, 5, n).astype(np.float32)
        b = rng.uniform(0, 5, n).astype(np.float32)
        angles = rng.uniform(-0.5, 0.5, 9).astype(np.float32)
        assert ob.rotational_match(a, b, 0.3, angles).tobytes() == rb.rotational_match(a, b, 0.3, angles).tobytes()
    z = np.zeros(8, np.float32)
    assert ob.rotational_match(z, z, 0.0, np.zeros(1, np.float32)).tobytes() == \
        rb.rotational_match(z, z, 0.0, np.zeros(1, np.float32)).tobytes()


def test_real_time_matcher3d_equals_reference():
    _, arr = gu.load3d()
    og = ob.OracleGrid3D(RES["fixture0"], arr["fixture0_xyz"], arr["fixture0_val"])
    rg = _grid(arr, "fixture0")
    rng = np.random.default_rng(9)
    xyz = arr["fixture0_xyz"]
    pts = (xyz[rng.choice(len(xyz), 60, replace=False)].astype(np.float32) * np.float32(RES["fixture0"]))
    opts = np.array([0.15, 0.05, 1e-1, 1e-1])
    for k in range(3):
        init = np.array([0.05 * k, -0.03, 0.02, 1, 0, 0, 0], np.float32)
        so, po, _, _ = ob.rt_match3d(og, opts, init, pts)
        sr, pr = rb.rt_match3d(rg, opts, init, pts)
        assert np.float32(so).tobytes() == np.float32(sr).tobytes()
        assert po.tobytes() == pr.tobytes()


def test_reference_reproduces_fullsize_c5_golden():
    """BASELINE.json's C5 size (40 x 40 x 6 m at 0.1 / 0.45 m, depth 8, full_resolution_depth 3):
    the results the oracle wrote into tests/golden/fullsize_golden.json, from the reference."""
    import json
    import os
    from concurrent.futures import ThreadPoolExecutor

    import fullsize_cases as fc
    doc = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "fullsize_golden.json")))["c5"]
    w3, vox, nodes = fc.c5_case()
    hi_xyz, hi_val, lo_xyz, lo_val = vox
    assert gu.sha(hi_xyz) + gu.sha(hi_val) == doc["voxels_sha1"]
    m = rb.RefFCSM3D(rb.RefGrid3D(0.1, hi_xyz, hi_val), rb.RefGrid3D(0.45, lo_xyz, lo_val), w3.submap_histogram(),
                     np.array(fc.C5_OPT))

    def one(nd):
        hi, lo, hist, grav, init, full = nd
        return m.match(hi, lo, hist, fc.C5_MIN_SCORE, init, ob.IDENTITY_POSE7, grav, bool(full))

    with ThreadPoolExecutor(len(nodes)) as ex:
        res = list(ex.map(one, nodes))
    for r, want in zip(res, doc["results"]):
        assert int(bool(r.found)) == want["found"]
        if want["found"]:
            assert gu.score_bits(r.score) == want["score_bits"]
            assert gu.score_bits(r.low_resolution_score) == want["low_resolution_score_bits"]
            assert gu.score_bits(r.rotational_score) == want["rotational_score_bits"]
            assert [float(x) for x in r.pose] == want["pose"]


def test_range_data_inserter3d_of_the_oracle_equals_the_reference():
    """RangeDataInserter3D: scans into an empty grid and into a filled one that has
    to grow, 0 / 2 / 5 free-space voxels; the non-default voxels, their values and grid_size()."""
    import synth3d
    w = synth3d.World3D(5, size_xy=8.0)
    rng = np.random.default_rng(5)
    for free in (0, 2, 5):
        for start in ("empty", "filled"):
            xyz, val = (None, None) if start == "empty" else w.voxels(0.2)
            res = 0.1 if start == "empty" else 0.2
            og, rg = ob.OracleGrid3D(res, xyz, val), rb.RefGrid3D(res, xyz, val)
            for k in range(4):
                o = np.array([rng.uniform(-2, 2), rng.uniform(-2, 2), rng.uniform(0.2, 1.5)], np.float32)
                pts = w.points[rng.choice(len(w.points), 300, replace=False)].astype(np.float32)
                if k == 3:
                    pts = np.concatenate([pts, np.array([[30.0, -25.0, 2.0]], np.float32)])  # grows the grid
                og.insert(o, pts, 0.6, 0.45, free)
                rg.insert(o, pts, 0.6, 0.45, free)
                (oxyz, oval), (rxyz, rval) = og.cells(), rg.cells()
                assert oxyz.tobytes() == rxyz.tobytes() and oval.tobytes() == rval.tobytes()
                assert og.grid_size == rg.grid_size
            assert og.grid_size > 128 and len(oval) > 500
