"""The 3D oracle against the committed golden vectors (CPU only): pins the checker."""
import numpy as np

import golden_util as gu
import oracle_binding as ob


def _grid(arr, key, res):
    return ob.OracleGrid3D(res, arr[key + "_xyz"], arr[key + "_val"])


RES = {"random": 2.0, "world_hi": 0.1, "world_lo": 0.45, "fixture0": 0.05, "fixture1": 0.05,
       "fixture2": 0.05, "far_voxel": 0.1}


def test_oracle_reproduces_golden_stacks3d():
    doc, arr = gu.load3d()
    for case in doc["stacks"]:
        g = _grid(arr, case["grid"], RES[case["grid"]])
        assert g.grid_size == case["grid_size"]
        m = ob.OracleFCSM3D(g, g, np.zeros(4, np.float32), ob.options3d(case["depth"], case["frd"]))
        for i, lv in enumerate(case["levels"]):
            lo, dense = gu.canonical_level(*m.level(i))
            assert lo.tolist() == lv["lo"] and list(dense.shape) == lv["shape"] and gu.sha(dense) == lv["sha1"]


def check_result(r, e, num_scans=None):
    assert int(r.found) == e["found"]
    assert (r.num_scans if num_scans is None else num_scans) == e["num_scans"]
    if e["found"]:
        assert [r.scan_index, list(r.offset)] == [e["scan_index"], e["offset"]]
        assert gu.score_bits(r.score) == e["score_bits"]
        assert gu.score_bits(r.low_resolution_score) == e["low_bits"]
        assert gu.score_bits(r.rotational_score) == e["rot_bits"]
        assert [gu.score_bits(v) for v in r.pose] == e["pose_bits"]


def test_oracle_reproduces_golden_matches3d():
    doc, arr = gu.load3d()
    for case in doc["matches"]:
        m = ob.OracleFCSM3D(_grid(arr, case["grid"], RES[case["grid"]]),
                            _grid(arr, case["low_grid"], RES[case["low_grid"]]), arr[case["submap_hist"]],
                            np.array(case["options"]))
        r = m.match(arr[case["hi"]], arr[case["lo"]], arr[case["hist"]], case["min_score"],
                    np.array(case["node_pose"], np.float32), np.array(case["submap_pose"], np.float32),
                    case["gravity"], bool(case["full"]))
        check_result(r, case["expect"])


def test_oracle_reproduces_golden_discrete_scans3d():
    doc, arr = gu.load3d()
    for d in doc["discrete"]:
        case = doc["matches"][d["match"]]
        m = ob.OracleFCSM3D(_grid(arr, case["grid"], RES[case["grid"]]),
                            _grid(arr, case["low_grid"], RES[case["low_grid"]]), arr[case["submap_hist"]],
                            np.array(case["options"]))
        poses, rot, cells = m.discrete_scans(arr[case["hi"]], arr[case["hist"]],
                                             np.array(case["node_pose"], np.float32),
                                             np.array(case["submap_pose"], np.float32), case["gravity"])
        assert len(rot) == d["num_scans"] and cells[0, :3].tolist() == d["cells_first"]
        assert (gu.sha(poses), gu.sha(rot), gu.sha(cells)) == (d["poses_sha1"], d["rot_sha1"], d["cells_sha1"])


def test_fixture_matches_meet_the_reference_tolerances():
    """FastCorrelativeScanMatcher3DTest (fast_correlative_scan_matcher_3d_test.cc:146-204):
    score above kMinScore, rotational > 0.09, low-resolution > 0.14, pose within 0.05."""
    doc, _ = gu.load3d()
    truth = {"fixture0": ([0.31, -0.22, 0.4], 0.12), "fixture1": ([-0.55, 0.6, -0.1], -0.17),
             "fixture2": ([0.0, 0.0, 0.0], 0.0), "fixture0_full": ([0.31, -0.22, 0.4], 0.12)}
    as_float = lambda bits: float(np.uint32(bits).view(np.float32))
    for case in doc["matches"]:
        e = case["expect"]
        if case["name"].endswith("_far"):
            assert not e["found"]
        if case["name"] not in truth:
            continue
        t, theta = truth[case["name"]]
        pose = [as_float(b) for b in e["pose_bits"]]
        assert e["found"] and as_float(e["score_bits"]) > 0.1
        assert as_float(e["rot_bits"]) > 0.09 and as_float(e["low_bits"]) > 0.14
        assert np.linalg.norm(np.array(pose[:3]) - t) < 0.05
        assert abs(2 * np.arctan2(pose[6], pose[3]) - theta) < 0.05


def test_rotational_matcher_reference_expectations(csm):
    """rotational_scan_matcher_test.cc:29-66 on the oracle and on the product's host
    class (pure host code, runs without a device), and the two bit-identical."""
    doc, arr = gu.load3d()
    ref = doc["reference_rotational"]
    n = ref["buckets"]
    unit3, unit2 = np.zeros(n, np.float32), np.zeros(n, np.float32)
    unit3[ref["submap_unit"]] = 1.0
    unit2[ref["scan_unit"]] = 1.0
    for impl in (ob.rotational_match, csm.rotational_match):
        t = np.float32(0.0)
        while t < 1.0:
            expected = t / np.hypot(t, 1 - t)
            got = impl(unit3, unit2, 0.0, [t * np.float32(np.pi / n)])
            assert abs(got[0] - expected) < ref["tolerance"]
            t = np.float32(t + np.float32(0.1))
        h = np.array([1.0, 43.0, 0.5, 0.3123, 23.0, 42.0, 0.0], np.float32)
        s = impl(h, h, 0.0, [0.0, 1.0])
        assert abs(s[0] - 1.0) < 1e-6 and s[1] < 1.0
    rng = np.random.default_rng(3)
    a, b = rng.random(120).astype(np.float32), rng.random(120).astype(np.float32)
    angles = np.linspace(-3.2, 3.2, 257).astype(np.float32)
    assert np.array_equal(ob.rotational_match(a, b, 0.37, angles), csm.rotational_match(a, b, 0.37, angles))
