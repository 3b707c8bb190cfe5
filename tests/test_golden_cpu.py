"""The oracle against the committed golden vectors (CPU only): pins the checker."""
import numpy as np

import golden_util as gu
import oracle_binding as ob


def test_oracle_reproduces_golden_stacks():
    doc, arr = gu.load()
    for case in doc["stacks"]:
        m = ob.OracleFCSM2D(ob.OracleGrid(arr[case["cells"]], 0.05, 1.0, 1.0), 1.0, 0.1, case["depth"])
        for i, lv in enumerate(case["levels"]):
            got = m.level(i)
            assert list(got.shape) == lv["shape"] and gu.sha(got) == lv["sha1"]


def test_oracle_reproduces_reference_discretize_fixture():
    doc, _ = gu.load()
    ref = doc["reference_discretize"]
    pts = np.array(ref["points"], dtype=np.float32)
    xy, _ = ob.discretize(ref["resolution"], ref["max"][0], ref["max"][1], ref["cells"][0],
                          ref["cells"][1], pts, (0.0, 0.0, 0.0), 0, 0, 0.0, shrink=False)
    assert xy.shape == (1, 7, 2)
    assert xy[0].tolist() == ref["expected_xy"]


def test_oracle_reproduces_golden_discretize():
    doc, arr = gu.load()
    for case in doc["discretize"]:
        w = doc["worlds"][case["world"]]
        cells = arr["world%d" % case["world"]]
        c = arr["cloud%d" % case["cloud"]]
        nap, step, ns, nlin = ob.search_parameters(case["linear"], case["angular"], c, w["res"])
        assert ns == case["num_scans"]
        xy, bounds = ob.discretize(w["res"], w["max"][0], w["max"][1], cells.shape[1], cells.shape[0],
                                   c, case["init"], nlin, nap, step)
        assert gu.sha(xy) == case["xy_sha1"] and gu.sha(bounds) == case["bounds_sha1"]


def test_oracle_reproduces_golden_matches():
    doc, arr = gu.load()
    for case in doc["matches"]:
        g = ob.OracleGrid(arr[case["cells"]], case["res"], case["max"][0], case["max"][1])
        m = ob.OracleFCSM2D(g, case["linear"], case["angular"], case["depth"])
        r = m.match(arr[case["cloud"]], case["min_score"],
                    initial=None if case["init"] is None else tuple(case["init"]))
        assert r.found == case["found"]
        if r.found:
            assert [r.scan_index, r.x_index_offset, r.y_index_offset] == case["winner"]
            assert gu.score_bits(r.score) == case["score_bits"]
            assert [r.pose_x, r.pose_y, r.pose_theta] == case["pose"]


def test_oracle_reproduces_golden_rt():
    doc, arr = gu.load()
    for case in doc["rt"]:
        w = doc["worlds"][case["world"]]
        g = ob.OracleGrid(arr["world%d" % case["world"]], w["res"], w["max"][0], w["max"][1])
        r = ob.rt_match(g, case["options"], tuple(case["init"]), arr["cloud%d" % case["cloud"]])
        assert [r.scan_index, r.x_index_offset, r.y_index_offset] == case["winner"]
        assert gu.score_bits(r.score) == case["score_bits"]
        assert [r.pose_x, r.pose_y, r.pose_theta] == case["pose"]
