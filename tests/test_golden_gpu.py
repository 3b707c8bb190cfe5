"""The CUDA path against the committed golden vectors (no generator, no oracle call)."""
import numpy as np
import pytest

import golden_util as gu

pytestmark = pytest.mark.gpu


def test_cuda_stacks_match_golden(csm, gpu_ctx):
    doc, arr = gu.load()
    for case in doc["stacks"]:
        g = csm.Grid(arr[case["cells"]], 0.05, 1.0, 1.0)
        m = csm.Matcher(gpu_ctx, g, 1.0, 0.1, case["depth"])
        for i, lv in enumerate(case["levels"]):
            got = m.level(i)
            assert list(got.shape) == lv["shape"] and gu.sha(got) == lv["sha1"], (case["cells"], i)


def test_cuda_discretize_matches_reference_fixture(csm, gpu_ctx):
    # correlative_scan_matcher_test.cc:72-96: exact cell indices, zero rotation.  The
    # test's SearchParameters(0, 0, 0., 0.) is one scan at angle 0; a zero angular
    # window yields exactly that.
    doc, _ = gu.load()
    ref = doc["reference_discretize"]
    pts = np.array(ref["points"], dtype=np.float32)
    xy, _ = csm.discretize(gpu_ctx, ref["resolution"], ref["max"][0], ref["max"][1],
                           ref["cells"][0], ref["cells"][1], pts, (0.0, 0.0, 0.0), 0.0, 0.0,
                           shrink=False)
    assert xy.shape == (1, 7, 2)
    assert xy[0].tolist() == ref["expected_xy"]


def test_cuda_discretize_matches_golden(csm, gpu_ctx):
    doc, arr = gu.load()
    for case in doc["discretize"]:
        w = doc["worlds"][case["world"]]
        cells = arr["world%d" % case["world"]]
        xy, bounds = csm.discretize(gpu_ctx, w["res"], w["max"][0], w["max"][1], cells.shape[1],
                                    cells.shape[0], arr["cloud%d" % case["cloud"]], case["init"],
                                    case["linear"], case["angular"])
        assert xy.shape[0] == case["num_scans"]
        assert xy[0, :4].tolist() == case["xy_first"] and bounds[0].tolist() == case["bounds_first"]
        assert gu.sha(xy) == case["xy_sha1"] and gu.sha(bounds) == case["bounds_sha1"]


def test_cuda_matches_golden(csm, gpu_ctx):
    doc, arr = gu.load()
    grids, matchers, jobs = {}, {}, []
    for case in doc["matches"]:
        key = (case["cells"], case["linear"], case["angular"], case["depth"])
        if case["cells"] not in grids:
            grids[case["cells"]] = csm.Grid(arr[case["cells"]], case["res"], case["max"][0],
                                            case["max"][1])
        if key not in matchers:
            matchers[key] = csm.Matcher(gpu_ctx, grids[case["cells"]], case["linear"],
                                        case["angular"], case["depth"])
        jobs.append((matchers[key], case))
    res = csm.match_batch(gpu_ctx, [m for m, _ in jobs], [arr[c["cloud"]] for _, c in jobs],
                          [c["min_score"] for _, c in jobs],
                          initial=[None if c["init"] is None else tuple(c["init"]) for _, c in jobs])
    for (_, case), r in zip(jobs, res):
        assert r.found == case["found"]
        if r.found:
            assert [r.scan_index, r.x_index_offset, r.y_index_offset] == case["winner"]
            assert gu.score_bits(r.score) == case["score_bits"]
            # poses: identical fp64 formulas on both sides; 1e-5 relative is the stated bar
            for got, want in zip([r.pose_x, r.pose_y, r.pose_theta], case["pose"]):
                assert abs(got - want) <= 1e-5 * max(1.0, abs(want))


def test_cuda_rt_matches_golden(csm, gpu_ctx):
    doc, arr = gu.load()
    for case in doc["rt"]:
        w = doc["worlds"][case["world"]]
        g = csm.Grid(arr["world%d" % case["world"]], w["res"], w["max"][0], w["max"][1])
        r = csm.rt_match_batch(gpu_ctx, [g], [arr["cloud%d" % case["cloud"]]], [tuple(case["init"])],
                               **case["options"])[0]
        assert [r.scan_index, r.x_index_offset, r.y_index_offset] == case["winner"]
        assert gu.score_bits(r.score) == case["score_bits"]
        for got, want in zip([r.pose_x, r.pose_y, r.pose_theta], case["pose"]):
            assert abs(got - want) <= 1e-5 * max(1.0, abs(want))
