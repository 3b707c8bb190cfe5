"""Oracle restatement of the TSDF2D branch of RealTimeCorrelativeScanMatcher2D against the
known answers of the reference's own tests (real_time_correlative_scan_matcher_2d_test.cc:
143-201: a perfectly aligned candidate scores ~1, a candidate one cell off scores
1 - 4/(7*6)) and of tsd_value_converter_test.cc (value range, clamping)."""
import numpy as np

import oracle_binding as ob
import tsdf_util

# The reference fixture's seven-point "L" (real_time_correlative_scan_matcher_2d_test.cc:55-61).
POINTS = np.array([[0.025, 0.175, 0.0], [-0.025, 0.175, 0.0], [-0.075, 0.175, 0.0],
                   [-0.125, 0.175, 0.0], [-0.125, 0.125, 0.0], [-0.125, 0.075, 0.0],
                   [-0.125, 0.025, 0.0]], dtype=np.float32)


def _fixture():
    # MapLimits(0.05, (0.3, 0.5), CellLimits(20, 20)), truncation 0.3, max weight 1.0 (:68-70);
    # range data origin (0.5, -0.5) (:90).
    tv, wv = tsdf_util.paint_from_points(20, 20, 0.05, 0.3, 0.5, POINTS[:, :2], (0.5, -0.5), 0.3, 1.0)
    return ob.OracleTsdf(tv, wv, 0.05, 0.3, 0.5, 0.3, 1.0)


def test_tsd_value_converter_range_and_clamping():
    # tsd_value_converter_test.cc: values live in [1, 32767]; out-of-range input clamps.
    tv, wv = ob.tsdf_values(0.3, 10.0, [-0.3, 0.3, -5.0, 5.0, 0.0], [0.0, 10.0, -1.0, 99.0, 5.0])
    assert tv.tolist() == [1, 32767, 1, 32767, 16384]
    assert wv.tolist() == [1, 32767, 1, 32767, 16384]


def test_score_perfect_candidate_tsdf():
    g = _fixture()
    score = g.score_candidate(POINTS, 0, 0)
    assert abs(score - 1.0) < 1e-1 and score > 0.95  # reference :160-161


def test_score_partially_correct_candidate_tsdf():
    g = _fixture()
    score = g.score_candidate(POINTS, 0, 1)
    # three points stay on the surface, four sit one cell (0.05 of 0.3) off: 1 - 4/(7*6)
    assert abs(score - (1.0 - 4.0 / (7.0 * 6.0))) < 2e-3 and score < 1.0  # reference :198-199


def test_unobserved_tsdf_scores_zero():
    z = np.zeros((20, 20), np.uint16)
    g = ob.OracleTsdf(z, z, 0.05, 0.3, 0.5, 0.3, 1.0)
    assert g.score_candidate(POINTS, 0, 0) == 0.0  # summed_weight == 0 (RT2D.cc:55)


def test_match_recovers_offset_on_tsdf():
    g = _fixture()
    opts = {"linear_search_window": 0.1, "angular_search_window": 0.1,
            "translation_delta_cost_weight": 0.0, "rotation_delta_cost_weight": 0.0}
    r = ob.rt_match_tsdf(g, opts, (0.05, -0.05, 0.0), POINTS)
    assert abs(r.pose_x) < 0.026 and abs(r.pose_y) < 0.026 and r.score > 0.95
