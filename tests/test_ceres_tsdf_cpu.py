"""Oracle restatement of TSDFMatchCostFunction2D / InterpolatedTSDF2D against the reference's
known answers (tsdf_match_cost_function_2d_test.cc:33-150): a wall at y = 1 seen from
(-0.5, -0.5).  The reference builds the fixture with TSDFRangeDataInserter2D (projected
distance to the scan normal, weight 1); here the same field is painted directly: tsd = 1 - y
in the truncation band above and below the wall."""
import numpy as np

import oracle_binding as ob
import synth
import tsdf_util

UNIT = {"occupied_space_weight": 1.0, "translation_weight": 1.0, "rotation_weight": 1.0}


def wall_fixture(empty=False):
    # TSDF2D(MapLimits(0.1, (2.05, 2.05), CellLimits(40, 40)), 0.3, 1.0)  (:36-37)
    yi, xi = np.meshgrid(np.arange(40), np.arange(40), indexing="ij")
    cx = 2.05 - (yi + 0.5) * 0.1   # world x of the cell centre
    cy = 2.05 - (xi + 0.5) * 0.1   # world y
    tsd = (1.0 - cy).astype(np.float32)
    known = (np.abs(tsd) <= 0.3 + 1e-6) & (cx > -0.65) & (cx < 0.55) & (not empty)
    tv, wv = ob.tsdf_values(0.3, 1.0, tsd, np.ones_like(tsd))
    tv[~known] = 0
    wv[~known] = 0
    return ob.OracleTsdf(tv, wv, 0.1, 2.05, 2.05, 0.3, 1.0)


CLOUD = np.array([[0.0, 1.0, 0.0]], np.float32)


def test_match_empty_tsdf_is_invalid():
    ok, _, _ = ob.ceres_evaluate_tsdf(wall_fixture(empty=True), UNIT, (0, 0), 0.0, (0.0, 0.0, 0.0),
                                      np.zeros((1, 3), np.float32))
    assert not ok                                                     # :74


def test_exact_initial_pose():
    ok, r, j = ob.ceres_evaluate_tsdf(wall_fixture(), UNIT, (0, 0), 0.0, (0.0, 0.0, 0.0), CLOUD)
    assert ok and abs(r[0]) < 1e-3                                    # :91-92
    assert abs(j[0][0]) < 1e-3 and abs(j[0][1] + 1.0) < 1e-3 and abs(j[0][2]) < 1e-3   # :93-95


def test_perturbed_initial_pose():
    g = wall_fixture()
    for dy, want in ((0.1, -0.1), (-0.1, 0.1)):                       # :112-128
        ok, r, j = ob.ceres_evaluate_tsdf(g, UNIT, (0, 0), 0.0, (0.0, dy, 0.0), CLOUD)
        assert ok and abs(r[0] - want) < 1e-3
        assert abs(j[0][0]) < 1e-3 and abs(j[0][1] + 1.0) < 1e-3 and abs(j[0][2]) < 1e-3


def test_jacobian_matches_finite_differences_and_match_converges():
    w = synth.World(200, 200, seed=79)
    tv, wv = tsdf_util.from_world(w, 0.3, 10.0, seed=2)
    g = ob.OracleTsdf(tv, wv, w.res, w.max_x, w.max_y, 0.3, 10.0)
    truth, pts = synth.make_scan(w, 9, max_range=6.0)
    pose = np.array([truth[0] + 0.012, truth[1] - 0.017, truth[2] + 0.004])
    ok, r, j = ob.ceres_evaluate_tsdf(g, None, pose[:2], pose[2], pose, pts)
    assert ok
    for k in range(3):
        h = 1e-7  # the field is piecewise bilinear: stay inside the pixel
        hi, lo = pose.copy(), pose.copy()
        hi[k] += h
        lo[k] -= h
        _, rh, _ = ob.ceres_evaluate_tsdf(g, None, pose[:2], pose[2], hi, pts)
        _, rl, _ = ob.ceres_evaluate_tsdf(g, None, pose[:2], pose[2], lo, pts)
        fd = (rh - rl) / (2 * h)
        bad = np.abs(fd - j[:, k]) > 1e-3 + 1e-3 * np.abs(j[:, k])
        assert bad.mean() < 0.02  # points within 1e-7 of a pixel border see two pixels
    opts = {"occupied_space_weight": 20.0, "translation_weight": 10.0, "rotation_weight": 1.0,
            "use_nonmonotonic_steps": True, "max_num_iterations": 10}
    out, summary = ob.ceres_match_tsdf(g, opts, pose[:2], pose, pts)
    assert summary["final_cost"] <= summary["initial_cost"] and summary["termination"] != 5
