"""Oracle restatement of CeresScanMatcher2D against the reference's own known answers
(ceres_scan_matcher_2d_test.cc:30-98) plus derivative and sanity checks of the restated
solver.  Ceres itself is not in this image: parity with it is pinned at the reference
tests' tolerance only."""
import numpy as np
import pytest

import oracle_binding as ob
import synth

REF_OPTS = {"occupied_space_weight": 1.0, "translation_weight": 0.1, "rotation_weight": 1.5,
            "use_nonmonotonic_steps": True, "max_num_iterations": 50}


def is_nearly(pose, expected, epsilon):
    """transform::IsNearly (rigid_transform_test_helpers.h:42-46): Eigen's isApprox on the
    homogeneous 3x3 matrices, |a - b|_F <= epsilon * min(|a|_F, |b|_F)."""
    def mat(p):
        c, s = np.cos(p[2]), np.sin(p[2])
        return np.array([[c, -s, p[0]], [s, c, p[1]], [0.0, 0.0, 1.0]])
    a, b = mat(pose), mat(expected)
    return np.linalg.norm(a - b) <= epsilon * min(np.linalg.norm(a), np.linalg.norm(b))


def reference_fixture():
    # ProbabilityGrid(MapLimits(1., (10, 10), CellLimits(20, 20))), one cell at
    # GetCellIndex(-3.5, 2.5) set to kMaxProbability; cloud = {(-3, 2, 0)}  (:34-42)
    cells = np.zeros((20, 20), np.uint16)
    x_index = int(np.round((10.0 - 2.5) / 1.0 - 0.5))   # from the world y
    y_index = int(np.round((10.0 + 3.5) / 1.0 - 0.5))   # from the world x
    cells[y_index, x_index] = ob.lib().oracle_probability_to_cost_value(np.float32(0.9))
    return ob.OracleGrid(cells, 1.0, 10.0, 10.0), np.array([[-3.0, 2.0, 0.0]], np.float32)


@pytest.mark.parametrize("start", [(-0.5, 0.5), (-0.3, 0.5), (-0.45, 0.3), (-0.3, 0.3)])
def test_reference_known_answers(start):
    grid, cloud = reference_fixture()
    init = (start[0], start[1], 0.0)
    pose, summary = ob.ceres_match(grid, REF_OPTS, start, init, cloud)
    assert abs(summary["final_cost"]) < 1e-2                       # :72
    assert is_nearly(pose, (-0.5, 0.5, 0.0), 1e-2)                 # :73


def test_jacobian_matches_finite_differences():
    w = synth.World(200, 200, seed=77)
    grid = ob.OracleGrid(w.cells, w.res, w.max_x, w.max_y)
    pose0, pts = synth.make_scan(w, 5, max_range=6.0)
    pose = np.array([pose0[0] + 0.03, pose0[1] - 0.02, pose0[2] + 0.01])
    r, j = ob.ceres_evaluate(grid, None, pose[:2], pose[2], pose, pts)
    for k in range(3):
        h = 1e-4  # coarser than the 1.2e-7-cell rounding of "+ kPadding" (2^29) in the cell coordinate
        hi, lo = pose.copy(), pose.copy()
        hi[k] += h
        lo[k] -= h
        rh, _ = ob.ceres_evaluate(grid, None, pose[:2], pose[2], hi, pts)
        rl, _ = ob.ceres_evaluate(grid, None, pose[:2], pose[2], lo, pts)
        fd = (rh - rl) / (2 * h)
        assert np.allclose(fd, j[:, k], rtol=2e-3, atol=2e-4)


def test_refinement_improves_a_perturbed_pose():
    w = synth.World(300, 300, seed=78)
    grid = ob.OracleGrid(w.cells, w.res, w.max_x, w.max_y)
    opts = {"occupied_space_weight": 20.0, "translation_weight": 10.0, "rotation_weight": 1.0,
            "use_nonmonotonic_steps": True, "max_num_iterations": 10}  # pose_graph.lua
    better = 0
    for i in range(6):
        truth, pts = synth.make_scan(w, 100 + i, max_range=8.0)
        init = (truth[0] + 0.04, truth[1] - 0.03, truth[2] + 0.01)
        pose, summary = ob.ceres_match(grid, opts, init[:2], init, pts)
        assert summary["final_cost"] <= summary["initial_cost"]
        e0 = np.hypot(init[0] - truth[0], init[1] - truth[1])
        e1 = np.hypot(pose[0] - truth[0], pose[1] - truth[1])
        better += e1 < e0
    assert better >= 5
