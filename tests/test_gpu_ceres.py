"""CeresScanMatcher2D on the device (csm_ceres_match2d_batch through the host class) against
the reference's known answers and against the CPU restatement.  Tolerances: the reference's
own 1e-2 (IsNearly) for its fixture; 1e-6 on the pose and 1e-9 relative on the costs against
the oracle (the device solves the damped normal equations, the oracle the QR of [J; D], and
the device sums the residual rows in another order: agreement is to rounding, not bitwise)."""
import numpy as np
import pytest

import oracle_binding as ob
import synth
from test_ceres_cpu import REF_OPTS, is_nearly, reference_fixture

pytestmark = pytest.mark.gpu

POSE_GRAPH = {"occupied_space_weight": 20.0, "translation_weight": 10.0, "rotation_weight": 1.0,
              "use_nonmonotonic_steps": True, "max_num_iterations": 10}    # pose_graph.lua
LOCAL_SLAM = {"occupied_space_weight": 1.0, "translation_weight": 10.0, "rotation_weight": 40.0,
              "use_nonmonotonic_steps": False, "max_num_iterations": 20}   # trajectory_builder_2d.lua


@pytest.mark.parametrize("start", [(-0.5, 0.5), (-0.3, 0.5), (-0.45, 0.3), (-0.3, 0.3)])
def test_reference_known_answers_on_device(csm, gpu_ctx, start):
    og, cloud = reference_fixture()
    g = csm.Grid(og.cells, 1.0, 10.0, 10.0)
    poses, summ = csm.ceres_match_batch(gpu_ctx, [g], [cloud], [start], [(start[0], start[1], 0.0)],
                                        **REF_OPTS)
    assert abs(summ[0][1]) < 1e-2                      # ceres_scan_matcher_2d_test.cc:72
    assert is_nearly(poses[0], (-0.5, 0.5, 0.0), 1e-2)  # :73
    want, ws = ob.ceres_match(og, REF_OPTS, start, (start[0], start[1], 0.0), cloud)
    assert np.allclose(poses[0], want, atol=1e-6)
    assert int(summ[0][2]) == ws["iterations"]


@pytest.mark.parametrize("opts", [POSE_GRAPH, LOCAL_SLAM])
def test_batch_matches_oracle(csm, gpu_ctx, opts):
    w = synth.World(300, 300, seed=91)
    og = ob.OracleGrid(w.cells, w.res, w.max_x, w.max_y)
    g = csm.Grid(w.cells, w.res, w.max_x, w.max_y)
    rng = np.random.default_rng(4)
    clouds, inits, targets = [], [], []
    for i in range(24):
        truth, pts = synth.make_scan(w, 700 + i, max_range=8.0)
        init = (truth[0] + rng.uniform(-0.06, 0.06), truth[1] + rng.uniform(-0.06, 0.06),
                truth[2] + rng.uniform(-0.03, 0.03))
        if i == 20:
            init = (truth[0] + 9.0, truth[1] + 9.0, truth[2])  # mostly outside the grid
        if i == 21:
            pts = pts[:1]                                      # single point
        clouds.append(np.ascontiguousarray(pts))
        inits.append(init)
        # the constraint builder passes the pose itself as the target (CB2D.cc:245-249);
        # local SLAM passes the extrapolated pose: shift it a little
        targets.append((init[0], init[1]) if opts is POSE_GRAPH else (init[0] + 0.01, init[1] - 0.01))
    poses, summ = csm.ceres_match_batch(gpu_ctx, [g] * len(clouds), clouds, targets, inits, **opts)
    same_iterations = 0
    for i in range(len(clouds)):
        want, ws = ob.ceres_match(og, opts, targets[i], inits[i], clouds[i])
        assert np.allclose(poses[i], want, atol=1e-6), (i, poses[i], want)
        assert abs(summ[i][0] - ws["initial_cost"]) <= 1e-9 * max(1.0, ws["initial_cost"])
        assert abs(summ[i][1] - ws["final_cost"]) <= 1e-7 * max(1.0, ws["final_cost"])
        assert summ[i][1] <= summ[i][0]
        same_iterations += int(summ[i][2]) == ws["iterations"]
    assert same_iterations >= len(clouds) - 2
