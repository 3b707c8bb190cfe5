"""The CUDA 3D path against the committed golden vectors (no oracle involved)."""
import numpy as np
import pytest

import golden_util as gu
from test_golden3d_cpu import RES, check_result

pytestmark = pytest.mark.gpu


def _matcher(csm, ctx, arr, grid, low_grid, hist, options):
    return csm.Matcher3D(ctx, RES[grid], arr[grid + "_xyz"], arr[grid + "_val"], RES[low_grid],
                         arr[low_grid + "_xyz"], arr[low_grid + "_val"], hist, np.array(options, np.float64))


def test_gpu_reproduces_golden_stacks3d(csm, gpu_ctx):
    doc, arr = gu.load3d()
    for case in doc["stacks"]:
        m = _matcher(csm, gpu_ctx, arr, case["grid"], case["grid"], np.zeros(4, np.float32),
                     [case["depth"], case["frd"], 0, 0, 1, 1, 0.1])
        assert m.grid_size == case["grid_size"]
        for i, lv in enumerate(case["levels"]):
            lo, dense = gu.canonical_level(*m.level(i))
            assert lo.tolist() == lv["lo"] and list(dense.shape) == lv["shape"] and gu.sha(dense) == lv["sha1"]


def test_gpu_reproduces_golden_matches3d(csm, gpu_ctx):
    doc, arr = gu.load3d()
    matchers, nodes, node_poses, submap_poses, full, min_scores = [], [], [], [], [], []
    for case in doc["matches"]:
        matchers.append(_matcher(csm, gpu_ctx, arr, case["grid"], case["low_grid"], arr[case["submap_hist"]],
                                 case["options"]))
        nodes.append(csm.Node3D(arr[case["hi"]], arr[case["lo"]], arr[case["hist"]], case["gravity"]))
        node_poses.append(case["node_pose"])
        submap_poses.append(case["submap_pose"])
        full.append(case["full"])
        min_scores.append(case["min_score"])
    res = csm.match3d_batch(gpu_ctx, matchers, nodes, min_scores, np.array(node_poses), np.array(submap_poses),
                            full)
    for case, r in zip(doc["matches"], res):
        # the device reports a float pose widened to double, like the reference
        r.pose[:] = [float(np.float32(v)) for v in r.pose]
        check_result(r, case["expect"])


def test_gpu_reproduces_golden_discrete_scans3d(csm, gpu_ctx):
    doc, arr = gu.load3d()
    for d in doc["discrete"]:
        case = doc["matches"][d["match"]]
        m = _matcher(csm, gpu_ctx, arr, case["grid"], case["low_grid"], arr[case["submap_hist"]], case["options"])
        node = csm.Node3D(arr[case["hi"]], arr[case["lo"]], arr[case["hist"]], case["gravity"])
        poses, rot, cells = csm.discrete_scans3d(gpu_ctx, m, node, np.array(case["node_pose"]),
                                                 np.array(case["submap_pose"]))
        assert len(rot) == d["num_scans"] and cells[0, :3].tolist() == d["cells_first"]
        assert (gu.sha(poses), gu.sha(rot), gu.sha(cells)) == (d["poses_sha1"], d["rot_sha1"], d["cells_sha1"])
