"""TSDFRangeDataInserter2D on the device (csm_tsdfgrid2d_insert) against the CPU restatement, which
tests/test_oracle_vs_ref.py holds to the reference's own tsdf_range_data_inserter_2d.cc and
normal_estimation_2d.cc: TSD cells, weight cells and limits bit-identical after the scans, on the
reference-signature path (one Insert per scan) and on the device-resident one."""
import numpy as np
import pytest

import oracle_binding as ob
import synth

pytestmark = pytest.mark.gpu

OPTION_SETS = [
    {},
    {"update_free_space": True},
    {"project_sdf_distance_to_scan_normal": False, "update_weight_angle_scan_normal_to_ray_kernel_bandwidth": 0.0},
    {"update_weight_range_exponent": 2, "update_weight_distance_cell_to_hit_kernel_bandwidth": 0.0},
    {"truncation_distance": 0.15, "maximum_weight": 3.0, "num_normal_samples": 8, "sample_radius": 0.2},
]


def _scans(seed, count):
    w = synth.World(200, 180, seed=seed)
    origins, returns = [], []
    for k in range(count):
        pose, pts = synth.make_scan(w, seed * 10 + k, max_range=6.0)
        c, s = np.cos(pose[2]), np.sin(pose[2])
        returns.append(np.stack([pose[0] + c * pts[:, 0] - s * pts[:, 1], pose[1] + s * pts[:, 0] + c * pts[:, 1],
                                 np.zeros(len(pts))], axis=1).astype(np.float32))
        origins.append(np.array([pose[0], pose[1]], np.float32))
    return origins, returns


@pytest.mark.parametrize("resident", [False, True])
@pytest.mark.parametrize("opts", OPTION_SETS)
def test_scans_into_a_small_grid_that_has_to_grow(csm, gpu_ctx, opts, resident):
    origins, returns = _scans(33, 5)
    trunc, maxw = opts.get("truncation_distance", 0.3), opts.get("maximum_weight", 10.0)
    zero = np.zeros((40, 40), np.uint16)
    og = ob.OracleTsdf(zero, zero, 0.05, 1.0, 1.0, trunc, maxw)
    for o, r in zip(origins, returns):
        ob.tsdf_insert(og, o, r, opts)
    g = csm.TsdfGrid(zero, zero, 0.05, 1.0, 1.0, trunc, maxw)
    tsd, weight, (max_x, max_y) = csm.tsdf_insert(gpu_ctx, g, origins, returns, ob.tsdf_inserter_options(opts),
                                                  resident=resident)
    assert (g.nx, g.ny, max_x, max_y) == (og.nx, og.ny, og.max_x, og.max_y)
    assert tsd.tobytes() == og.tsd_cells.tobytes()
    assert weight.tobytes() == og.weight_cells.tobytes()
    assert (weight > 0).sum() > 1500


def test_repeated_insertion_converges_like_the_oracle(csm, gpu_ctx):
    """The same wall seen 30 times: weights saturate at maximum_weight, TSD values settle."""
    origins, returns = _scans(41, 1)
    zero = np.zeros((64, 64), np.uint16)
    og = ob.OracleTsdf(zero, zero, 0.05, 1.6, 1.6, 0.3, 10.0)
    g = csm.TsdfGrid(zero, zero, 0.05, 1.6, 1.6, 0.3, 10.0)
    for _ in range(30):
        ob.tsdf_insert(og, origins[0], returns[0])
    tsd, weight, _ = csm.tsdf_insert(gpu_ctx, g, origins * 30, returns * 30, ob.tsdf_inserter_options(), resident=True)
    assert tsd.tobytes() == og.tsd_cells.tobytes() and weight.tobytes() == og.weight_cells.tobytes()
    assert weight.max() == 32767  # WeightToValue(maximum_weight)


def test_matcher_on_a_tsdf_built_on_the_device(csm, gpu_ctx):
    """The inserter feeds the matchers: a TSDF2D built by device insertions scores a scan of the same
    world like the oracle's TSDF built by the oracle's inserter."""
    origins, returns = _scans(33, 5)
    zero = np.zeros((40, 40), np.uint16)
    og = ob.OracleTsdf(zero, zero, 0.05, 1.0, 1.0, 0.3, 10.0)
    for o, r in zip(origins, returns):
        ob.tsdf_insert(og, o, r)
    g = csm.TsdfGrid(zero, zero, 0.05, 1.0, 1.0, 0.3, 10.0)
    csm.tsdf_insert(gpu_ctx, g, origins, returns, ob.tsdf_inserter_options(), resident=True)
    w = synth.World(200, 180, seed=33)
    pose, pts = synth.make_scan(w, 335, max_range=6.0)
    opts = {"linear_search_window": 0.1, "angular_search_window": float(np.deg2rad(10.0)),
            "translation_delta_cost_weight": 0.1, "rotation_delta_cost_weight": 0.1}
    init = (pose[0] + 0.03, pose[1] - 0.02, pose[2] + 0.02)
    r_g = csm.rt_match_batch(gpu_ctx, [g], [pts], [init], **opts)[0]
    og2 = ob.OracleTsdf(og.tsd_cells, og.weight_cells, 0.05, og.max_x, og.max_y, 0.3, 10.0)
    r = ob.OracleResult2D()
    ob.lib().oracle_rt_match2d_tsdf(og2.h, opts["linear_search_window"], opts["angular_search_window"], 0.1, 0.1,
                                    init[0], init[1], init[2], np.ascontiguousarray(pts, np.float32), len(pts),
                                    ob.C.byref(r))
    assert np.float32(r_g.score).tobytes() == np.float32(r.score).tobytes()
    assert (r_g.scan_index, r_g.x_index_offset, r_g.y_index_offset) == (r.scan_index, r.x_index_offset, r.y_index_offset)
