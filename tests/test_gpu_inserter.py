"""ProbabilityGridRangeDataInserter2D on the device against the CPU restatement: cells
bit-identical after every insertion, on the reference's fixture (range_data_inserter_2d_test.cc)
and on synthetic scans with misses, growth and repeated insertion."""
import numpy as np
import pytest

import oracle_binding as ob
import synth
from test_inserter_cpu import EXPECTED, ORIGIN, RETURNS, H, M, U

pytestmark = pytest.mark.gpu


def test_reference_fixture_on_device(csm, gpu_ctx):
    g = csm.Grid(np.zeros((5, 5), np.uint16), 1.0, 1.0, 5.0)
    og = ob.OracleGrid(np.zeros((5, 5), np.uint16), 1.0, 1.0, 5.0)
    kw = dict(hit_probability=0.7, miss_probability=0.4)
    cells, mx, my = csm.grid_insert(gpu_ctx, g, [ORIGIN + (0.0,)], [RETURNS], **kw)
    ob.grid_insert(og, ORIGIN, RETURNS, **kw)
    assert cells.shape == (5, 5) and (mx, my) == (1.0, 5.0)
    assert np.array_equal(cells, og.cells)
    for row in range(5):
        for column in range(5):
            assert (cells[column, row] == 0) == (EXPECTED[column][row] == U)
    for _ in range(50):  # ProbabilityProgression: repeated insertion stays identical
        cells, _, _ = csm.grid_insert(gpu_ctx, g, [ORIGIN + (0.0,)], [RETURNS], **kw)
        ob.grid_insert(og, ORIGIN, RETURNS, **kw)
    assert np.array_equal(cells, og.cells)


def _scans(seed, count, max_range):
    w = synth.World(240, 200, seed=seed)
    origins, returns, misses = [], [], []
    rng = np.random.default_rng(seed)
    for i in range(count):
        pose, pts = synth.make_scan(w, seed * 100 + i, max_range=max_range)
        c, s = np.cos(pose[2]), np.sin(pose[2])
        world = np.stack([pose[0] + c * pts[:, 0] - s * pts[:, 1], pose[1] + s * pts[:, 0] + c * pts[:, 1],
                          np.zeros(len(pts))], axis=1).astype(np.float32)
        ang = rng.uniform(-np.pi, np.pi, size=12)
        miss = np.stack([pose[0] + max_range * np.cos(ang), pose[1] + max_range * np.sin(ang),
                         np.zeros(12)], axis=1).astype(np.float32)
        origins.append((np.float32(pose[0]), np.float32(pose[1]), 0.0))
        returns.append(world)
        misses.append(miss)
    return origins, returns, misses


@pytest.mark.parametrize("resident", [False, True])
@pytest.mark.parametrize("free_space", [True, False])
def test_scans_bit_identical(csm, gpu_ctx, resident, free_space):
    origins, returns, misses = _scans(3, 10, 5.0)
    # a small grid off-centre: the inserter has to grow it
    start = np.zeros((40, 30), np.uint16)
    g = csm.Grid(start, 0.05, 0.5, 0.25)
    og = ob.OracleGrid(start, 0.05, 0.5, 0.25)
    kw = dict(hit_probability=0.55, miss_probability=0.49, insert_free_space=free_space)
    cells, mx, my = csm.grid_insert(gpu_ctx, g, origins, returns, misses, resident=resident, **kw)
    for o, r, m in zip(origins, returns, misses):
        ob.grid_insert(og, o[:2], r, m, **kw)
    assert cells.shape == og.cells.shape and (mx, my) == (og.max_x, og.max_y)
    assert np.array_equal(cells, og.cells)
    assert (cells > 0).sum() > 1000


def test_degenerate_rays(csm, gpu_ctx):
    # vertical, horizontal, single-cell and exact-corner rays; a return in the origin's cell
    start = np.zeros((64, 64), np.uint16)
    g = csm.Grid(start, 0.1, 3.2, 3.2)
    og = ob.OracleGrid(start, 0.1, 3.2, 3.2)
    origin = (0.05, 0.05, 0.0)
    rets = np.array([[0.05, 2.0, 0], [2.0, 0.05, 0], [0.06, 0.06, 0], [1.05, 1.05, 0], [-1.95, -1.95, 0],
                     [-2.0, 0.05, 0], [0.05, -2.0, 0], [1.0, -2.0, 0], [-3.0, 2.95, 0]], np.float32)
    cells, _, _ = csm.grid_insert(gpu_ctx, g, [origin], [rets])
    ob.grid_insert(og, origin[:2], rets)
    assert np.array_equal(cells, og.cells)
