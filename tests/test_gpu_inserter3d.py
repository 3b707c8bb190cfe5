"""RangeDataInserter3D on the device (csm_hgrid3d_insert) against the CPU restatement, which
tests/test_oracle_vs_ref3d.py holds to the reference's own range_data_inserter_3d.cc: the set of
non-default voxels, their values and grid_size() are identical after every insertion."""
import numpy as np
import pytest

import oracle_binding as ob
import synth3d

pytestmark = pytest.mark.gpu


def _scans(seed, count, max_range=5.0, n=400):
    w = synth3d.World3D(seed, size_xy=10.0)
    rng = np.random.default_rng(seed)
    origins, returns = [], []
    for _ in range(count):
        o = np.array([rng.uniform(-3, 3), rng.uniform(-3, 3), rng.uniform(0.2, 1.5)], np.float32)
        d = np.linalg.norm(w.points - o, axis=1)
        near = np.nonzero(d < max_range)[0]
        pick = rng.choice(near, min(n, len(near)), replace=False)
        origins.append(o)
        returns.append(np.ascontiguousarray(w.points[pick].astype(np.float32)))
    return origins, returns


def _oracle(resolution, xyz, values, origins, returns, **kw):
    g = ob.OracleGrid3D(resolution, xyz, values)
    for o, r in zip(origins, returns):
        g.insert(o, r, **kw)
    oxyz, oval = g.cells()
    return oxyz, oval, g.grid_size


@pytest.mark.parametrize("resident", [False, True])
@pytest.mark.parametrize("free", [0, 2, 5])
def test_scans_into_an_empty_grid(csm, gpu_ctx, resident, free):
    origins, returns = _scans(7, 6)
    kw = dict(hit_probability=0.55, miss_probability=0.49, num_free_space_voxels=free)
    got = csm.grid3d_insert(gpu_ctx, 0.1, np.zeros((0, 3), np.int32), np.zeros(0, np.uint16), origins, returns,
                            resident=resident, **kw)
    want = _oracle(0.1, None, None, origins, returns, **kw)
    assert len(want[1]) > 1000
    assert got[0].tobytes() == want[0].tobytes() and got[1].tobytes() == want[1].tobytes() and got[2] == want[2]


def test_repeated_insertion_into_a_filled_grid_that_has_to_grow(csm, gpu_ctx):
    w = synth3d.World3D(3, size_xy=6.0)
    xyz, val = w.voxels(0.2)
    origins, returns = _scans(11, 4, max_range=8.0)
    far = np.array([[40.0, -35.0, 3.0], [-60.0, 20.0, -4.0]], np.float32)  # far outside the box: it has to grow
    origins.append(np.array([0.5, 0.5, 0.5], np.float32))
    returns.append(far)
    kw = dict(hit_probability=0.7, miss_probability=0.4, num_free_space_voxels=3)
    want = _oracle(0.2, xyz, val, origins * 3, returns * 3, **kw)
    for resident in (False, True):
        got = csm.grid3d_insert(gpu_ctx, 0.2, xyz, val, origins * 3, returns * 3, resident=resident, **kw)
        assert got[0].tobytes() == want[0].tobytes() and got[1].tobytes() == want[1].tobytes()
        assert got[2] == want[2] and want[2] > 128  # the reference's grid doubled


def test_edge_cases(csm, gpu_ctx):
    kw = dict(hit_probability=0.6, miss_probability=0.45, num_free_space_voxels=4)
    o = np.array([0.04, 0.0, 0.0], np.float32)
    cases = [
        np.array([[0.04, 0.01, 0.0]], np.float32),                      # hit in the origin's voxel: no samples
        np.array([[0.35, 0.0, 0.0], [0.35, 0.0, 0.0]], np.float32),     # duplicate hit: applied once
        np.array([[0.35, 0.0, 0.0], [0.65, 0.0, 0.0]], np.float32),     # a ray through another hit: hit wins
        np.array([[-0.25, -0.25, -0.25], [0.15, -0.45, 0.05]], np.float32),  # negative indices, halfway rounding
        np.zeros((0, 3), np.float32),                                   # no returns
    ]
    for returns in cases:
        want = _oracle(0.1, None, None, [o], [returns], **kw)
        got = csm.grid3d_insert(gpu_ctx, 0.1, np.zeros((0, 3), np.int32), np.zeros(0, np.uint16), [o], [returns],
                                resident=True, **kw)
        assert got[0].tobytes() == want[0].tobytes() and got[1].tobytes() == want[1].tobytes()


@pytest.mark.parametrize("depth,frd", [(5, 3), (6, 1), (3, 5)])
def test_stack_straight_from_the_device_grid(csm, gpu_ctx, depth, frd):
    """csm_stack3d_build_from_hgrid: the precomputation stack built from the device-resident grid
    (voxels never leave HBM) equals the one built from the exported voxel list, box by box and
    byte by byte; the latter is what test_gpu_parity3d holds to the oracle."""
    w = synth3d.World3D(13, size_xy=6.0)
    xyz, val = w.voxels(0.1)
    origins, returns = _scans(13, 3, max_range=4.0, n=300)
    assert csm.grid3d_resident_stack_check(gpu_ctx, 0.1, xyz, val, origins, returns, depth, frd) == 0
    assert csm.grid3d_resident_stack_check(gpu_ctx, 0.1, np.zeros((0, 3), np.int32), np.zeros(0, np.uint16),
                                           origins, returns, depth, frd) == 0
