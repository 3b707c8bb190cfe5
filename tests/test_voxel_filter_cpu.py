"""The node-cloud filters of the host layer (host/cartographer/sensor/internal/voxel_filter.h)
against this fork's own sensor/internal/voxel_filter.cc compiled into oracle/_ref, and against the
reference's unit tests (sensor/internal/voxel_filter_test.cc).  Host only."""
import numpy as np
import pytest

import ref_binding as rb

needs_ref = pytest.mark.skipif(not rb.available(), reason="oracle/_ref is not built and /root/reference is absent")


def _scan_like(rng, n, spread, clusters=True):
    pts = rng.uniform(-spread, spread, (n, 3)).astype(np.float32)
    if clusters:  # dense clumps: many collisions per voxel, so the reservoir's draws matter
        centres = rng.uniform(-spread, spread, (8, 3))
        pick = rng.integers(0, 8, n // 2)
        pts[: n // 2] = (centres[pick] + rng.normal(0, 0.05, (n // 2, 3))).astype(np.float32)
    return pts


def test_returns_one_point_in_each_voxel(csm):
    # voxel_filter_test.cc:30-42: four points, two of them share the voxel of size 0.3 around (0, 0, 0)
    pts = np.array([[0, 0, 0], [0.1, -0.1, 0.1], [0.3, -0.1, 0], [0, 0, 0.1]], np.float32)
    out = csm.voxel_filter(pts, 0.3)
    assert len(out) == 2
    assert any((out == pts[2]).all(axis=1))  # the lone point of its voxel always survives
    assert all(any((p == q).all() for q in pts) for p in out)


def test_handles_large_coordinates(csm):
    # voxel_filter_test.cc:63-73
    pts = np.array([[100000.0, 0, 0], [100000.001, -0.0001, 0.0001], [100000.003, -0.0001, 0], [-200000.0, 0, 0]],
                   np.float32)
    out = csm.voxel_filter(pts, 0.01)
    assert len(out) == 2 and any((out == pts[3]).all(axis=1))


def test_many_points_in_one_voxel_leave_one(csm):
    # voxel_filter_test.cc:75-83 (IgnoresTime, positions only here)
    pts = np.tile(np.array([[-100.0, 0.3, 0.4]], np.float32), (100, 1))
    assert len(csm.voxel_filter(pts, 0.3)) == 1


def test_ignores_points_beyond_max_range_and_reaches_the_minimum(csm):
    # AdaptiveVoxelFilter (voxel_filter.cc:32-76,264-269): max_range crops first
    pts = np.array([[0, 0, 0], [1, 0, 0], [0, 50, 0]], np.float32)
    out = csm.adaptive_voxel_filter(pts, max_length=0.1, min_num_points=1, max_range=10.0)
    assert len(out) == 2 and (np.linalg.norm(out, axis=1) <= 10).all()
    rng = np.random.default_rng(0)
    cloud = _scan_like(rng, 3000, 5.0, clusters=False)
    out = csm.adaptive_voxel_filter(cloud, max_length=2.0, min_num_points=400, max_range=50.0)
    assert 400 <= len(out) < 3000  # coarse edge first, then refined until enough points remain
    assert len(csm.adaptive_voxel_filter(cloud[:300], 2.0, 400, 50.0)) == 300  # sparse enough: untouched


@needs_ref
@pytest.mark.parametrize("n,spread,resolution", [(0, 1.0, 0.1), (1, 1.0, 0.1), (2000, 3.0, 0.05), (4000, 10.0, 0.5),
                                                 (1500, 40.0, 0.45), (3000, 2.0, 1.5)])
def test_voxel_filter_keeps_the_points_the_reference_keeps(csm, n, spread, resolution):
    rng = np.random.default_rng(n + int(100 * resolution))
    pts = _scan_like(rng, n, spread)
    got, want = csm.voxel_filter(pts, resolution), rb.voxel_filter(0, pts, resolution)
    assert got.tobytes() == want.tobytes()
    if n > 1:
        assert len(got) < n  # collisions happened: the random choices were exercised


@needs_ref
@pytest.mark.parametrize("max_length,min_num_points,max_range", [(0.5, 200, 50.0), (0.9, 1000, 8.0), (4.0, 1500, 60.0),
                                                                 (0.05, 10, 3.0), (2.0, 100000, 60.0)])
def test_adaptive_voxel_filter_matches_the_reference(csm, max_length, min_num_points, max_range):
    rng = np.random.default_rng(int(10 * max_length) + min_num_points)
    pts = _scan_like(rng, 5000, 12.0)
    got = csm.adaptive_voxel_filter(pts, max_length, min_num_points, max_range)
    want = rb.voxel_filter(1, pts, max_length, min_num_points, max_range)
    assert got.tobytes() == want.tobytes()


@needs_ref
@pytest.mark.parametrize("rad,ratio", [(0.2, 0.5), (0.5, 0.3), (1.0, 0.9), (0.1, 0.0)])
def test_voxel_filter_edge_matches_the_reference(csm, rad, ratio):
    rng = np.random.default_rng(int(100 * rad))
    pts = _scan_like(rng, 3000, 4.0)
    got, want = csm.voxel_filter_edge(pts, rad, ratio), rb.voxel_filter(2, pts, rad, ratio)
    assert got.tobytes() == want.tobytes()
    if ratio > 0:
        assert 0 < len(got) < len(pts)


@needs_ref
def test_negative_coordinates_share_keys_like_the_reference(csm):
    # the voxel key packs sign-extended indices with wrapping adds: points on both sides of every axis
    rng = np.random.default_rng(5)
    pts = (rng.integers(-6, 6, (4000, 3)) * 0.25 + rng.normal(0, 0.01, (4000, 3))).astype(np.float32)
    assert csm.voxel_filter(pts, 0.25).tobytes() == rb.voxel_filter(0, pts, 0.25).tobytes()
