"""csm_discretize2d_noalloc (no counterpart in the
reference): the variant of csm_discretize2d whose scratch is a fixed device array inside the library,
for callers that must not allocate.  Called through the C ABI, twice in a row on different inputs
(the second smaller, so stale contents of the shared array would show), and compared bit for bit
with the existing entry point -- integer cell indices and window bounds, so the tolerance is zero.
The file name keeps it last in pytest's collection order."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


class Job2D(C.Structure):  # csm_job2d of include/csm_b200.h
    _fields_ = [("stack", C.c_void_p), ("points_xyz", C.c_void_p), ("num_points", C.c_int32),
                ("num_scans", C.c_int32), ("scan_quat", C.c_void_p), ("init_quat_w", C.c_float),
                ("init_quat_s", C.c_float), ("init_tx", C.c_float), ("init_ty", C.c_float),
                ("max_x", C.c_double), ("max_y", C.c_double), ("resolution", C.c_double),
                ("linear_bound", C.c_int32), ("min_score", C.c_float), ("score_offset", C.c_float),
                ("score_scale", C.c_float)]


def _job(seed, n, num_scans, yaw, tx, ty):
    rng = np.random.default_rng(seed)
    pts = np.ascontiguousarray(np.c_[rng.uniform(-6, 6, (n, 2)), np.zeros(n)], dtype=np.float32)
    half = (0.5 * np.linspace(-1.0, 1.0, num_scans)).astype(np.float32)
    quat = np.ascontiguousarray(np.stack([np.cos(half), np.sin(half)], 1), dtype=np.float32)
    job = Job2D()
    job.points_xyz, job.num_points = pts.ctypes.data, n
    job.scan_quat, job.num_scans = quat.ctypes.data, num_scans
    job.init_quat_w, job.init_quat_s = float(np.cos(np.float32(yaw / 2))), float(np.sin(np.float32(yaw / 2)))
    job.init_tx, job.init_ty = tx, ty
    job.max_x, job.max_y, job.resolution, job.linear_bound = 10.0, 10.0, 0.05, 140
    return job, (pts, quat)


def _run(abi, fn, raw, job, nx, ny, shrink):
    xy = np.full((job.num_scans, job.num_points, 2), -12345, np.int32)
    bounds = np.full((job.num_scans, 4), -12345, np.int32)
    rc = fn(raw, C.byref(job), nx, ny, shrink, xy.ctypes.data_as(C.c_void_p), bounds.ctypes.data_as(C.c_void_p))
    assert rc == 0, abi.csm_last_error()
    return xy, bounds


def test_no_alloc_discretize_matches_the_allocating_entry_point(csm, gpu_ctx):
    abi = csm.abi()
    raw = C.c_void_p(gpu_ctx.raw)
    for fn in (abi.csm_discretize2d, abi.csm_discretize2d_noalloc):
        fn.argtypes = [C.c_void_p, C.POINTER(Job2D), C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        fn.restype = C.c_int
    cases = [(1, 360, 1259, 0.3, 1.5, -2.0, 1), (2, 97, 31, -1.1, -3.0, 0.25, 0), (3, 1, 1, 0.0, 0.0, 0.0, 1)]
    for seed, n, scans, yaw, tx, ty, shrink in cases:  # consecutive calls share the scratch array
        job, keep = _job(seed, n, scans, yaw, tx, ty)
        got_xy, got_b = _run(abi, abi.csm_discretize2d_noalloc, raw, job, 400, 400, shrink)
        want_xy, want_b = _run(abi, abi.csm_discretize2d, raw, job, 400, 400, shrink)
        assert np.array_equal(got_xy, want_xy) and np.array_equal(got_b, want_b)
        assert (got_xy != -12345).all() and (got_b != -12345).all()
        del keep
    # the same input again after the smaller ones: nothing stale survives
    job, keep = _job(1, 360, 1259, 0.3, 1.5, -2.0)
    again_xy, again_b = _run(abi, abi.csm_discretize2d_noalloc, raw, job, 400, 400, 1)
    want_xy, want_b = _run(abi, abi.csm_discretize2d, raw, job, 400, 400, 1)
    assert np.array_equal(again_xy, want_xy) and np.array_equal(again_b, want_b)


def test_no_alloc_discretize_rejects_sizes_beyond_its_scratch(csm, gpu_ctx):
    abi = csm.abi()
    fn = abi.csm_discretize2d_noalloc
    fn.argtypes = [C.c_void_p, C.POINTER(Job2D), C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    fn.restype = C.c_int
    for n, scans in ((4097, 4), (8, 4097), (1024, 1024)):
        job, keep = _job(5, n, scans, 0.0, 0.0, 0.0)
        xy = np.zeros((scans, n, 2), np.int32)
        bounds = np.zeros((scans, 4), np.int32)
        rc = fn(C.c_void_p(gpu_ctx.raw), C.byref(job), 400, 400, 1, xy.ctypes.data_as(C.c_void_p),
                bounds.ctypes.data_as(C.c_void_p))
        assert rc == -1  # CSM_ERR_ARG
        assert b"exceed the fixed scratch" in abi.csm_last_error()
