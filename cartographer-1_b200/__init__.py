"""ctypes view of the B200 correlative scan matcher (test / bench driver only).

The product is native: `libcsm_b200.so` (C ABI of include/csm_b200.h, CUDA kernels
for sm_100a) and `libcsm_host.so` (C++ host layer with the reference's class
interfaces).  This module only loads them; it contains no algorithmic code and no
fallback: importing it without the built libraries raises, and creating a context
without a CUDA device raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_ABI = os.path.join(_HERE, "libcsm_b200.so")
LIB_HOST = os.path.join(_HERE, "libcsm_host.so")


class Result2D(C.Structure):
    _fields_ = [
        ("found", C.c_int32),
        ("scan_index", C.c_int32),
        ("x_index_offset", C.c_int32),
        ("y_index_offset", C.c_int32),
        ("score", C.c_float),
        ("tie_resolution", C.c_int32),
        ("pose_x", C.c_double),
        ("pose_y", C.c_double),
        ("pose_theta", C.c_double),
        ("candidates_scored", C.c_uint64),
    ]


class Result3D(C.Structure):
    _fields_ = [
        ("found", C.c_int32),
        ("scan_index", C.c_int32),
        ("offset", C.c_int32 * 3),
        ("num_scans", C.c_int32),
        ("score", C.c_float),
        ("rotational_score", C.c_float),
        ("low_resolution_score", C.c_float),
        ("tie_resolution", C.c_int32),
        ("pose", C.c_double * 7),
        ("candidates_scored", C.c_uint64),
    ]


# Symbols include/csm_b200.h declares; tests check every one is exported.
ABI_SYMBOLS = [
    "csm_stack2d_build_many", "csm_host_register", "csm_host_unregister",
    "csm_last_error", "csm_device_count", "csm_ctx_create", "csm_ctx_destroy",
    "csm_ctx_synchronize", "csm_ctx_stream", "csm_ctx_set_profiling", "csm_ctx_get_profile",
    "csm_ctx_get_tile_histogram", "csm_stack2d_build", "csm_stack2d_build_dev",
    "csm_stack2d_rebuild_dev", "csm_stack2d_cells_dev", "csm_ctx_set_keep_cells", "csm_stack2d_level", "csm_stack2d_quad", "csm_stack2d_depth",
    "csm_stack2d_device_bytes", "csm_stack2d_destroy", "csm_match2d_batch",
    "csm_batch2d_upload", "csm_batch2d_run", "csm_batch2d_fetch", "csm_batch2d_destroy",
    "csm_discretize2d", "csm_rt_match2d_batch", "csm_ctx_launch_count", "csm_measure_gather_rate",
    "csm_stack3d_build", "csm_stack3d_level", "csm_stack3d_depth", "csm_stack3d_device_bytes",
    "csm_stack3d_destroy", "csm_grid3d_build", "csm_grid3d_destroy", "csm_match3d_batch",
    "csm_batch3d_upload", "csm_batch3d_run", "csm_batch3d_fetch", "csm_batch3d_destroy",
    "csm_discretize3d", "csm_rt_match3d_batch", "csm_ceres_match2d_batch", "csm_grid2d_create", "csm_grid2d_insert",
    "csm_grid2d_download", "csm_grid2d_cells_dev", "csm_grid2d_destroy",
    "csm_grid2d_broadcast_dev", "csm_stack2d_broadcast",
    "csm_tsdfgrid2d_create", "csm_tsdfgrid2d_insert", "csm_tsdfgrid2d_download", "csm_tsdfgrid2d_tsd_dev",
    "csm_tsdfgrid2d_weight_dev", "csm_tsdfgrid2d_destroy",
    "csm_hgrid3d_create", "csm_hgrid3d_insert", "csm_hgrid3d_export", "csm_hgrid3d_box", "csm_hgrid3d_destroy",
    "csm_stack3d_build_from_hgrid",
    "csm_measure_gather_ceiling", "csm_ctx_get_depth_profile", "csm_precomputation_grid2d",
    "csm_rotate_scans2d", "csm_discretize_points2d", "csm_rt_score_candidates2d",
    "csm_discretize2d_noalloc",
]

_abi = None
_host = None


def abi():
    global _abi
    if _abi is None:
        if not os.path.exists(LIB_ABI):
            raise RuntimeError("libcsm_b200.so is not built (run __graft_entry__.build()); "
                               "there is no CPU fallback")
        _abi = C.CDLL(LIB_ABI, mode=C.RTLD_GLOBAL)
        _abi.csm_last_error.restype = C.c_char_p
        _abi.csm_ctx_stream.restype = C.c_void_p
        _abi.csm_ctx_stream.argtypes = [C.c_void_p]
        _abi.csm_stack2d_device_bytes.restype = C.c_size_t
        _abi.csm_ctx_launch_count.restype = C.c_ulonglong
        _abi.csm_ctx_launch_count.argtypes = [C.c_void_p]
        _abi.csm_measure_gather_rate.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_int,
                                                 C.c_uint, C.c_int, C.POINTER(C.c_double)]
        _abi.csm_measure_gather_ceiling.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_uint, C.c_int, C.c_int,
                                                    C.POINTER(C.c_double)]
        _abi.csm_ctx_get_depth_profile.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
    return _abi


def host():
    global _host
    if _host is not None:
        return _host
    abi()
    if not os.path.exists(LIB_HOST):
        raise RuntimeError("libcsm_host.so is not built (run __graft_entry__.build())")
    L = C.CDLL(LIB_HOST)
    u16p = np.ctypeslib.ndpointer(np.uint16, flags="C_CONTIGUOUS")
    f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
    f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
    i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
    i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
    u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")
    L.csmh_context_create.restype = C.c_void_p
    L.csmh_context_create.argtypes = [C.c_int]
    L.csmh_context_destroy.argtypes = [C.c_void_p]
    L.csmh_context_raw.restype = C.c_void_p
    L.csmh_context_raw.argtypes = [C.c_void_p]
    L.csmh_grid_create.restype = C.c_void_p
    L.csmh_grid_create.argtypes = [u16p, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double]
    L.csmh_grid_destroy.argtypes = [C.c_void_p]
    L.csmh_tsdf_create.restype = C.c_void_p
    L.csmh_tsdf_create.argtypes = [u16p, u16p, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double,
                                   C.c_float, C.c_float]
    L.csmh_quantization_table.argtypes = [C.c_void_p, u8p]
    L.csmh_matcher_create.restype = C.c_void_p
    L.csmh_matcher_create.argtypes = [C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_int]
    L.csmh_matcher_destroy.argtypes = [C.c_void_p]
    L.csmh_matcher_stack.restype = C.c_void_p
    L.csmh_matcher_stack.argtypes = [C.c_void_p]
    L.csmh_matcher_level.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                     C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.csmh_match_batch.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), i32p, f64p, f32p,
                                   i64p, f32p, C.POINTER(Result2D)]
    L.csmh_batch_create.restype = C.c_void_p
    L.csmh_batch_create.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), i32p, f64p, f32p,
                                    i64p, f32p]
    L.csmh_batch_run.argtypes = [C.c_void_p]
    L.csmh_batch_fetch.argtypes = [C.c_void_p, C.POINTER(Result2D)]
    L.csmh_batch_destroy.argtypes = [C.c_void_p]
    L.csmh_matcher_rebuild_dev.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.csmh_rt_match_batch.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), f64p, f64p, f32p,
                                      i64p, C.POINTER(Result2D)]
    L.csmh_discretize.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_int, C.c_int,
                                  f32p, C.c_int, C.c_double, C.c_double, C.c_double, C.c_double,
                                  C.c_double, C.c_int, C.POINTER(C.c_int), C.c_void_p, C.c_void_p]
    L.csmh_constraint_builder_run.restype = C.c_double
    L.csmh_constraint_builder_run.argtypes = [
        C.c_int, i32p, C.c_int, C.c_int, C.c_int, f64p, C.c_int, C.POINTER(C.c_void_p), f64p,
        C.c_int, i32p, i32p, f64p, f32p, i64p, i32p, f64p, f64p]
    L.csmh_constraint_builder_run2.restype = C.c_double
    L.csmh_constraint_builder_run2.argtypes = L.csmh_constraint_builder_run.argtypes + [C.c_int, C.c_int, f64p]
    L.csmh_fcsm3d_create.restype = C.c_void_p
    L.csmh_fcsm3d_create.argtypes = [C.c_void_p, C.c_float, i32p, u16p, C.c_int, C.c_float, i32p, u16p,
                                     C.c_int, f32p, C.c_int, f64p]
    L.csmh_fcsm3d_destroy.argtypes = [C.c_void_p]
    L.csmh_fcsm3d_grid_size.argtypes = [C.c_void_p]
    L.csmh_fcsm3d_level.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, i32p, i32p]
    L.csmh_node3d_create.restype = C.c_void_p
    L.csmh_node3d_create.argtypes = [f32p, C.c_int, f32p, C.c_int, f32p, C.c_int, f64p]
    L.csmh_node3d_destroy.argtypes = [C.c_void_p]
    L.csmh_batch3d_create.restype = C.c_void_p
    L.csmh_batch3d_create.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                      i32p, f64p, f64p, f32p]
    L.csmh_batch3d_run.argtypes = [C.c_void_p]
    L.csmh_batch3d_fetch.argtypes = [C.c_void_p, C.POINTER(Result3D)]
    L.csmh_batch3d_destroy.argtypes = [C.c_void_p]
    L.csmh_fcsm3d_match.argtypes = [C.c_void_p, C.c_void_p, C.c_int, f64p, f64p, C.c_float,
                                    C.POINTER(Result3D)]
    L.csmh_constraint_builder3d_run.restype = C.c_double
    L.csmh_constraint_builder3d_run.argtypes = [
        C.c_int, i32p, C.c_int, C.c_int, f64p, f64p, C.c_int, C.POINTER(C.c_void_p), C.c_int, i32p, i32p,
        C.POINTER(C.c_void_p), f64p, f64p, i32p, f64p, f64p]
    L.csmh_rt_match3d.restype = C.c_float
    L.csmh_rt_match3d.argtypes = [C.c_float, i32p, u16p, C.c_int, f64p, f64p, f32p, C.c_int, f64p]
    L.csmh_rt_match3d_batch.argtypes = [C.c_void_p, C.c_float, i32p, u16p, C.c_int, f64p, C.c_int, f64p, f32p,
                                        i64p, f32p, f64p, i32p, i32p]
    L.csmh_rotational_match.argtypes = [f32p, f32p, C.c_int, C.c_float, f32p, C.c_int, f32p]
    L.csmh_fcsm3d_discrete_scans.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, f64p, f64p, C.c_void_p,
                                             C.c_void_p, C.c_void_p]
    _host = L
    return L


class Context:
    def __init__(self, device=0):
        self.h = host().csmh_context_create(device)
        if not self.h:
            raise RuntimeError("no CUDA device %d: the scan matcher has no CPU fallback" % device)
        self.raw = host().csmh_context_raw(self.h)

    def close(self):
        if self.h:
            host().csmh_context_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()


class Grid:
    """ProbabilityGrid holder: cells uint16 [ny, nx]."""

    def __init__(self, cells, res, max_x, max_y):
        cells = np.ascontiguousarray(cells, dtype=np.uint16)
        self.ny, self.nx = cells.shape
        self.res, self.max_x, self.max_y = res, max_x, max_y
        self.h = host().csmh_grid_create(cells, self.nx, self.ny, res, max_x, max_y)

    def quantization_table(self):
        out = np.empty(65536, dtype=np.uint8)
        host().csmh_quantization_table(self.h, out)
        return out

    def pin(self):
        """Page-locks the cells where they are: Matcher.create_many then copies them to the device
        with an asynchronous transfer instead of through a staging block."""
        L = host()
        L.csmh_grid_pin.argtypes = [C.c_void_p]
        rc = L.csmh_grid_pin(self.h)
        if rc != 0:
            raise RuntimeError(abi().csm_last_error().decode())
        return self

    def __del__(self):
        if getattr(self, "h", None):
            host().csmh_grid_destroy(self.h)
            self.h = None


class TsdfGrid:
    """A TSDF2D (truncated signed distance + weight cells); RealTimeCorrelativeScanMatcher2D
    scores it with the reference's weighted TSD score (RT2D.cc:38-59)."""

    def __init__(self, tsd_cells, weight_cells, res, max_x, max_y, truncation_distance, max_weight):
        tsd_cells = np.ascontiguousarray(tsd_cells, dtype=np.uint16)
        weight_cells = np.ascontiguousarray(weight_cells, dtype=np.uint16)
        assert tsd_cells.shape == weight_cells.shape
        self.ny, self.nx = tsd_cells.shape
        self.h = host().csmh_tsdf_create(tsd_cells, weight_cells, self.nx, self.ny, res, max_x, max_y,
                                         truncation_distance, max_weight)

    def __del__(self):
        if getattr(self, "h", None):
            host().csmh_grid_destroy(self.h)
            self.h = None


class Matcher:
    """FastCorrelativeScanMatcher2D on the device."""

    def __init__(self, ctx, grid, linear_window, angular_window, depth):
        self.ctx, self.grid, self.depth = ctx, grid, depth
        self.h = host().csmh_matcher_create(ctx.h, grid.h, linear_window, angular_window, depth)

    def level(self, i):
        wx, wy = C.c_int(), C.c_int()
        rc = host().csmh_matcher_level(self.ctx.h, self.h, i, None, C.byref(wx), C.byref(wy))
        assert rc == 0, abi().csm_last_error()
        out = np.empty((wy.value, wx.value), dtype=np.uint8)
        rc = host().csmh_matcher_level(self.ctx.h, self.h, i, out.ctypes.data_as(C.c_void_p),
                                       C.byref(wx), C.byref(wy))
        assert rc == 0, abi().csm_last_error()
        return out

    def quad(self, i):
        """Quad-phase words of level i as a [2s, 2s, jn, in] uint32 array (pv, pu, j, i)."""
        A = abi()
        A.csm_stack2d_quad.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                       C.POINTER(C.c_int), C.POINTER(C.c_int)]
        stack = C.c_void_p(host().csmh_matcher_stack(self.h))
        n_i, n_j = C.c_int(), C.c_int()
        rc = A.csm_stack2d_quad(C.c_void_p(self.ctx.raw), stack, i, None, C.byref(n_i), C.byref(n_j))
        assert rc == 0, A.csm_last_error()
        out = np.empty((2 << i, 2 << i, n_j.value, n_i.value), dtype=np.uint32)
        rc = A.csm_stack2d_quad(C.c_void_p(self.ctx.raw), stack, i, out.ctypes.data_as(C.c_void_p),
                                C.byref(n_i), C.byref(n_j))
        assert rc == 0, A.csm_last_error()
        return out

    def __del__(self):
        if getattr(self, "h", None):
            host().csmh_matcher_destroy(self.h)
            self.h = None

    @classmethod
    def create_many(cls, ctx, grids, linear_window, angular_window, depth):
        """FastCorrelativeScanMatcher2D::CreateMany: one matcher per grid, their stacks built in one
        pipelined call (csm_stack2d_build_many)."""
        L = host()
        L.csmh_matchers_create.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.c_double, C.c_double,
                                           C.c_int, C.POINTER(C.c_void_p)]
        L.csmh_matchers_create.restype = None
        n = len(grids)
        handles = (C.c_void_p * n)(*[g.h for g in grids])
        out = (C.c_void_p * n)()
        L.csmh_matchers_create(ctx.h, n, handles, linear_window, angular_window, depth, out)
        matchers = []
        for i in range(n):
            m = cls.__new__(cls)
            m.ctx, m.grid, m.depth, m.h = ctx, grids[i], depth, out[i]
            matchers.append(m)
        return matchers


class PackedClouds:
    """Point clouds already in the layout the C entry points take: xyz float32 [total, 3] and
    offsets int64 [n + 1] (cloud i = xyz[offsets[i]:offsets[i + 1]]).  Behaves like a list of
    clouds; match_batch passes the two arrays on without a copy."""

    def __init__(self, clouds):
        if isinstance(clouds, PackedClouds):
            self.xyz, self.offsets = clouds.xyz, clouds.offsets
            return
        self.offsets = np.zeros(len(clouds) + 1, dtype=np.int64)
        np.cumsum([len(c) for c in clouds], out=self.offsets[1:])
        if len(clouds) == 0:
            self.xyz = np.zeros((0, 3), np.float32)
        elif all(isinstance(c, np.ndarray) and c.dtype == np.float32 and c.ndim == 2 and c.shape[1] == 3 for c in clouds):
            self.xyz = np.concatenate(clouds, axis=0)  # one copy
        else:
            self.xyz = np.ascontiguousarray(np.concatenate([np.asarray(c, np.float32).reshape(-1, 3) for c in clouds],
                                                           axis=0), dtype=np.float32)

    def __len__(self):
        return len(self.offsets) - 1

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[k] for k in range(*i.indices(len(self)))]
        return self.xyz[self.offsets[i]:self.offsets[i + 1]]

    def __iter__(self):
        return (self[i] for i in range(len(self)))


def _pack(matchers, clouds, min_scores, initial):
    n = len(clouds)
    handles = (C.c_void_p * n)(*[m.h for m in matchers])
    init = np.zeros((n, 3), dtype=np.float64)
    if initial is None:
        full = np.ones(n, dtype=np.int32)
    else:
        full = np.zeros(n, dtype=np.int32)
        for i in range(n):
            if initial[i] is None:
                full[i] = 1
            else:
                init[i] = initial[i]
    if isinstance(clouds, PackedClouds):
        return n, handles, full, init.reshape(-1), clouds.xyz, clouds.offsets, \
            np.ascontiguousarray(min_scores, dtype=np.float32)
    offsets = np.zeros(n + 1, dtype=np.int64)
    np.cumsum([len(c) for c in clouds], out=offsets[1:])
    if n and all(isinstance(c, np.ndarray) and c.dtype == np.float32 and c.ndim == 2 and c.shape[1] == 3 for c in clouds):
        xyz = np.concatenate(clouds, axis=0)  # already [k, 3] float32: one copy
    else:
        xyz = np.ascontiguousarray(np.concatenate([np.asarray(c, np.float32).reshape(-1, 3)
                                                   for c in clouds], axis=0), dtype=np.float32)
    return n, handles, full, init.reshape(-1), xyz, offsets, \
        np.ascontiguousarray(min_scores, dtype=np.float32)


def match_batch(ctx, matchers, clouds, min_scores, initial=None):
    """initial[i] = None -> MatchFullSubmap, else (x, y, theta) -> Match."""
    n, handles, full, init, xyz, offsets, ms = _pack(matchers, clouds, min_scores, initial)
    out = (Result2D * n)()
    rc = host().csmh_match_batch(ctx.h, n, handles, full, init, xyz, offsets, ms, out)
    assert rc == 0
    return out


def rt_match_batch(ctx, grids, clouds, initial, linear_search_window=0.1,
                   angular_search_window=float(np.deg2rad(20.0)),
                   translation_delta_cost_weight=0.1, rotation_delta_cost_weight=0.1):
    """RealTimeCorrelativeScanMatcher2D::Match for every (grid, cloud, initial pose)."""
    n = len(clouds)
    handles = (C.c_void_p * n)(*[g.h for g in grids])
    opts = np.array([linear_search_window, angular_search_window, translation_delta_cost_weight,
                     rotation_delta_cost_weight], dtype=np.float64)
    init = np.ascontiguousarray(initial, dtype=np.float64).reshape(-1)
    packed = clouds if isinstance(clouds, PackedClouds) else PackedClouds(clouds)
    xyz, offsets = packed.xyz, packed.offsets
    out = (Result2D * n)()
    rc = host().csmh_rt_match_batch(ctx.h, n, handles, opts, init, xyz, offsets, out)
    assert rc == 0
    return out


def grid_insert(ctx, grid, origins, returns, misses=None, hit_probability=0.55, miss_probability=0.49,
                insert_free_space=True, resident=False):
    """ProbabilityGridRangeDataInserter2D::Insert of every scan into `grid` (a Grid; it may grow).
    Returns (cells [ny, nx] uint16, max_x, max_y) of the grid afterwards."""
    L = host()
    f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
    i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
    u16p = np.ctypeslib.ndpointer(np.uint16, flags="C_CONTIGUOUS")
    L.csmh_grid_insert.argtypes = [C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_int, C.c_int, C.c_int,
                                   f32p, f32p, i64p, f32p, i64p]
    L.csmh_grid_info.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_double),
                                 C.POINTER(C.c_double)]
    L.csmh_grid_info.restype = None
    L.csmh_grid_cells.argtypes = [C.c_void_p, u16p]
    L.csmh_grid_cells.restype = None
    n = len(origins)
    misses = misses if misses is not None else [np.zeros((0, 3), np.float32)] * n

    def pack(clouds):
        off = np.zeros(n + 1, dtype=np.int64)
        arrs = []
        for i, c in enumerate(clouds):
            a = np.asarray(c, np.float32).reshape(-1, 3)
            arrs.append(a)
            off[i + 1] = off[i] + len(a)
        flat = np.concatenate(arrs, axis=0) if arrs else np.zeros((0, 3), np.float32)
        if len(flat) == 0:
            flat = np.zeros((1, 3), np.float32)
        return np.ascontiguousarray(flat, dtype=np.float32), off

    r, roff = pack(returns)
    m, moff = pack(misses)
    rc = L.csmh_grid_insert(ctx.h, grid.h, hit_probability, miss_probability, 1 if insert_free_space else 0,
                            1 if resident else 0, n,
                            np.ascontiguousarray(np.asarray(origins, np.float32)[:, :2]).reshape(-1), r, roff, m, moff)
    assert rc == 0
    nx, ny = C.c_int(), C.c_int()
    mx, my = C.c_double(), C.c_double()
    L.csmh_grid_info(grid.h, C.byref(nx), C.byref(ny), C.byref(mx), C.byref(my))
    cells = np.empty((ny.value, nx.value), dtype=np.uint16)
    L.csmh_grid_cells(grid.h, cells)
    return cells, mx.value, my.value


def ceres_match_batch(ctx, grids, clouds, targets, initial, occupied_space_weight=1.0,
                      translation_weight=10.0, rotation_weight=40.0, use_nonmonotonic_steps=False,
                      max_num_iterations=20):
    """CeresScanMatcher2D::Match for every (grid, cloud, target translation, initial pose):
    returns (poses [n, 3], summaries [n, 6] = initial_cost, final_cost, iterations, successful,
    unsuccessful steps, termination type)."""
    n = len(clouds)
    L = host()
    f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
    f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
    i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
    L.csmh_ceres_match_batch.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), f64p, f64p, f64p,
                                         f32p, i64p, f64p, f64p]
    handles = (C.c_void_p * n)(*[g.h for g in grids])
    opts = np.array([occupied_space_weight, translation_weight, rotation_weight,
                     1.0 if use_nonmonotonic_steps else 0.0, max_num_iterations], dtype=np.float64)
    offsets = np.zeros(n + 1, dtype=np.int64)
    for i, c in enumerate(clouds):
        offsets[i + 1] = offsets[i] + len(c)
    xyz = np.ascontiguousarray(np.concatenate([np.asarray(c, np.float32).reshape(-1, 3)
                                               for c in clouds], axis=0), dtype=np.float32)
    poses, summaries = np.zeros((n, 3)), np.zeros((n, 6))
    rc = L.csmh_ceres_match_batch(ctx.h, n, handles, opts,
                                  np.ascontiguousarray(targets, np.float64).reshape(-1),
                                  np.ascontiguousarray(initial, np.float64).reshape(-1), xyz, offsets,
                                  poses.reshape(-1), summaries.reshape(-1))
    assert rc == 0
    return poses, summaries


def discretize(ctx, res, max_x, max_y, nx, ny, pts, init, linear_window, angular_window,
               shrink=True):
    """Device GenerateRotatedScans + DiscretizeScans (+ ShrinkToFit); returns
    (xy[int32 S, n, 2], bounds[int32 S, 4])."""
    pts = np.ascontiguousarray(pts, dtype=np.float32)
    ns = C.c_int()
    args = (ctx.h, res, max_x, max_y, nx, ny, pts, len(pts), init[0], init[1], init[2],
            linear_window, angular_window, 1 if shrink else 0, C.byref(ns))
    rc = host().csmh_discretize(*args, None, None)
    assert rc == 0, abi().csm_last_error()
    xy = np.empty((ns.value, len(pts), 2), dtype=np.int32)
    bounds = np.empty((ns.value, 4), dtype=np.int32)
    rc = host().csmh_discretize(*args, xy.ctypes.data_as(C.c_void_p),
                                bounds.ctypes.data_as(C.c_void_p))
    assert rc == 0, abi().csm_last_error()
    return xy, bounds


def constraint_builder_run(grids, submap_poses, search_submap, search_full, rel_poses, clouds,
                           devices=(0,), pool_threads=4, flush_threshold=256, depth=7,
                           sampling_ratio=1.0, max_constraint_distance=15.0, min_score=0.55,
                           global_min_score=0.6, linear_window=7.0,
                           angular_window=float(np.deg2rad(30.0)), translation_weight=1.1e4,
                           rotation_weight=1e5, ceres=None, burst=0, extended=False):
    """Drives ConstraintBuilder2D (MaybeAdd*Constraint / NotifyEndOfNode / WhenDone) over the
    given searches; returns (seconds, found[int32 n], transform[n, 3], stats[4]).  burst > 0:
    the searches enter through MaybeAddConstraints in bursts of that many requests.  extended:
    stats is the 32-double record of csmh_constraint_builder_run2 (metrics, per-device counts)."""
    n = len(clouds)
    handles = (C.c_void_p * len(grids))(*[g.h for g in grids])
    dev = np.ascontiguousarray(devices, dtype=np.int32)
    c = ceres or {}  # ceres_scan_matcher options; None = no refinement
    opts = np.array([sampling_ratio, max_constraint_distance, min_score, global_min_score,
                     linear_window, angular_window, translation_weight, rotation_weight,
                     c.get("occupied_space_weight", 0.0), c.get("translation_weight", 0.0),
                     c.get("rotation_weight", 0.0), 1.0 if c.get("use_nonmonotonic_steps") else 0.0,
                     c.get("max_num_iterations", 0)], dtype=np.float64)
    offsets = np.zeros(n + 1, dtype=np.int64)
    for i, c in enumerate(clouds):
        offsets[i + 1] = offsets[i] + len(c)
    xyz = np.ascontiguousarray(np.concatenate([np.asarray(c, np.float32).reshape(-1, 3)
                                               for c in clouds], axis=0), dtype=np.float32)
    found = np.zeros(n, dtype=np.int32)
    transform = np.zeros((n, 3), dtype=np.float64)
    stats = np.zeros(4, dtype=np.float64)
    ext = np.zeros(32, dtype=np.float64)
    secs = host().csmh_constraint_builder_run2(
        len(dev), dev, pool_threads, flush_threshold, depth, opts, len(grids), handles,
        np.ascontiguousarray(submap_poses, dtype=np.float64).reshape(-1), n,
        np.ascontiguousarray(search_submap, dtype=np.int32),
        np.ascontiguousarray(search_full, dtype=np.int32),
        np.ascontiguousarray(rel_poses, dtype=np.float64).reshape(-1), xyz, offsets, found,
        transform.reshape(-1), stats, 1 if burst > 0 else 0, int(burst), ext)
    return secs, found, transform, (ext if extended else stats)


class PreparedBatch:
    """Host parameters computed and inputs uploaded; run() is asynchronous."""

    def __init__(self, ctx, matchers, clouds, min_scores, initial=None):
        self.n, handles, full, init, xyz, offsets, ms = _pack(matchers, clouds, min_scores, initial)
        self.keep = (matchers, ctx)
        self.h = host().csmh_batch_create(ctx.h, self.n, handles, full, init, xyz, offsets, ms)
        self.h2d_bytes = xyz.nbytes

    def run(self):
        host().csmh_batch_run(self.h)

    def fetch(self):
        out = (Result2D * self.n)()
        host().csmh_batch_fetch(self.h, out)
        return out

    def close(self):
        if self.h:
            host().csmh_batch_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()


# ---- 3D -----------------------------------------------------------------------------
IDENTITY_POSE7 = np.array([0, 0, 0, 1, 0, 0, 0], np.float64)


class Matcher3D:
    """FastCorrelativeScanMatcher3D over voxel lists (cell index, uint16 value)."""

    def __init__(self, ctx, resolution, xyz, values, low_resolution, low_xyz, low_values, histogram,
                 options7):
        self.ctx = ctx
        xyz = np.ascontiguousarray(xyz, np.int32).reshape(-1)
        low_xyz = np.ascontiguousarray(low_xyz, np.int32).reshape(-1)
        values = np.ascontiguousarray(values, np.uint16)
        low_values = np.ascontiguousarray(low_values, np.uint16)
        histogram = np.ascontiguousarray(histogram, np.float32)
        self.depth = int(options7[0])
        self.h = host().csmh_fcsm3d_create(ctx.h, resolution, xyz, values, len(values), low_resolution,
                                           low_xyz, low_values, len(low_values), histogram,
                                           len(histogram), np.ascontiguousarray(options7, np.float64))

    def level(self, i):
        """(lo[3], uint8 [nz, ny, nx]) of precomputation level i."""
        lo, dims = np.zeros(3, np.int32), np.zeros(3, np.int32)
        rc = host().csmh_fcsm3d_level(self.ctx.h, self.h, i, None, lo, dims)
        if rc != 0:
            raise RuntimeError(abi().csm_last_error().decode())
        out = np.zeros((int(dims[2]), int(dims[1]), int(dims[0])), np.uint8)
        if out.size:
            host().csmh_fcsm3d_level(self.ctx.h, self.h, i, out.ctypes.data_as(C.c_void_p), lo, dims)
        return lo, out

    @property
    def grid_size(self):
        return host().csmh_fcsm3d_grid_size(self.h)

    def close(self):
        if getattr(self, "h", None):
            host().csmh_fcsm3d_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()


class Node3D:
    """TrajectoryNode::Data of one node (3D fields)."""

    def __init__(self, high_points, low_points, histogram, gravity=(1., 0., 0., 0.)):
        hi = np.ascontiguousarray(high_points, np.float32).reshape(-1, 3)
        lo = np.ascontiguousarray(low_points, np.float32).reshape(-1, 3)
        hist = np.ascontiguousarray(histogram, np.float32)
        self.num_high, self.num_low = len(hi), len(lo)
        self.h = host().csmh_node3d_create(hi.reshape(-1), len(hi), lo.reshape(-1), len(lo), hist, len(hist),
                                           np.ascontiguousarray(gravity, np.float64))

    def close(self):
        if getattr(self, "h", None):
            host().csmh_node3d_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()


def match3d(matcher, node, min_score, node_pose7=IDENTITY_POSE7, submap_pose7=IDENTITY_POSE7, full=False):
    """The reference's Match / MatchFullSubmap: Result3D or None."""
    r = Result3D()
    ok = host().csmh_fcsm3d_match(matcher.h, node.h, 1 if full else 0,
                                  np.ascontiguousarray(node_pose7, np.float64),
                                  np.ascontiguousarray(submap_pose7, np.float64), min_score, C.byref(r))
    return r if ok else None


def discrete_scans3d(ctx, matcher, node, node_pose7=IDENTITY_POSE7, submap_pose7=IDENTITY_POSE7):
    args = (ctx.h, matcher.h, node.h, np.ascontiguousarray(node_pose7, np.float64),
            np.ascontiguousarray(submap_pose7, np.float64))
    n = host().csmh_fcsm3d_discrete_scans(*args, None, None, None)
    poses = np.zeros((n, 7), np.float32)
    rot = np.zeros(n, np.float32)
    cells = np.zeros((n, node.num_high, 3), np.int32)
    if n:
        rc = host().csmh_fcsm3d_discrete_scans(*args, poses.ctypes.data_as(C.c_void_p),
                                               rot.ctypes.data_as(C.c_void_p),
                                               cells.ctypes.data_as(C.c_void_p))
        if rc < 0:
            raise RuntimeError(abi().csm_last_error().decode())
    return poses, rot, cells


class PreparedBatch3D:
    """Host-side per-search work done and inputs uploaded; run() is asynchronous."""

    def __init__(self, ctx, matchers, nodes, min_scores, node_poses=None, submap_poses=None, full=None):
        n = len(matchers)
        self.n = n
        self.keep = (matchers, nodes, ctx)
        mh = (C.c_void_p * n)(*[m.h for m in matchers])
        nh = (C.c_void_p * n)(*[d.h for d in nodes])
        ident = np.tile(IDENTITY_POSE7, (n, 1))
        node_poses = ident if node_poses is None else np.ascontiguousarray(node_poses, np.float64)
        submap_poses = ident if submap_poses is None else np.ascontiguousarray(submap_poses, np.float64)
        full = np.zeros(n, np.int32) if full is None else np.ascontiguousarray(full, np.int32)
        self.h = host().csmh_batch3d_create(ctx.h, n, mh, nh, full, node_poses.reshape(-1),
                                            submap_poses.reshape(-1),
                                            np.ascontiguousarray(min_scores, np.float32))

    def run(self):
        host().csmh_batch3d_run(self.h)

    def fetch(self):
        out = (Result3D * self.n)()
        host().csmh_batch3d_fetch(self.h, out)
        return out

    def close(self):
        if getattr(self, "h", None):
            host().csmh_batch3d_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()


def match3d_batch(ctx, matchers, nodes, min_scores, node_poses=None, submap_poses=None, full=None):
    b = PreparedBatch3D(ctx, matchers, nodes, min_scores, node_poses, submap_poses, full)
    b.run()
    out = b.fetch()
    b.close()
    return out


def rotational_match(submap_hist, hist, initial_angle, angles):
    """RotationalScanMatcher::Match of the host layer (no device needed)."""
    submap_hist = np.ascontiguousarray(submap_hist, np.float32)
    hist = np.ascontiguousarray(hist, np.float32)
    angles = np.ascontiguousarray(angles, np.float32)
    out = np.zeros(len(angles), np.float32)
    host().csmh_rotational_match(submap_hist, hist, len(hist), initial_angle, angles, len(angles), out)
    return out


def constraint_builder3d_run(submap_matchers, search_submap, search_full, nodes, node_poses, submap_poses,
                             fcsm_options7, sampling_ratio=1.0, max_constraint_distance=15.0, min_score=0.55,
                             global_localization_min_score=0.6, devices=(0,), pool_threads=4,
                             flush_threshold=256, burst=0):
    """Drives ConstraintBuilder3D (MaybeAdd(Global)Constraint / NotifyEndOfNode / WhenDone).
    burst > 0: the searches enter through MaybeAddConstraints in bursts of that many requests.
    Returns (found[n], transform[n, 7], stats dict, seconds)."""
    n = len(nodes)
    sm = (C.c_void_p * len(submap_matchers))(*[m.h for m in submap_matchers])
    nh = (C.c_void_p * n)(*[d.h for d in nodes])
    found = np.zeros(n, np.int32)
    transform = np.zeros((n, 7), np.float64)
    stats = np.zeros(4, np.float64)
    opts = np.array([sampling_ratio, max_constraint_distance, min_score, global_localization_min_score,
                     1.1e4, 1e5], np.float64)
    L = host()
    L.csmh_constraint_builder3d_run2.restype = C.c_double
    L.csmh_constraint_builder3d_run2.argtypes = L.csmh_constraint_builder3d_run.argtypes + [C.c_int]
    secs = L.csmh_constraint_builder3d_run2(
        len(devices), np.ascontiguousarray(devices, np.int32), pool_threads, flush_threshold, opts,
        np.ascontiguousarray(fcsm_options7, np.float64), len(submap_matchers), sm, n,
        np.ascontiguousarray(search_submap, np.int32), np.ascontiguousarray(search_full, np.int32), nh,
        np.ascontiguousarray(node_poses, np.float64).reshape(-1),
        np.ascontiguousarray(submap_poses, np.float64).reshape(-1), found, transform.reshape(-1), stats, int(burst))
    return found, transform, dict(searches=int(stats[0]), found=int(stats[1]),
                                  candidates_scored=int(stats[2]), flushes=int(stats[3])), secs


def rt_options(linear_search_window=0.15, angular_search_window=np.radians(1.0),
               translation_delta_cost_weight=1e-1, rotation_delta_cost_weight=1e-1):
    return np.array([linear_search_window, angular_search_window, translation_delta_cost_weight,
                     rotation_delta_cost_weight], np.float64)


def rt_match3d(resolution, xyz, values, options4, initial_pose7, points):
    """RealTimeCorrelativeScanMatcher3D::Match (reference signature): (score, pose7)."""
    xyz = np.ascontiguousarray(xyz, np.int32).reshape(-1)
    values = np.ascontiguousarray(values, np.uint16)
    pts = np.ascontiguousarray(points, np.float32).reshape(-1, 3)
    pose = np.zeros(7, np.float64)
    score = host().csmh_rt_match3d(resolution, xyz, values, len(values), np.ascontiguousarray(options4, np.float64),
                                   np.ascontiguousarray(initial_pose7, np.float64), pts.reshape(-1), len(pts), pose)
    return score, pose


class DeviceGrid3D:
    """A HybridGrid (voxel list) resident on the device for RealTimeCorrelativeScanMatcher3D."""

    def __init__(self, ctx, resolution, xyz, values):
        L = host()
        L.csmh_device_grid3d_create.restype = C.c_void_p
        L.csmh_device_grid3d_create.argtypes = [C.c_void_p, C.c_float,
                                                np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS"),
                                                np.ctypeslib.ndpointer(np.uint16, flags="C_CONTIGUOUS"), C.c_int]
        L.csmh_device_grid3d_destroy.argtypes = [C.c_void_p]
        self.ctx = ctx
        xyz = np.ascontiguousarray(xyz, np.int32).reshape(-1)
        values = np.ascontiguousarray(values, np.uint16)
        self.h = L.csmh_device_grid3d_create(ctx.h, resolution, xyz, values, len(values))

    def match(self, options4, initial_poses7, clouds):
        """(scores, poses[n, 7], best_index, num_candidates) for every cloud against this grid."""
        L = host()
        f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
        f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
        i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
        i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
        L.csmh_rt_match3d_resident.argtypes = [C.c_void_p, C.c_void_p, f64p, C.c_int, f64p, f32p, i64p, f32p,
                                               f64p, i32p, i32p]
        n = len(clouds)
        offsets = np.zeros(n + 1, np.int64)
        offsets[1:] = np.cumsum([len(c) for c in clouds])
        pts = np.ascontiguousarray(np.concatenate([np.asarray(c, np.float32).reshape(-1, 3) for c in clouds]))
        scores = np.zeros(n, np.float32)
        poses = np.zeros((n, 7), np.float64)
        best = np.zeros(n, np.int32)
        num = np.zeros(n, np.int32)
        L.csmh_rt_match3d_resident(self.ctx.h, self.h, np.ascontiguousarray(options4, np.float64), n,
                                   np.ascontiguousarray(initial_poses7, np.float64).reshape(-1),
                                   pts.reshape(-1), offsets, scores, poses.reshape(-1), best, num)
        return scores, poses, best, num

    def __del__(self):
        if getattr(self, "h", None):
            host().csmh_device_grid3d_destroy(self.h)
            self.h = None


def rt_match3d_batch(ctx, resolution, xyz, values, options4, initial_poses7, clouds):
    """Several clouds against one grid state: (scores, poses[n, 7], best_index, num_candidates)."""
    xyz = np.ascontiguousarray(xyz, np.int32).reshape(-1)
    values = np.ascontiguousarray(values, np.uint16)
    n = len(clouds)
    offsets = np.zeros(n + 1, np.int64)
    offsets[1:] = np.cumsum([len(c) for c in clouds])
    pts = np.ascontiguousarray(np.concatenate([np.asarray(c, np.float32).reshape(-1, 3) for c in clouds]))
    scores = np.zeros(n, np.float32)
    poses = np.zeros((n, 7), np.float64)
    best = np.zeros(n, np.int32)
    num = np.zeros(n, np.int32)
    host().csmh_rt_match3d_batch(ctx.h, resolution, xyz, values, len(values),
                                 np.ascontiguousarray(options4, np.float64), n,
                                 np.ascontiguousarray(initial_poses7, np.float64).reshape(-1), pts.reshape(-1),
                                 offsets, scores, poses.reshape(-1), best, num)
    return scores, poses, best, num


def grid3d_insert(ctx, resolution, xyz, values, origins, returns, hit_probability=0.55, miss_probability=0.49,
                  num_free_space_voxels=2, resident=False):
    """RangeDataInserter3D::Insert of every scan (origins[i], returns[i]) into the HybridGrid given by
    its non-default voxels.  resident=False: the reference-signature call per scan; True: one
    device-resident grid for all scans.  Returns (xyz [n, 3] int32, values [n] uint16, grid_size),
    voxels sorted by (z, y, x)."""
    L = host()
    f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
    f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
    i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
    i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
    u16p = np.ctypeslib.ndpointer(np.uint16, flags="C_CONTIGUOUS")
    L.csmh_grid3d_insert.argtypes = [C.c_void_p, C.c_float, i32p, u16p, C.c_int, f64p, C.c_int, f32p, f32p, i64p,
                                     C.c_int]
    L.csmh_grid3d_insert.restype = C.c_void_p
    L.csmh_grid3d_result_size.argtypes = [C.c_void_p]
    L.csmh_grid3d_result_grid_size.argtypes = [C.c_void_p]
    L.csmh_grid3d_result_copy.argtypes = [C.c_void_p, i32p, u16p]
    L.csmh_grid3d_result_copy.restype = None
    L.csmh_grid3d_result_destroy.argtypes = [C.c_void_p]
    L.csmh_grid3d_result_destroy.restype = None
    xyz = np.ascontiguousarray(np.asarray(xyz, np.int32).reshape(-1, 3))
    values = np.ascontiguousarray(values, np.uint16)
    n = len(origins)
    offsets = np.zeros(n + 1, dtype=np.int64)
    for i, r in enumerate(returns):
        offsets[i + 1] = offsets[i] + len(r)
    pts = np.ascontiguousarray(np.concatenate([np.asarray(r, np.float32).reshape(-1, 3) for r in returns], axis=0)
                               if n else np.zeros((0, 3), np.float32), dtype=np.float32)
    h = L.csmh_grid3d_insert(ctx.h, resolution, xyz.reshape(-1), values, len(values),
                             np.array([hit_probability, miss_probability, num_free_space_voxels], np.float64), n,
                             np.ascontiguousarray(origins, np.float32).reshape(-1), pts.reshape(-1), offsets,
                             1 if resident else 0)
    assert h
    m = L.csmh_grid3d_result_size(h)
    out_xyz = np.zeros((m, 3), np.int32)
    out_val = np.zeros(m, np.uint16)
    if m:
        L.csmh_grid3d_result_copy(h, out_xyz.reshape(-1), out_val)
    size = L.csmh_grid3d_result_grid_size(h)
    L.csmh_grid3d_result_destroy(h)
    order = np.lexsort((out_xyz[:, 0], out_xyz[:, 1], out_xyz[:, 2]))
    return out_xyz[order], out_val[order], size


class ProtoGrid:
    """A grid handle made from a serialized proto::Grid2D message (io/grid_proto_reader.h): usable
    wherever a Grid / TsdfGrid is (Matcher, rt_match_batch, ...)."""

    def __init__(self, serialized):
        L = host()
        L.csmh_grid_from_proto.argtypes = [C.c_char_p, C.c_int64, C.POINTER(C.c_double)]
        L.csmh_grid_from_proto.restype = C.c_void_p
        info = (C.c_double * 6)()
        self.h = L.csmh_grid_from_proto(serialized, len(serialized), info)
        self.nx, self.ny = int(info[0]), int(info[1])
        self.res, self.max_x, self.max_y = info[2], info[3], info[4]
        self.is_tsdf = bool(info[5])

    def cells(self):
        u16p = np.ctypeslib.ndpointer(np.uint16, flags="C_CONTIGUOUS")
        host().csmh_grid_cells.argtypes = [C.c_void_p, u16p]
        host().csmh_grid_cells.restype = None
        out = np.empty((self.ny, self.nx), np.uint16)
        host().csmh_grid_cells(self.h, out)
        return out

    def weight_cells(self):
        u16p = np.ctypeslib.ndpointer(np.uint16, flags="C_CONTIGUOUS")
        host().csmh_grid_weight_cells.argtypes = [C.c_void_p, u16p]
        host().csmh_grid_weight_cells.restype = None
        out = np.empty((self.ny, self.nx), np.uint16)
        host().csmh_grid_weight_cells(self.h, out)
        return out

    def quantization_table(self):
        out = np.empty(65536, dtype=np.uint8)
        host().csmh_quantization_table(self.h, out)
        return out

    def __del__(self):
        if getattr(self, "h", None):
            host().csmh_grid_destroy(self.h)
            self.h = None


def hybrid_grid_from_proto(serialized):
    """(resolution, grid_size, xyz [n, 3] int32 sorted by (z, y, x), values [n] uint16) of a serialized
    proto::HybridGrid."""
    L = host()
    L.csmh_hybrid_grid_from_proto.argtypes = [C.c_char_p, C.c_int64, C.POINTER(C.c_float), C.POINTER(C.c_int),
                                              C.c_void_p, C.c_void_p]
    res, size = C.c_float(), C.c_int()
    n = L.csmh_hybrid_grid_from_proto(serialized, len(serialized), C.byref(res), C.byref(size), None, None)
    xyz = np.zeros((n, 3), np.int32)
    values = np.zeros(n, np.uint16)
    if n:
        L.csmh_hybrid_grid_from_proto(serialized, len(serialized), C.byref(res), C.byref(size),
                                      xyz.ctypes.data_as(C.c_void_p), values.ctypes.data_as(C.c_void_p))
    order = np.lexsort((xyz[:, 0], xyz[:, 1], xyz[:, 2]))
    return res.value, size.value, xyz[order], values[order]


def grid3d_resident_stack_check(ctx, resolution, xyz, values, origins, returns, depth, full_resolution_depth,
                                hit_probability=0.55, miss_probability=0.49, num_free_space_voxels=2):
    """Inserts the scans into a device-resident HybridGrid, builds the matcher's stack straight from it and
    from its exported voxel list; returns the number of differing level bytes (0 = identical)."""
    L = host()
    f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
    f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
    i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
    i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
    u16p = np.ctypeslib.ndpointer(np.uint16, flags="C_CONTIGUOUS")
    L.csmh_grid3d_resident_stack_check.argtypes = [C.c_void_p, C.c_float, i32p, u16p, C.c_int, f64p, C.c_int, f32p,
                                                   f32p, i64p, C.c_int, C.c_int]
    L.csmh_grid3d_resident_stack_check.restype = C.c_int64
    xyz = np.ascontiguousarray(np.asarray(xyz, np.int32).reshape(-1, 3))
    values = np.ascontiguousarray(values, np.uint16)
    n = len(origins)
    offsets = np.zeros(n + 1, dtype=np.int64)
    for i, r in enumerate(returns):
        offsets[i + 1] = offsets[i] + len(r)
    pts = np.ascontiguousarray(np.concatenate([np.asarray(r, np.float32).reshape(-1, 3) for r in returns], axis=0),
                               dtype=np.float32)
    return L.csmh_grid3d_resident_stack_check(
        ctx.h, resolution, xyz.reshape(-1), values, len(values),
        np.array([hit_probability, miss_probability, num_free_space_voxels], np.float64), n,
        np.ascontiguousarray(origins, np.float32).reshape(-1), pts.reshape(-1), offsets, depth, full_resolution_depth)


def tsdf_insert(ctx, grid, origins, returns, options9, resident=False):
    """TSDFRangeDataInserter2D::Insert of every scan (origins[i] = (x, y), returns[i] = [n, 3]) into a
    TsdfGrid (it may grow).  options9: the 9 doubles of TSDFRangeDataInserterOptions2D (see
    host_capi.cc).  Returns (tsd cells, weight cells) uint16 [ny, nx] and (max_x, max_y)."""
    L = host()
    f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
    f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
    i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
    u16p = np.ctypeslib.ndpointer(np.uint16, flags="C_CONTIGUOUS")
    L.csmh_tsdf_insert.argtypes = [C.c_void_p, C.c_void_p, f64p, C.c_int, f32p, f32p, i64p, C.c_int]
    L.csmh_grid_info.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_double),
                                 C.POINTER(C.c_double)]
    L.csmh_grid_info.restype = None
    L.csmh_grid_cells.argtypes = [C.c_void_p, u16p]
    L.csmh_grid_cells.restype = None
    L.csmh_grid_weight_cells.argtypes = [C.c_void_p, u16p]
    L.csmh_grid_weight_cells.restype = None
    n = len(origins)
    offsets = np.zeros(n + 1, dtype=np.int64)
    for i, r in enumerate(returns):
        offsets[i + 1] = offsets[i] + len(r)
    pts = np.ascontiguousarray(np.concatenate([np.asarray(r, np.float32).reshape(-1, 3) for r in returns], axis=0),
                               dtype=np.float32)
    rc = L.csmh_tsdf_insert(ctx.h, grid.h, np.ascontiguousarray(options9, np.float64), n,
                            np.ascontiguousarray(np.asarray(origins, np.float32)[:, :2]).reshape(-1), pts.reshape(-1),
                            offsets, 1 if resident else 0)
    assert rc == 0
    nx, ny = C.c_int(), C.c_int()
    mx, my = C.c_double(), C.c_double()
    L.csmh_grid_info(grid.h, C.byref(nx), C.byref(ny), C.byref(mx), C.byref(my))
    grid.nx, grid.ny = nx.value, ny.value
    tsd = np.empty((grid.ny, grid.nx), np.uint16)
    weight = np.empty((grid.ny, grid.nx), np.uint16)
    L.csmh_grid_cells(grid.h, tsd)
    L.csmh_grid_weight_cells(grid.h, weight)
    return tsd, weight, (mx.value, my.value)


# ---------------------------------------------------------------------------------------------
# .pbstream ingestion (host/cartographer/io/proto_stream*.h)
# ---------------------------------------------------------------------------------------------
def _io():
    L = host()
    if getattr(L, "_io_bound", False):
        return L
    u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")
    u16p = np.ctypeslib.ndpointer(np.uint16, flags="C_CONTIGUOUS")
    f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
    f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
    i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
    i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
    L.csmh_pbstream_write.argtypes = [C.c_char_p, C.c_int, u8p, i64p]
    L.csmh_pbstream_read.argtypes = [C.c_char_p, u8p, C.c_int64, i64p, C.c_int]
    L.csmh_pbstream_read.restype = C.c_int64
    L.csmh_pbstream_count.argtypes = [C.c_char_p]
    L.csmh_pbstream_count.restype = C.c_int64
    L.csmh_state_load.argtypes = [C.c_char_p, C.c_int, i32p]
    L.csmh_state_load.restype = C.c_void_p
    L.csmh_state_destroy.argtypes = [C.c_void_p]
    L.csmh_state_destroy.restype = None
    L.csmh_state_submap.argtypes = [C.c_void_p, C.c_int, i32p, f64p]
    L.csmh_state_submap.restype = None
    L.csmh_state_submap_grid.argtypes = [C.c_void_p, C.c_int, f64p]
    L.csmh_state_submap_grid.restype = C.c_void_p
    L.csmh_state_submap_hybrid_grid.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_float),
                                                C.POINTER(C.c_int), C.c_void_p, C.c_void_p]
    L.csmh_state_submap_histogram.argtypes = [C.c_void_p, C.c_int, f32p]
    L.csmh_state_submap_histogram.restype = None
    L.csmh_state_node.argtypes = [C.c_void_p, C.c_int, i32p, C.POINTER(C.c_int64), f64p]
    L.csmh_state_node.restype = None
    L.csmh_state_node_cloud.argtypes = [C.c_void_p, C.c_int, C.c_int, f32p]
    L.csmh_state_node_cloud.restype = None
    L.csmh_state_node_histogram.argtypes = [C.c_void_p, C.c_int, f32p]
    L.csmh_state_node_histogram.restype = None
    L.csmh_state_constraints.argtypes = [C.c_void_p, i32p, f64p]
    L.csmh_state_constraints.restype = None
    L._io_bound = True
    return L


def pbstream_write(path, messages):
    """io::ProtoStreamWriter: serialized messages (bytes) -> a .pbstream file."""
    offsets = np.zeros(len(messages) + 1, dtype=np.int64)
    for i, m in enumerate(messages):
        offsets[i + 1] = offsets[i] + len(m)
    blob = np.frombuffer(b"".join(messages) or b"\0", dtype=np.uint8).copy()
    rc = _io().csmh_pbstream_write(os.fsencode(path), len(messages), blob, offsets)
    if rc != 0:
        raise IOError("could not write %s" % path)


def pbstream_read(path):
    """io::ProtoStreamReader: every message of a .pbstream file, decompressed."""
    L = _io()
    count = L.csmh_pbstream_count(os.fsencode(path))
    if count < 0:
        raise IOError("truncated proto stream %s" % path)
    offsets = np.zeros(count + 1, dtype=np.int64)
    need = L.csmh_pbstream_read(os.fsencode(path), np.zeros(1, np.uint8), 0, offsets, count)
    size = -(need + 1) if need < 0 else 0
    blob = np.zeros(max(1, size), dtype=np.uint8)
    got = L.csmh_pbstream_read(os.fsencode(path), blob, size, offsets, count)
    assert got == count
    return [blob[offsets[i]:offsets[i + 1]].tobytes() for i in range(count)]


class _StateGrid(ProtoGrid):
    """The grid of a loaded 2D submap (a handle like ProtoGrid's)."""

    def __init__(self, handle, info):
        self.h = handle
        self.nx, self.ny = int(info[0]), int(info[1])
        self.res, self.max_x, self.max_y = info[2], info[3], info[4]
        self.is_tsdf = bool(info[5])


class MappingState:
    """A saved mapping state (.pbstream) read through io::ProtoStreamDeserializer and the
    kSubmap / kNode cases of MapBuilder::LoadState.  `submaps` and `nodes` are lists of dicts in
    stream order; poses are (tx, ty, tz, qw, qx, qy, qz).  A 2D submap's "grid" is a grid handle
    every matcher entry point takes; a 3D submap carries "high" / "low" = (resolution, grid_size,
    xyz[n, 3] int32, values[n] uint16) sorted by voxel index, and its "histogram"."""

    def __init__(self, path, expect_3d=False):
        L = _io()
        counts = np.zeros(4, dtype=np.int32)
        self.h = L.csmh_state_load(os.fsencode(path), 1 if expect_3d else 0, counts)
        self.format_version = int(counts[3])
        self.submaps, self.nodes = [], []
        for i in range(int(counts[0])):
            ints, poses = np.zeros(8, np.int32), np.zeros(14, np.float64)
            L.csmh_state_submap(self.h, i, ints, poses)
            s = {"trajectory_id": int(ints[0]), "submap_index": int(ints[1]), "is_3d": bool(ints[3]),
                 "num_range_data": int(ints[4]), "finished": bool(ints[5]),
                 "local_pose": poses[:7].copy(), "global_pose": poses[7:].copy(), "grid": None}
            if ints[2] and ints[6]:
                info = np.zeros(6, np.float64)
                s["grid"] = _StateGrid(L.csmh_state_submap_grid(self.h, i, info), info)
            if ints[3]:
                for which, key in ((0, "high"), (1, "low")):
                    s[key] = self._hybrid(i, which) if ints[6] & (1 << which) else None
                hist = np.zeros(int(ints[7]), np.float32)
                if len(hist):
                    L.csmh_state_submap_histogram(self.h, i, hist)
                s["histogram"] = hist
            self.submaps.append(s)
        for i in range(int(counts[1])):
            ints, doubles, stamp = np.zeros(8, np.int32), np.zeros(18, np.float64), C.c_int64()
            L.csmh_state_node(self.h, i, ints, C.byref(stamp), doubles)
            n = {"trajectory_id": int(ints[0]), "node_index": int(ints[1]), "timestamp": stamp.value,
                 "gravity_alignment": doubles[:4].copy(), "local_pose": doubles[4:11].copy(),
                 "global_pose": doubles[11:].copy()}
            for which, key in ((0, "filtered_gravity_aligned_point_cloud"), (1, "high_resolution_point_cloud"),
                               (2, "low_resolution_point_cloud")):
                cloud = np.zeros((int(ints[2 + which]), 3), np.float32)
                if len(cloud):
                    L.csmh_state_node_cloud(self.h, i, which, cloud.reshape(-1))
                n[key] = cloud
            hist = np.zeros(int(ints[5]), np.float32)
            if len(hist):
                L.csmh_state_node_histogram(self.h, i, hist)
            n["histogram"] = hist
            self.nodes.append(n)
        nc = int(counts[2])
        ints, doubles = np.zeros(max(1, 5 * nc), np.int32), np.zeros(max(1, 9 * nc), np.float64)
        if nc:
            L.csmh_state_constraints(self.h, ints, doubles)
        self.constraints = [{"submap_id": (int(ints[5 * c]), int(ints[5 * c + 1])),
                             "node_id": (int(ints[5 * c + 2]), int(ints[5 * c + 3])),
                             "tag": "INTER_SUBMAP" if ints[5 * c + 4] else "INTRA_SUBMAP",
                             "relative_pose": doubles[9 * c:9 * c + 7].copy(),
                             "translation_weight": float(doubles[9 * c + 7]),
                             "rotation_weight": float(doubles[9 * c + 8])} for c in range(nc)]

    def _hybrid(self, i, which):
        L = _io()
        res, size = C.c_float(), C.c_int()
        n = L.csmh_state_submap_hybrid_grid(self.h, i, which, C.byref(res), C.byref(size), None, None)
        xyz, values = np.zeros((max(1, n), 3), np.int32), np.zeros(max(1, n), np.uint16)
        L.csmh_state_submap_hybrid_grid(self.h, i, which, C.byref(res), C.byref(size),
                                        xyz.ctypes.data_as(C.c_void_p), values.ctypes.data_as(C.c_void_p))
        xyz, values = xyz[:n], values[:n]
        order = np.lexsort((xyz[:, 0], xyz[:, 1], xyz[:, 2]))
        return float(res.value), int(size.value), xyz[order], values[order]

    def close(self):
        if getattr(self, "h", None):
            host().csmh_state_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()


# ---------------------------------------------------------------------------------------------
# Voxel filters of the node clouds (host/cartographer/sensor/internal/voxel_filter.h; host code)
# ---------------------------------------------------------------------------------------------
def _voxel_filter(which, points, a, b=0.0, c=0.0):
    L = host()
    f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
    L.csmh_voxel_filter.argtypes = [C.c_int, f32p, C.c_int, C.c_float, C.c_float, C.c_float, f32p]
    pts = np.ascontiguousarray(np.asarray(points, np.float32).reshape(-1, 3))
    out = np.zeros((max(1, len(pts)), 3), np.float32)
    n = L.csmh_voxel_filter(which, pts.reshape(-1) if len(pts) else np.zeros(3, np.float32), len(pts), a, b, c,
                            out.reshape(-1))
    return out[:n]


def voxel_filter(points, resolution):
    """sensor::VoxelFilter: one point per voxel of edge `resolution`, reservoir-sampled in point order."""
    return _voxel_filter(0, points, resolution)


def adaptive_voxel_filter(points, max_length, min_num_points, max_range):
    """sensor::AdaptiveVoxelFilter."""
    return _voxel_filter(1, points, max_length, min_num_points, max_range)


def voxel_filter_edge(points, rad, voxeledgeratio):
    """sensor::VoxelFilterEdge (this fork's addition)."""
    return _voxel_filter(2, points, rad, voxeledgeratio)


def compress_point_cloud(points):
    """sensor::CompressedPointCloud(points): the int32 words of proto::CompressedPointCloud.point_data."""
    L = host()
    f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
    L.csmh_compress_point_cloud.argtypes = [f32p, C.c_int, C.c_void_p, C.c_int64]
    L.csmh_compress_point_cloud.restype = C.c_int64
    pts = np.ascontiguousarray(np.asarray(points, np.float32).reshape(-1, 3))
    flat = pts.reshape(-1) if len(pts) else np.zeros(3, np.float32)
    n = L.csmh_compress_point_cloud(flat, len(pts), None, 0)
    words = np.zeros(max(1, n), dtype=np.int32)
    L.csmh_compress_point_cloud(flat, len(pts), words.ctypes.data_as(C.c_void_p), n)
    return words[:n]


# ---- writing grids and states back (host/cartographer/io/grid_proto_writer.h) ----
def _sized_bytes(call):
    n = call(None, 0)
    if n < 0:
        raise RuntimeError("serialization failed (%d)" % n)
    buf = np.zeros(max(1, n), dtype=np.uint8)
    got = call(buf.ctypes.data_as(C.c_void_p), n)
    assert got == n
    return buf[:n].tobytes()


def grid_to_proto(grid):
    """Grid2D::ToProto().SerializeAsString() of a grid handle (Grid, TsdfGrid, ProtoGrid, a state's grid)."""
    L = host()
    L.csmh_grid_to_proto.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
    L.csmh_grid_to_proto.restype = C.c_int64
    return _sized_bytes(lambda out, cap: L.csmh_grid_to_proto(grid.h, out, cap))


def hybrid_grid_to_proto(resolution, xyz, values):
    """HybridGrid::ToProto().SerializeAsString() of the grid holding the given voxels."""
    L = host()
    i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
    u16p = np.ctypeslib.ndpointer(np.uint16, flags="C_CONTIGUOUS")
    L.csmh_hybrid_grid_to_proto.argtypes = [C.c_float, i32p, u16p, C.c_int, C.c_void_p, C.c_int64]
    L.csmh_hybrid_grid_to_proto.restype = C.c_int64
    xyz = np.ascontiguousarray(xyz, np.int32).reshape(-1)
    values = np.ascontiguousarray(values, np.uint16)
    if len(values) == 0:
        xyz, values = np.zeros(3, np.int32), np.zeros(1, np.uint16)
        n = 0
    else:
        n = len(values)
    return _sized_bytes(lambda out, cap: L.csmh_hybrid_grid_to_proto(resolution, xyz, values, n, out, cap))


def reencode_pose_graph(record):
    """A serialized SerializedData{pose_graph} record parsed and written again."""
    L = host()
    L.csmh_reencode_pose_graph.argtypes = [C.c_char_p, C.c_int64, C.c_void_p, C.c_int64]
    L.csmh_reencode_pose_graph.restype = C.c_int64
    return _sized_bytes(lambda out, cap: L.csmh_reencode_pose_graph(record, len(record), out, cap))


def _state_write(self, path):
    """WriteStateToFile: submap and node records re-encoded from the state, the others as read."""
    L = _io()
    L.csmh_state_write.argtypes = [C.c_void_p, C.c_char_p]
    L.csmh_state_write.restype = None
    L.csmh_state_write(self.h, os.fsencode(path))


def _state_set_submap_cells(self, i, cells):
    """Replaces the cells of 2D submap i's grid (same limits), e.g. after insertions on the device."""
    L = _io()
    L.csmh_state_set_submap_cells.argtypes = [C.c_void_p, C.c_int, np.ctypeslib.ndpointer(np.uint16, flags="C_CONTIGUOUS"),
                                              C.c_int64]
    cells = np.ascontiguousarray(cells, np.uint16)
    if L.csmh_state_set_submap_cells(self.h, i, cells.reshape(-1), cells.size) != 0:
        raise ValueError("submap %d has no grid of %d cells" % (i, cells.size))


MappingState.write = _state_write
MappingState.set_submap_cells = _state_set_submap_cells
