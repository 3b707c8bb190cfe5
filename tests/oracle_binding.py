"""ctypes binding of oracle/liboracle.so (TEST INFRASTRUCTURE ONLY).

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import
this module; the product path never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_ORACLE_DIR = os.path.join(_ROOT, "oracle")
_LIB_PATH = os.path.join(_ORACLE_DIR, "liboracle.so")


class OracleResult2D(C.Structure):
    _fields_ = [
        ("found", C.c_int32),
        ("scan_index", C.c_int32),
        ("x_index_offset", C.c_int32),
        ("y_index_offset", C.c_int32),
        ("score", C.c_float),
        ("num_scans", C.c_int32),
        ("pose_x", C.c_double),
        ("pose_y", C.c_double),
        ("pose_theta", C.c_double),
        ("scored_top", C.c_uint64),
        ("scored_lower", C.c_uint64),
    ]


class OracleResult3D(C.Structure):
    _fields_ = [
        ("found", C.c_int32),
        ("scan_index", C.c_int32),
        ("offset", C.c_int32 * 3),
        ("num_scans", C.c_int32),
        ("score", C.c_float),
        ("rotational_score", C.c_float),
        ("low_resolution_score", C.c_float),
        ("pose", C.c_float * 7),
        ("candidates_scored", C.c_uint64),
    ]


def build_oracle(force=False):
    srcs = [os.path.join(_ORACLE_DIR, f) for f in ("oracle_capi.cc", "csm_oracle.h", "csm_oracle3d.h", "cpu_pool.h")]
    stale = (not os.path.exists(_LIB_PATH)) or any(
        os.path.exists(s) and os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in srcs)
    if force or stale:
        subprocess.check_call(["make", "-C", _ORACLE_DIR, "all"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    build_oracle()
    L = C.CDLL(_LIB_PATH)
    u16p = np.ctypeslib.ndpointer(np.uint16, flags="C_CONTIGUOUS")
    f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
    i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
    i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
    f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
    L.oracle_correspondence_cost_table.argtypes = [f32p]
    L.oracle_probability_to_cost_value.argtypes = [C.c_float]
    L.oracle_probability_to_cost_value.restype = C.c_uint16
    L.oracle_grid2d_create.argtypes = [u16p, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double]
    L.oracle_grid2d_create.restype = C.c_void_p
    L.oracle_grid2d_destroy.argtypes = [C.c_void_p]
    L.oracle_fcsm2d_create.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_int]
    L.oracle_fcsm2d_create.restype = C.c_void_p
    L.oracle_fcsm2d_create_tsdf.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_int]
    L.oracle_fcsm2d_create_tsdf.restype = C.c_void_p
    L.oracle_fcsm2d_destroy.argtypes = [C.c_void_p]
    L.oracle_fcsm2d_level.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.POINTER(C.c_int),
                                      C.POINTER(C.c_int)]
    L.oracle_fcsm2d_match.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_double, f32p,
                                      C.c_int, C.c_float, C.POINTER(OracleResult2D)]
    L.oracle_search_parameters2d.argtypes = [C.c_double, C.c_double, f32p, C.c_int, C.c_double,
                                             C.POINTER(C.c_int), C.POINTER(C.c_double),
                                             C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.oracle_discretize2d.argtypes = [C.c_double, C.c_double, C.c_double, C.c_int, C.c_int, f32p,
                                      C.c_int, C.c_double, C.c_double, C.c_double, C.c_int, C.c_int,
                                      C.c_double, C.c_int, i32p, i32p]
    L.oracle_rt_match2d.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_double,
                                    C.c_double, C.c_double, C.c_double, f32p, C.c_int,
                                    C.POINTER(OracleResult2D)]
    L.oracle_rt_match2d.restype = C.c_double
    L.oracle_grid2d_insert.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_int, f32p, f32p, C.c_int,
                                       f32p, C.c_int]
    L.oracle_grid2d_insert.restype = None
    L.oracle_grid2d_info.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                     C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double)]
    L.oracle_grid2d_info.restype = None
    L.oracle_grid2d_cells.argtypes = [C.c_void_p, u16p]
    L.oracle_grid2d_cells.restype = None
    L.oracle_apply_odds_table.argtypes = [C.c_double, u16p]
    L.oracle_apply_odds_table.restype = None
    L.oracle_ceres_match2d.argtypes = [C.c_void_p, f64p, f64p, f64p, f32p, C.c_int, f64p, f64p]
    L.oracle_ceres_match2d.restype = None
    L.oracle_ceres_evaluate2d.argtypes = [C.c_void_p, f64p, f64p, C.c_double, f64p, f32p, C.c_int,
                                          f64p, f64p]
    L.oracle_ceres_evaluate2d.restype = None
    L.oracle_ceres_match2d_tsdf.argtypes = L.oracle_ceres_match2d.argtypes
    L.oracle_ceres_match2d_tsdf.restype = None
    L.oracle_ceres_evaluate2d_tsdf.argtypes = L.oracle_ceres_evaluate2d.argtypes
    L.oracle_ceres_evaluate2d_tsdf.restype = C.c_int
    L.oracle_tsdf2d_create.argtypes = [u16p, u16p, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double,
                                       C.c_float, C.c_float]
    L.oracle_tsdf2d_create.restype = C.c_void_p
    L.oracle_tsdf2d_destroy.argtypes = [C.c_void_p]
    L.oracle_tsdf2d_insert.argtypes = [C.c_void_p, np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS"), f32p, f32p,
                                       C.c_int]
    L.oracle_tsdf2d_insert.restype = None
    L.oracle_tsdf2d_info.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_double),
                                     C.POINTER(C.c_double), C.POINTER(C.c_double)]
    L.oracle_tsdf2d_info.restype = None
    L.oracle_tsdf2d_cells.argtypes = [C.c_void_p, u16p, u16p]
    L.oracle_tsdf2d_cells.restype = None
    L.oracle_tsdf2d_values.argtypes = [C.c_float, C.c_float, f32p, f32p, C.c_int, u16p, u16p]
    L.oracle_rt_match2d_tsdf.argtypes = L.oracle_rt_match2d.argtypes
    L.oracle_rt_match2d_tsdf.restype = C.c_double
    L.oracle_rt_score_candidate_tsdf.argtypes = [C.c_void_p, f32p, C.c_int, C.c_int, C.c_int]
    L.oracle_rt_score_candidate_tsdf.restype = C.c_float
    L.oracle_constraint_batch.argtypes = [
        C.c_int, C.POINTER(C.c_void_p), f64p, C.c_int, C.c_int, i32p, i32p, f64p, f32p, i64p,
        C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int,
        C.POINTER(OracleResult2D), f64p]
    L.oracle_constraint_batch.restype = C.c_double
    u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")
    L.oracle_grid3d_create.argtypes = [C.c_float, i32p, u16p, C.c_int]
    L.oracle_grid3d_create.restype = C.c_void_p
    L.oracle_grid3d_destroy.argtypes = [C.c_void_p]
    L.oracle_grid3d_size.argtypes = [C.c_void_p]
    L.oracle_grid3d_insert.argtypes = [C.c_void_p, f32p, f32p, C.c_int, C.c_float, C.c_float, C.c_int]
    L.oracle_grid3d_cells.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.oracle_fcsm3d_create.argtypes = [C.c_void_p, C.c_void_p, f32p, C.c_int, f64p]
    L.oracle_fcsm3d_create.restype = C.c_void_p
    L.oracle_fcsm3d_destroy.argtypes = [C.c_void_p]
    L.oracle_fcsm3d_level.argtypes = [C.c_void_p, C.c_int, i32p, i32p, C.c_void_p]
    L.oracle_fcsm3d_match.argtypes = [C.c_void_p, C.c_int, f32p, f32p, f32p, C.c_int, f32p, C.c_int,
                                      f32p, C.c_int, f64p, C.c_float, C.POINTER(OracleResult3D)]
    L.oracle_fcsm3d_discrete_scans.argtypes = [C.c_void_p, f64p, C.c_float, f32p, f32p, f32p, C.c_int,
                                               f32p, C.c_int, f64p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.oracle_rt_match3d.restype = C.c_float
    L.oracle_rt_match3d.argtypes = [C.c_void_p, f64p, f32p, f32p, C.c_int, f32p, C.POINTER(C.c_int),
                                    C.POINTER(C.c_int)]
    L.oracle_rotational_match.argtypes = [f32p, f32p, C.c_int, C.c_float, f32p, C.c_int, f32p]
    _lib = L
    return L


class OracleGrid:
    def __init__(self, cells, res, max_x, max_y):
        cells = np.ascontiguousarray(cells, dtype=np.uint16)
        self.ny, self.nx = cells.shape
        self.res, self.max_x, self.max_y = res, max_x, max_y
        self.cells = cells
        self.h = lib().oracle_grid2d_create(cells, self.nx, self.ny, res, max_x, max_y)

    def __del__(self):
        if getattr(self, "h", None):
            lib().oracle_grid2d_destroy(self.h)
            self.h = None


def _xyz(points):
    a = np.asarray(points, dtype=np.float32).reshape(-1, 3) if len(points) else np.zeros((0, 3), np.float32)
    return np.ascontiguousarray(a)


def grid_insert(grid, origin, returns, misses=(), hit_probability=0.55, miss_probability=0.49,
                insert_free_space=True):
    """ProbabilityGridRangeDataInserter2D::Insert on an OracleGrid (it may grow); refreshes
    grid.cells / nx / ny / max_x / max_y."""
    r, m = _xyz(returns), _xyz(misses)
    lib().oracle_grid2d_insert(grid.h, hit_probability, miss_probability, 1 if insert_free_space else 0,
                               np.asarray(origin, np.float32)[:2].copy(), r, len(r), m, len(m))
    nx, ny = C.c_int(), C.c_int()
    mx, my, res = C.c_double(), C.c_double(), C.c_double()
    lib().oracle_grid2d_info(grid.h, C.byref(nx), C.byref(ny), C.byref(mx), C.byref(my), C.byref(res))
    grid.nx, grid.ny, grid.max_x, grid.max_y = nx.value, ny.value, mx.value, my.value
    cells = np.empty((grid.ny, grid.nx), dtype=np.uint16)
    lib().oracle_grid2d_cells(grid.h, cells)
    grid.cells = cells
    return grid


def apply_odds_table(probability):
    t = np.empty(32768, dtype=np.uint16)
    lib().oracle_apply_odds_table(probability, t)
    return t


class OracleTsdf:
    """TSDF2D restatement (tsd cells, weight cells, truncation distance, max weight)."""

    def __init__(self, tsd_cells, weight_cells, res, max_x, max_y, truncation_distance, max_weight):
        self.tsd_cells = np.ascontiguousarray(tsd_cells, dtype=np.uint16)
        self.weight_cells = np.ascontiguousarray(weight_cells, dtype=np.uint16)
        self.ny, self.nx = self.tsd_cells.shape
        self.res, self.max_x, self.max_y = res, max_x, max_y
        self.truncation_distance, self.max_weight = truncation_distance, max_weight
        self.h = lib().oracle_tsdf2d_create(self.tsd_cells, self.weight_cells, self.nx, self.ny, res,
                                            max_x, max_y, truncation_distance, max_weight)

    def __del__(self):
        if getattr(self, "h", None):
            lib().oracle_tsdf2d_destroy(self.h)
            self.h = None

    def score_candidate(self, pts, x_offset, y_offset):
        pts = np.ascontiguousarray(pts, dtype=np.float32)
        return float(lib().oracle_rt_score_candidate_tsdf(self.h, pts, len(pts), x_offset, y_offset))


def tsdf_values(truncation_distance, max_weight, tsd, weight):
    """TSDValueConverter::TSDToValue / WeightToValue, elementwise."""
    tsd = np.ascontiguousarray(tsd, dtype=np.float32)
    weight = np.ascontiguousarray(weight, dtype=np.float32)
    tv = np.empty(tsd.size, dtype=np.uint16)
    wv = np.empty(tsd.size, dtype=np.uint16)
    lib().oracle_tsdf2d_values(truncation_distance, max_weight, tsd.reshape(-1), weight.reshape(-1),
                               tsd.size, tv, wv)
    return tv.reshape(tsd.shape), wv.reshape(tsd.shape)


class OracleFCSM2D:
    def __init__(self, grid, linear_window, angular_window, depth):
        self.grid = grid
        self.depth = depth
        create = lib().oracle_fcsm2d_create_tsdf if isinstance(grid, OracleTsdf) else lib().oracle_fcsm2d_create
        self.h = create(grid.h, linear_window, angular_window, depth)

    def __del__(self):
        if getattr(self, "h", None):
            lib().oracle_fcsm2d_destroy(self.h)
            self.h = None

    def level(self, i):
        wx, wy = C.c_int(), C.c_int()
        assert lib().oracle_fcsm2d_level(self.h, i, None, C.byref(wx), C.byref(wy)) == 0
        out = np.empty((wy.value, wx.value), dtype=np.uint8)
        lib().oracle_fcsm2d_level(self.h, i, out.ctypes.data_as(C.c_void_p), C.byref(wx), C.byref(wy))
        return out

    def match(self, pts, min_score, initial=None):
        pts = np.ascontiguousarray(pts, dtype=np.float32)
        r = OracleResult2D()
        full = initial is None
        ix, iy, it = (0.0, 0.0, 0.0) if full else initial
        lib().oracle_fcsm2d_match(self.h, 1 if full else 0, ix, iy, it, pts, len(pts),
                                  C.c_float(min_score), C.byref(r))
        return r


def search_parameters(linear_window, angular_window, pts, res):
    pts = np.ascontiguousarray(pts, dtype=np.float32)
    nap, step, ns, nlin = C.c_int(), C.c_double(), C.c_int(), C.c_int()
    lib().oracle_search_parameters2d(linear_window, angular_window, pts, len(pts), res,
                                     C.byref(nap), C.byref(step), C.byref(ns), C.byref(nlin))
    return nap.value, step.value, ns.value, nlin.value


def discretize(res, max_x, max_y, nx, ny, pts, init, nlin, nap, step, shrink=True):
    pts = np.ascontiguousarray(pts, dtype=np.float32)
    s = 2 * nap + 1
    xy = np.empty((s, len(pts), 2), dtype=np.int32)
    bounds = np.empty((s, 4), dtype=np.int32)
    lib().oracle_discretize2d(res, max_x, max_y, nx, ny, pts, len(pts), init[0], init[1], init[2],
                              nlin, nap, step, 1 if shrink else 0, xy, bounds)
    return xy, bounds


def rt_match(grid, opts, init, pts):
    pts = np.ascontiguousarray(pts, dtype=np.float32)
    r = OracleResult2D()
    lib().oracle_rt_match2d(grid.h, opts["linear_search_window"], opts["angular_search_window"],
                            opts["translation_delta_cost_weight"], opts["rotation_delta_cost_weight"],
                            init[0], init[1], init[2], pts, len(pts), C.byref(r))
    return r


CERES_DEFAULTS = {"occupied_space_weight": 1.0, "translation_weight": 10.0, "rotation_weight": 40.0,
                  "use_nonmonotonic_steps": False, "max_num_iterations": 20}


def _ceres_options(opts):
    o = dict(CERES_DEFAULTS)
    o.update(opts or {})
    return np.array([o["occupied_space_weight"], o["translation_weight"], o["rotation_weight"],
                     1.0 if o["use_nonmonotonic_steps"] else 0.0, o["max_num_iterations"]], np.float64)


def ceres_match(grid, opts, target, init, pts):
    """CeresScanMatcher2D::Match restatement: (pose[3], summary dict)."""
    pts = np.ascontiguousarray(pts, dtype=np.float32)
    pose, summ = np.zeros(3), np.zeros(6)
    lib().oracle_ceres_match2d(grid.h, _ceres_options(opts), np.asarray(target, np.float64),
                               np.asarray(init, np.float64), pts, len(pts), pose, summ)
    return pose, {"initial_cost": summ[0], "final_cost": summ[1], "iterations": int(summ[2]),
                  "num_successful_steps": int(summ[3]), "num_unsuccessful_steps": int(summ[4]),
                  "termination": int(summ[5])}


def ceres_match_tsdf(grid, opts, target, init, pts):
    """CeresScanMatcher2D::Match restatement on an OracleTsdf: (pose[3], summary dict)."""
    pts = np.ascontiguousarray(pts, dtype=np.float32)
    pose, summ = np.zeros(3), np.zeros(6)
    lib().oracle_ceres_match2d_tsdf(grid.h, _ceres_options(opts), np.asarray(target, np.float64),
                                    np.asarray(init, np.float64), pts, len(pts), pose, summ)
    return pose, {"initial_cost": summ[0], "final_cost": summ[1], "iterations": int(summ[2]),
                  "num_successful_steps": int(summ[3]), "num_unsuccessful_steps": int(summ[4]),
                  "termination": int(summ[5])}


def ceres_evaluate_tsdf(grid, opts, target, initial_angle, pose, pts):
    """(valid, residuals, jacobian) of the TSDF cost function + the two delta functors."""
    pts = np.ascontiguousarray(pts, dtype=np.float32)
    r, j = np.zeros(len(pts) + 3), np.zeros((len(pts) + 3, 3))
    ok = lib().oracle_ceres_evaluate2d_tsdf(grid.h, _ceres_options(opts), np.asarray(target, np.float64),
                                            float(initial_angle), np.asarray(pose, np.float64), pts,
                                            len(pts), r, j)
    return bool(ok), r, j


def ceres_evaluate(grid, opts, target, initial_angle, pose, pts):
    pts = np.ascontiguousarray(pts, dtype=np.float32)
    r, j = np.zeros(len(pts) + 3), np.zeros((len(pts) + 3, 3))
    lib().oracle_ceres_evaluate2d(grid.h, _ceres_options(opts), np.asarray(target, np.float64),
                                  float(initial_angle), np.asarray(pose, np.float64), pts, len(pts), r, j)
    return r, j


def rt_match_tsdf(grid, opts, init, pts):
    pts = np.ascontiguousarray(pts, dtype=np.float32)
    r = OracleResult2D()
    lib().oracle_rt_match2d_tsdf(grid.h, opts["linear_search_window"], opts["angular_search_window"],
                                 opts["translation_delta_cost_weight"],
                                 opts["rotation_delta_cost_weight"], init[0], init[1], init[2], pts,
                                 len(pts), C.byref(r))
    return r


def constraint_batch(num_threads, grids, submap_poses, search_submap, search_full, rel_poses, clouds,
                     sampling_ratio=1.0, max_constraint_distance=15.0, min_score=0.55,
                     global_min_score=0.6, linear_window=7.0, angular_window=np.deg2rad(30.0),
                     depth=7):
    """Thread-pooled ConstraintBuilder2D restatement; returns (seconds, results, stats)."""
    n = len(clouds)
    handles = (C.c_void_p * len(grids))(*[g.h for g in grids])
    offsets = np.zeros(n + 1, dtype=np.int64)
    for i, c in enumerate(clouds):
        offsets[i + 1] = offsets[i] + len(c)
    xyz = np.ascontiguousarray(np.concatenate(clouds, axis=0), dtype=np.float32) if n else \
        np.zeros((0, 3), np.float32)
    out = (OracleResult2D * n)()
    stats = np.zeros(4, dtype=np.float64)
    secs = lib().oracle_constraint_batch(
        num_threads, handles, np.ascontiguousarray(submap_poses, dtype=np.float64).reshape(-1),
        len(grids), n, np.ascontiguousarray(search_submap, dtype=np.int32),
        np.ascontiguousarray(search_full, dtype=np.int32),
        np.ascontiguousarray(rel_poses, dtype=np.float64).reshape(-1), xyz, offsets,
        sampling_ratio, max_constraint_distance, min_score, global_min_score, linear_window,
        angular_window, depth, out, stats)
    return secs, out, stats


# ---- 3D ---------------------------------------------------------------------------
IDENTITY_POSE7 = np.array([0, 0, 0, 1, 0, 0, 0], np.float32)


def options3d(depth=8, full_resolution_depth=3, min_rotational_score=0.77, min_low_resolution_score=0.55,
              linear_xy_search_window=5., linear_z_search_window=1., angular_search_window=np.radians(15.)):
    """FastCorrelativeScanMatcherOptions3D as the 7 doubles the C entries take."""
    return np.array([depth, full_resolution_depth, min_rotational_score, min_low_resolution_score,
                     linear_xy_search_window, linear_z_search_window, angular_search_window], np.float64)


class OracleGrid3D:
    """HybridGrid: sparse voxel -> uint16 value."""

    def __init__(self, resolution, xyz=None, values=None):
        xyz = np.zeros((0, 3), np.int32) if xyz is None else np.ascontiguousarray(xyz, np.int32)
        values = np.zeros(0, np.uint16) if values is None else np.ascontiguousarray(values, np.uint16)
        self.resolution = float(np.float32(resolution))
        self.h = lib().oracle_grid3d_create(resolution, xyz.reshape(-1), values, len(values))

    def insert(self, origin, returns, hit_probability=0.7, miss_probability=0.4, num_free_space_voxels=5):
        returns = np.ascontiguousarray(returns, np.float32)
        lib().oracle_grid3d_insert(self.h, np.ascontiguousarray(origin, np.float32), returns.reshape(-1),
                                   len(returns), hit_probability, miss_probability, num_free_space_voxels)

    def cells(self):
        n = lib().oracle_grid3d_cells(self.h, None, None)
        xyz = np.zeros((n, 3), np.int32)
        values = np.zeros(n, np.uint16)
        if n:
            lib().oracle_grid3d_cells(self.h, xyz.ctypes.data_as(C.c_void_p), values.ctypes.data_as(C.c_void_p))
        order = np.lexsort((xyz[:, 0], xyz[:, 1], xyz[:, 2]))
        return xyz[order], values[order]

    @property
    def grid_size(self):
        return lib().oracle_grid3d_size(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            lib().oracle_grid3d_destroy(self.h)
            self.h = None


class OracleFCSM3D:
    def __init__(self, grid, low_grid, histogram, options):
        self.grid, self.low_grid = grid, low_grid
        self.options = np.ascontiguousarray(options, np.float64)
        self.histogram = np.ascontiguousarray(histogram, np.float32)
        self.h = lib().oracle_fcsm3d_create(grid.h, low_grid.h, self.histogram, len(self.histogram), self.options)

    def __del__(self):
        if getattr(self, "h", None):
            lib().oracle_fcsm3d_destroy(self.h)
            self.h = None

    def level(self, i):
        """(lo[3], dense uint8 [nz, ny, nx]) over the bounding box of stored cells."""
        lo, hi = np.zeros(3, np.int32), np.zeros(3, np.int32)
        lib().oracle_fcsm3d_level(self.h, i, lo, hi, None)
        shape = tuple(int(v) for v in (hi - lo + 1)[::-1])
        out = np.zeros(shape, np.uint8)
        lib().oracle_fcsm3d_level(self.h, i, lo, hi, out.ctypes.data_as(C.c_void_p))
        return lo, out

    def match(self, hi_pts, lo_pts, histogram, min_score, node_pose7=IDENTITY_POSE7,
              submap_pose7=IDENTITY_POSE7, gravity=(1., 0., 0., 0.), full=False):
        hi_pts = np.ascontiguousarray(hi_pts, np.float32)
        lo_pts = np.ascontiguousarray(lo_pts, np.float32)
        histogram = np.ascontiguousarray(histogram, np.float32)
        r = OracleResult3D()
        lib().oracle_fcsm3d_match(self.h, 1 if full else 0, np.ascontiguousarray(node_pose7, np.float32),
                                  np.ascontiguousarray(submap_pose7, np.float32), hi_pts.reshape(-1),
                                  len(hi_pts), lo_pts.reshape(-1), len(lo_pts), histogram, len(histogram),
                                  np.ascontiguousarray(gravity, np.float64), min_score, C.byref(r))
        return r

    def discrete_scans(self, hi_pts, histogram, node_pose7=IDENTITY_POSE7, submap_pose7=IDENTITY_POSE7,
                       gravity=(1., 0., 0., 0.)):
        hi_pts = np.ascontiguousarray(hi_pts, np.float32)
        histogram = np.ascontiguousarray(histogram, np.float32)
        args = (self.h, self.options, self.grid.resolution, np.ascontiguousarray(node_pose7, np.float32),
                np.ascontiguousarray(submap_pose7, np.float32), hi_pts.reshape(-1), len(hi_pts), histogram,
                len(histogram), np.ascontiguousarray(gravity, np.float64))
        n = lib().oracle_fcsm3d_discrete_scans(*args, None, None, None)
        poses = np.zeros((n, 7), np.float32)
        rot = np.zeros(n, np.float32)
        cells = np.zeros((n, len(hi_pts), 3), np.int32)
        if n:
            lib().oracle_fcsm3d_discrete_scans(*args, poses.ctypes.data_as(C.c_void_p),
                                               rot.ctypes.data_as(C.c_void_p),
                                               cells.ctypes.data_as(C.c_void_p))
        return poses, rot, cells


def rotational_match(submap_hist, hist, initial_angle, angles):
    submap_hist = np.ascontiguousarray(submap_hist, np.float32)
    hist = np.ascontiguousarray(hist, np.float32)
    angles = np.ascontiguousarray(angles, np.float32)
    out = np.zeros(len(angles), np.float32)
    lib().oracle_rotational_match(submap_hist, hist, len(hist), initial_angle, angles, len(angles), out)
    return out


def rt_match3d(grid, options4, initial_pose7, points):
    """RealTimeCorrelativeScanMatcher3D::Match: (score, pose7 float32, best_index, num_candidates)."""
    pts = np.ascontiguousarray(points, np.float32).reshape(-1, 3)
    pose = np.zeros(7, np.float32)
    best, num = C.c_int(), C.c_int()
    score = lib().oracle_rt_match3d(grid.h, np.ascontiguousarray(options4, np.float64),
                                    np.ascontiguousarray(initial_pose7, np.float32), pts.reshape(-1), len(pts),
                                    pose, C.byref(best), C.byref(num))
    return score, pose, best.value, num.value


TSDF_INSERTER_DEFAULTS = {"truncation_distance": 0.3, "maximum_weight": 10.0, "update_free_space": False,
                          "num_normal_samples": 4, "sample_radius": 0.5,
                          "project_sdf_distance_to_scan_normal": True, "update_weight_range_exponent": 0,
                          "update_weight_angle_scan_normal_to_ray_kernel_bandwidth": 0.5,
                          "update_weight_distance_cell_to_hit_kernel_bandwidth": 0.5}


def tsdf_inserter_options(opts=None):
    """TSDFRangeDataInserterOptions2D (defaults of trajectory_builder_2d.lua:100-112) as 9 doubles."""
    o = dict(TSDF_INSERTER_DEFAULTS)
    o.update(opts or {})
    return np.array([o["truncation_distance"], o["maximum_weight"], float(o["update_free_space"]),
                     o["num_normal_samples"], o["sample_radius"], float(o["project_sdf_distance_to_scan_normal"]),
                     o["update_weight_range_exponent"], o["update_weight_angle_scan_normal_to_ray_kernel_bandwidth"],
                     o["update_weight_distance_cell_to_hit_kernel_bandwidth"]], np.float64)


def _tsdf_insert(L, grid, origin, returns, opts):
    r = np.ascontiguousarray(np.asarray(returns, np.float32).reshape(-1, 3))
    L.oracle_tsdf2d_insert(grid.h, tsdf_inserter_options(opts), np.asarray(origin, np.float32)[:2].copy(), r, len(r))
    nx, ny = C.c_int(), C.c_int()
    mx, my, res = C.c_double(), C.c_double(), C.c_double()
    L.oracle_tsdf2d_info(grid.h, C.byref(nx), C.byref(ny), C.byref(mx), C.byref(my), C.byref(res))
    grid.nx, grid.ny, grid.max_x, grid.max_y = nx.value, ny.value, mx.value, my.value
    tsd = np.empty((grid.ny, grid.nx), np.uint16)
    weight = np.empty((grid.ny, grid.nx), np.uint16)
    L.oracle_tsdf2d_cells(grid.h, tsd, weight)
    grid.tsd_cells, grid.weight_cells = tsd, weight
    return grid


def tsdf_insert(grid, origin, returns, opts=None):
    """TSDFRangeDataInserter2D::Insert on an OracleTsdf (it may grow); refreshes its cells and limits."""
    return _tsdf_insert(lib(), grid, origin, returns, opts)
