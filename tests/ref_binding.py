"""ctypes binding of oracle/_ref/libref.so (TEST INFRASTRUCTURE ONLY): the reference's own 2D
scan-matcher sources compiled unmodified by `make -C oracle ref` (see oracle/ref_capi.cc).

The library only exists where it was built: in the container that has /root/reference, and
on the GPU box as a prebuilt file that travelled with the snapshot.  available() says which.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from oracle_binding import IDENTITY_POSE7, OracleResult2D, OracleResult3D

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_ORACLE_DIR = os.path.join(_ROOT, "oracle")
_LIB_PATH = os.path.join(_ORACLE_DIR, "_ref", "libref.so")
_REFERENCE = "/root/reference"

_lib = None


def available():
    return os.path.exists(_LIB_PATH) or os.path.isdir(os.path.join(_REFERENCE, "cartographer"))


def build_ref():
    """Builds oracle/_ref/libref.so from the sources under /root/reference, where they exist."""
    if os.path.isdir(os.path.join(_REFERENCE, "cartographer")):
        subprocess.check_call(["make", "-s", "-C", _ORACLE_DIR, "-j8", "ref"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is not None:
        return _lib
    build_ref()
    L = C.CDLL(_LIB_PATH)
    u16p = np.ctypeslib.ndpointer(np.uint16, flags="C_CONTIGUOUS")
    f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
    i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
    assert L.oracle_is_reference() == 1
    L.oracle_correspondence_cost_table.argtypes = [f32p]
    L.oracle_probability_to_cost_value.argtypes = [C.c_float]
    L.oracle_probability_to_cost_value.restype = C.c_uint16
    L.oracle_grid2d_create.argtypes = [u16p, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double]
    L.oracle_grid2d_create.restype = C.c_void_p
    L.oracle_grid2d_destroy.argtypes = [C.c_void_p]
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
    L.oracle_fcsm2d_create.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_int]
    L.oracle_fcsm2d_create.restype = C.c_void_p
    L.oracle_fcsm2d_destroy.argtypes = [C.c_void_p]
    L.oracle_stack2d_build_seconds.argtypes = [C.c_void_p, C.c_int]
    L.oracle_stack2d_build_seconds.restype = C.c_double
    L.oracle_fcsm2d_level.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.oracle_fcsm2d_match.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_double, f32p, C.c_int,
                                      C.c_float, C.POINTER(OracleResult2D)]
    L.oracle_search_parameters2d.argtypes = [C.c_double, C.c_double, f32p, C.c_int, C.c_double,
                                             C.POINTER(C.c_int), C.POINTER(C.c_double), C.POINTER(C.c_int),
                                             C.POINTER(C.c_int)]
    L.oracle_discretize2d.argtypes = [C.c_double, C.c_double, C.c_double, C.c_int, C.c_int, f32p, C.c_int,
                                      C.c_double, C.c_double, C.c_double, C.c_int, C.c_int, C.c_double, C.c_int,
                                      i32p, i32p]
    L.oracle_rt_match2d.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double,
                                    C.c_double, C.c_double, f32p, C.c_int, C.POINTER(OracleResult2D)]
    L.oracle_rt_match2d.restype = C.c_double
    L.oracle_grid2d_insert.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_int, f32p, f32p, C.c_int, f32p,
                                       C.c_int]
    L.oracle_grid2d_insert.restype = None
    L.oracle_grid2d_info.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_double),
                                     C.POINTER(C.c_double), C.POINTER(C.c_double)]
    L.oracle_grid2d_info.restype = None
    L.oracle_grid2d_cells.argtypes = [C.c_void_p, u16p]
    L.oracle_grid2d_cells.restype = None
    L.oracle_apply_odds_table.argtypes = [C.c_double, u16p]
    L.oracle_apply_odds_table.restype = None
    i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
    f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
    L.oracle_constraint_batch.argtypes = [
        C.c_int, C.POINTER(C.c_void_p), f64p, C.c_int, C.c_int, i32p, i32p, f64p, f32p, i64p,
        C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int,
        C.POINTER(OracleResult2D), f64p]
    L.oracle_constraint_batch.restype = C.c_double
    L.oracle_grid3d_create.argtypes = [C.c_float, i32p, u16p, C.c_int]
    L.oracle_grid3d_create.restype = C.c_void_p
    L.oracle_grid3d_destroy.argtypes = [C.c_void_p]
    L.oracle_grid3d_size.argtypes = [C.c_void_p]
    L.oracle_grid3d_cells.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.oracle_grid3d_insert.argtypes = [C.c_void_p, f32p, f32p, C.c_int, C.c_float, C.c_float, C.c_int]
    L.oracle_grid3d_insert.restype = None
    L.oracle_fcsm3d_create.argtypes = [C.c_void_p, C.c_void_p, f32p, C.c_int, f64p]
    L.oracle_fcsm3d_create.restype = C.c_void_p
    L.oracle_fcsm3d_destroy.argtypes = [C.c_void_p]
    L.oracle_fcsm3d_level.argtypes = [C.c_void_p, C.c_int, i32p, i32p, C.c_void_p]
    L.oracle_fcsm3d_match.argtypes = [C.c_void_p, C.c_int, f32p, f32p, f32p, C.c_int, f32p, C.c_int, f32p, C.c_int,
                                      f64p, C.c_float, C.POINTER(OracleResult3D)]
    L.oracle_rt_match3d.restype = C.c_float
    L.oracle_rt_match3d.argtypes = [C.c_void_p, f64p, f32p, f32p, C.c_int, f32p, C.POINTER(C.c_int),
                                    C.POINTER(C.c_int)]
    L.oracle_rotational_match.argtypes = [f32p, f32p, C.c_int, C.c_float, f32p, C.c_int, f32p]
    _lib = L
    return L


class RefGrid:
    """cartographer::mapping::ProbabilityGrid holding the given uint16 cells."""

    def __init__(self, cells, res, max_x, max_y):
        cells = np.ascontiguousarray(cells, dtype=np.uint16)
        self.ny, self.nx = cells.shape
        self.h = lib().oracle_grid2d_create(cells, self.nx, self.ny, res, max_x, max_y)

    def cells(self):
        out = np.empty((self.ny, self.nx), dtype=np.uint16)
        lib().oracle_grid2d_cells(self.h, out)
        return out

    def insert(self, origin, returns, misses=(), hit_probability=0.55, miss_probability=0.49,
               insert_free_space=True):
        """ProbabilityGridRangeDataInserter2D::Insert (the grid may grow); returns (cells, max_x, max_y)."""
        def xyz(points):
            a = np.asarray(points, dtype=np.float32).reshape(-1, 3) if len(points) else np.zeros((0, 3), np.float32)
            return np.ascontiguousarray(a)
        r, m = xyz(returns), xyz(misses)
        lib().oracle_grid2d_insert(self.h, hit_probability, miss_probability, 1 if insert_free_space else 0,
                                   np.asarray(origin, np.float32)[:2].copy(), r, len(r), m, len(m))
        nx, ny = C.c_int(), C.c_int()
        mx, my, res = C.c_double(), C.c_double(), C.c_double()
        lib().oracle_grid2d_info(self.h, C.byref(nx), C.byref(ny), C.byref(mx), C.byref(my), C.byref(res))
        self.nx, self.ny = nx.value, ny.value
        return self.cells(), mx.value, my.value

    def __del__(self):
        if getattr(self, "h", None):
            lib().oracle_grid2d_destroy(self.h)
            self.h = None


class RefTsdf:
    """cartographer::mapping::TSDF2D holding the given tsd / weight cells."""

    def __init__(self, tsd_cells, weight_cells, res, max_x, max_y, truncation_distance, max_weight):
        tsd_cells = np.ascontiguousarray(tsd_cells, dtype=np.uint16)
        weight_cells = np.ascontiguousarray(weight_cells, dtype=np.uint16)
        self.ny, self.nx = tsd_cells.shape
        self.h = lib().oracle_tsdf2d_create(tsd_cells, weight_cells, self.nx, self.ny, res, max_x, max_y,
                                            truncation_distance, max_weight)

    def __del__(self):
        if getattr(self, "h", None):
            lib().oracle_tsdf2d_destroy(self.h)
            self.h = None


class RefFCSM2D:
    """cartographer::mapping::scan_matching::FastCorrelativeScanMatcher2D on a RefGrid / RefTsdf."""

    def __init__(self, grid, linear_window, angular_window, depth):
        self.grid = grid
        self.h = lib().oracle_fcsm2d_create(grid.h, linear_window, angular_window, depth)

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
        lib().oracle_fcsm2d_match(self.h, 1 if full else 0, ix, iy, it, pts, len(pts), C.c_float(min_score),
                                  C.byref(r))
        return r


def stack_build_seconds(grid, depth):
    """Wall seconds of one PrecomputationGridStack2D construction on a RefGrid."""
    return lib().oracle_stack2d_build_seconds(grid.h, depth)


def search_parameters(linear_window, angular_window, pts, res):
    pts = np.ascontiguousarray(pts, dtype=np.float32)
    nap, step, ns, nlin = C.c_int(), C.c_double(), C.c_int(), C.c_int()
    lib().oracle_search_parameters2d(linear_window, angular_window, pts, len(pts), res, C.byref(nap),
                                     C.byref(step), C.byref(ns), C.byref(nlin))
    return nap.value, step.value, ns.value, nlin.value


def discretize(res, max_x, max_y, nx, ny, pts, init, nlin, nap, step, shrink=True):
    pts = np.ascontiguousarray(pts, dtype=np.float32)
    s = 2 * nap + 1
    xy = np.empty((s, len(pts), 2), dtype=np.int32)
    bounds = np.empty((s, 4), dtype=np.int32)
    lib().oracle_discretize2d(res, max_x, max_y, nx, ny, pts, len(pts), init[0], init[1], init[2], nlin, nap,
                              step, 1 if shrink else 0, xy, bounds)
    return xy, bounds


def rt_match(grid, opts, init, pts):
    pts = np.ascontiguousarray(pts, dtype=np.float32)
    r = OracleResult2D()
    lib().oracle_rt_match2d(grid.h, opts["linear_search_window"], opts["angular_search_window"],
                            opts["translation_delta_cost_weight"], opts["rotation_delta_cost_weight"],
                            init[0], init[1], init[2], pts, len(pts), C.byref(r))
    return r


def correspondence_cost_table():
    t = np.empty(65536, dtype=np.float32)
    lib().oracle_correspondence_cost_table(t)
    return t


def apply_odds_table(probability):
    t = np.empty(32768, dtype=np.uint16)
    lib().oracle_apply_odds_table(probability, t)
    return t


def constraint_batch(num_threads, grids, submap_poses, search_submap, search_full, rel_poses, clouds,
                     sampling_ratio=1.0, max_constraint_distance=15.0, min_score=0.55, global_min_score=0.6,
                     linear_window=7.0, angular_window=np.deg2rad(30.0), depth=7):
    """The constraint searches of ConstraintBuilder2D on the reference's own matchers, thread-pooled
    (same arguments and return as oracle_binding.constraint_batch; stats are zeros)."""
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
        num_threads, handles, np.ascontiguousarray(submap_poses, dtype=np.float64).reshape(-1), len(grids), n,
        np.ascontiguousarray(search_submap, dtype=np.int32), np.ascontiguousarray(search_full, dtype=np.int32),
        np.ascontiguousarray(rel_poses, dtype=np.float64).reshape(-1), xyz, offsets, sampling_ratio,
        max_constraint_distance, min_score, global_min_score, linear_window, angular_window, depth, out, stats)
    return secs, out, stats


# ---- 3D ---------------------------------------------------------------------------
class RefGrid3D:
    """cartographer::mapping::HybridGrid holding the given uint16 values."""

    def __init__(self, resolution, xyz=None, values=None):
        xyz = np.zeros((0, 3), np.int32) if xyz is None else np.ascontiguousarray(xyz, np.int32)
        values = np.zeros(0, np.uint16) if values is None else np.ascontiguousarray(values, np.uint16)
        self.resolution = float(np.float32(resolution))
        self.h = lib().oracle_grid3d_create(resolution, xyz.reshape(-1), values, len(values))

    def cells(self, iteration_order=False):
        """Non-default voxels, sorted by (z, y, x) or in the order the reference's own iterator visits them."""
        n = lib().oracle_grid3d_cells(self.h, None, None)
        xyz = np.zeros((n, 3), np.int32)
        values = np.zeros(n, np.uint16)
        if n:
            lib().oracle_grid3d_cells(self.h, xyz.ctypes.data_as(C.c_void_p), values.ctypes.data_as(C.c_void_p))
        if iteration_order:
            return xyz, values
        order = np.lexsort((xyz[:, 0], xyz[:, 1], xyz[:, 2]))
        return xyz[order], values[order]

    def insert(self, origin, returns, hit_probability=0.7, miss_probability=0.4, num_free_space_voxels=5):
        returns = np.ascontiguousarray(returns, np.float32)
        lib().oracle_grid3d_insert(self.h, np.ascontiguousarray(origin, np.float32), returns.reshape(-1),
                                   len(returns), hit_probability, miss_probability, num_free_space_voxels)

    @property
    def grid_size(self):
        return lib().oracle_grid3d_size(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            lib().oracle_grid3d_destroy(self.h)
            self.h = None


class RefFCSM3D:
    """cartographer::mapping::scan_matching::FastCorrelativeScanMatcher3D."""

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
        lo, hi = np.zeros(3, np.int32), np.zeros(3, np.int32)
        lib().oracle_fcsm3d_level(self.h, i, lo, hi, None)
        shape = tuple(int(v) for v in (hi - lo + 1)[::-1])
        out = np.zeros(shape, np.uint8)
        lib().oracle_fcsm3d_level(self.h, i, lo, hi, out.ctypes.data_as(C.c_void_p))
        return lo, out

    def match(self, hi_pts, lo_pts, histogram, min_score, node_pose7=IDENTITY_POSE7, submap_pose7=IDENTITY_POSE7,
              gravity=(1., 0., 0., 0.), full=False):
        hi_pts = np.ascontiguousarray(hi_pts, np.float32)
        lo_pts = np.ascontiguousarray(lo_pts, np.float32)
        histogram = np.ascontiguousarray(histogram, np.float32)
        r = OracleResult3D()
        lib().oracle_fcsm3d_match(self.h, 1 if full else 0, np.ascontiguousarray(node_pose7, np.float32),
                                  np.ascontiguousarray(submap_pose7, np.float32), hi_pts.reshape(-1), len(hi_pts),
                                  lo_pts.reshape(-1), len(lo_pts), histogram, len(histogram),
                                  np.ascontiguousarray(gravity, np.float64), min_score, C.byref(r))
        return r


def rotational_match(submap_hist, hist, initial_angle, angles):
    submap_hist = np.ascontiguousarray(submap_hist, np.float32)
    hist = np.ascontiguousarray(hist, np.float32)
    angles = np.ascontiguousarray(angles, np.float32)
    out = np.zeros(len(angles), np.float32)
    lib().oracle_rotational_match(submap_hist, hist, len(hist), initial_angle, angles, len(angles), out)
    return out


def rt_match3d(grid, options4, initial_pose7, points):
    """RealTimeCorrelativeScanMatcher3D::Match: (score, pose7 float32)."""
    pts = np.ascontiguousarray(points, np.float32).reshape(-1, 3)
    pose = np.zeros(7, np.float32)
    best, num = C.c_int(), C.c_int()
    score = lib().oracle_rt_match3d(grid.h, np.ascontiguousarray(options4, np.float64),
                                    np.ascontiguousarray(initial_pose7, np.float32), pts.reshape(-1), len(pts), pose,
                                    C.byref(best), C.byref(num))
    return score, pose


def tsdf_insert(grid, origin, returns, opts=None):
    """TSDFRangeDataInserter2D::Insert on a RefTsdf (it may grow); sets grid.tsd_cells / weight_cells / limits."""
    import oracle_binding as ob
    return ob._tsdf_insert(lib(), grid, origin, returns, opts)


# ---------------------------------------------------------------------------------------------
# The reference's own .pbstream container and point-cloud compression (oracle/ref_capi_io.cc)
# ---------------------------------------------------------------------------------------------
def _io():
    L = lib()
    if getattr(L, "_io_bound", False):
        return L
    u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")
    i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
    f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
    L.oracle_pbstream_write.argtypes = [C.c_char_p, C.c_int, u8p, i64p]
    L.oracle_pbstream_read.argtypes = [C.c_char_p, u8p, C.c_int64, i64p, C.c_int]
    L.oracle_pbstream_read.restype = C.c_int64
    L.oracle_gzip.argtypes = [u8p, C.c_int64, C.c_int, u8p, C.c_int64]
    L.oracle_gzip.restype = C.c_int64
    L.oracle_compress_point_cloud.argtypes = [f32p, C.c_int, C.c_void_p, C.c_int64]
    L.oracle_compress_point_cloud.restype = C.c_int64
    L.oracle_decompress_point_cloud.argtypes = [np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS"), C.c_int64,
                                                C.c_int, f32p]
    L.oracle_decompress_point_cloud.restype = None
    L._io_bound = True
    return L


def pbstream_write(path, messages):
    """The reference's ProtoStreamWriter on serialized messages."""
    offsets = np.zeros(len(messages) + 1, dtype=np.int64)
    for i, m in enumerate(messages):
        offsets[i + 1] = offsets[i] + len(m)
    blob = np.frombuffer(b"".join(messages) or b"\0", dtype=np.uint8).copy()
    assert _io().oracle_pbstream_write(os.fsencode(path), len(messages), blob, offsets) == 0


def pbstream_read(path, max_messages=1 << 16):
    """The reference's ProtoStreamReader: every message, decompressed."""
    L = _io()
    offsets = np.zeros(max_messages + 1, dtype=np.int64)
    need = L.oracle_pbstream_read(os.fsencode(path), np.zeros(1, np.uint8), 0, offsets, max_messages)
    size = -(need + 1) if need < 0 else 0
    blob = np.zeros(max(1, size), dtype=np.uint8)
    count = L.oracle_pbstream_read(os.fsencode(path), blob, size, offsets, max_messages)
    assert count >= 0
    return [blob[offsets[i]:offsets[i + 1]].tobytes() for i in range(count)]


def gzip_string(data, decompress=False):
    """common::FastGzipString / FastGunzipString."""
    L = _io()
    src = np.frombuffer(data or b"\0", dtype=np.uint8).copy()
    need = L.oracle_gzip(src, len(data), 1 if decompress else 0, np.zeros(1, np.uint8), 0)
    size = -(need + 1) if need < 0 else need
    out = np.zeros(max(1, size), dtype=np.uint8)
    got = L.oracle_gzip(src, len(data), 1 if decompress else 0, out, size)
    assert got == size
    return out[:size].tobytes()


def compress_point_cloud(points):
    """sensor::CompressedPointCloud(points).ToProto().point_data() as int32 words."""
    L = _io()
    pts = np.ascontiguousarray(np.asarray(points, np.float32).reshape(-1, 3))
    n = L.oracle_compress_point_cloud(pts.reshape(-1), len(pts), None, 0)
    words = np.zeros(max(1, n), dtype=np.int32)
    L.oracle_compress_point_cloud(pts.reshape(-1), len(pts), words.ctypes.data_as(C.c_void_p), n)
    return words[:n]


def decompress_point_cloud(num_points, words):
    out = np.zeros((max(1, num_points), 3), dtype=np.float32)
    w = np.ascontiguousarray(words, dtype=np.int32)
    if len(w) == 0:
        w = np.zeros(1, np.int32)
    _io().oracle_decompress_point_cloud(w, len(words), num_points, out.reshape(-1))
    return out[:num_points]


def voxel_filter(which, points, a, b=0.0, c=0.0):
    """The reference's own VoxelFilter (which = 0), AdaptiveVoxelFilter (1), VoxelFilterEdge (2)."""
    L = lib()
    f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
    L.oracle_voxel_filter.argtypes = [C.c_int, f32p, C.c_int, C.c_float, C.c_float, C.c_float, f32p]
    pts = np.ascontiguousarray(np.asarray(points, np.float32).reshape(-1, 3))
    out = np.zeros((max(1, len(pts)), 3), np.float32)
    n = L.oracle_voxel_filter(which, pts.reshape(-1) if len(pts) else np.zeros(3, np.float32), len(pts), a, b, c,
                              out.reshape(-1))
    return out[:n]
