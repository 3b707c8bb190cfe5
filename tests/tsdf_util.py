"""TSDF2D fixtures for the real-time matcher tests: cells painted directly (the
reference builds its fixture with TSDFRangeDataInserter2D, which is outside the path;
SURVEY.md section 8(c) allows a direct cell-painting fixture)."""
import numpy as np

import oracle_binding as ob


def paint_from_points(nx, ny, res, max_x, max_y, hits, origin, truncation, max_weight, weight=1.0):
    """tsd(cell) = signed distance from the cell centre to the nearest hit (positive on
    the sensor's side), weight `weight`, for cells closer than the truncation distance;
    all other cells stay unknown (0, 0).  Returns (tsd_cells, weight_cells) uint16 [ny, nx]."""
    yi, xi = np.meshgrid(np.arange(ny), np.arange(nx), indexing="ij")
    # cell (x index, y index) has its centre at (max_x - (y + .5) res, max_y - (x + .5) res)
    cx = max_x - (yi + 0.5) * res
    cy = max_y - (xi + 0.5) * res
    hits = np.asarray(hits, dtype=np.float64)
    d2 = (cx[..., None] - hits[:, 0]) ** 2 + (cy[..., None] - hits[:, 1]) ** 2
    k = np.argmin(d2, axis=-1)
    d = np.sqrt(np.take_along_axis(d2, k[..., None], axis=-1)[..., 0])
    r_cell = np.hypot(cx - origin[0], cy - origin[1])
    r_hit = np.hypot(hits[k, 0] - origin[0], hits[k, 1] - origin[1])
    tsd = np.where(r_cell <= r_hit, d, -d).astype(np.float32)
    known = d < truncation
    tv, wv = ob.tsdf_values(truncation, max_weight, tsd, np.full(tsd.shape, weight, np.float32))
    tv[~known] = 0
    wv[~known] = 0
    return tv, wv


def from_world(world, truncation=0.3, max_weight=10.0, seed=0):
    """A TSDF of a synth.World: distance to the nearest occupied cell (exact Euclidean
    over a window), random weights, unknown beyond the truncation distance."""
    occ = np.asarray(world.occ, dtype=bool)
    ny, nx = occ.shape
    r = int(np.ceil(truncation / world.res))
    best = np.full(occ.shape, np.inf)
    for dy in range(-r, r + 1):
        for dx in range(-r, r + 1):
            sh = np.zeros_like(occ)
            ys, yd = (slice(max(0, -dy), ny - max(0, dy)), slice(max(0, dy), ny - max(0, -dy)))
            xs, xd = (slice(max(0, -dx), nx - max(0, dx)), slice(max(0, dx), nx - max(0, -dx)))
            sh[yd, xd] = occ[ys, xs]
            dist = np.hypot(dx, dy) * world.res
            best = np.where(sh & (dist < best), dist, best)
    rng = np.random.default_rng(seed)
    sign = np.where(np.asarray(world.free, dtype=bool), 1.0, -1.0)
    tsd = (np.minimum(best, truncation) * sign).astype(np.float32)
    weight = rng.uniform(0.0, max_weight, size=occ.shape).astype(np.float32)
    known = best < truncation
    tv, wv = ob.tsdf_values(truncation, max_weight, tsd, weight)
    tv[~known] = 0
    wv[~known] = 0
    return np.ascontiguousarray(tv), np.ascontiguousarray(wv)
