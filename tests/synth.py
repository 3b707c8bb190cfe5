"""Deterministic synthetic worlds and scans for tests and bench.

Not product code: it only manufactures inputs in the reference's data layout --
a uint16 correspondence-cost cell array (flat = num_x_cells * y + x,
/root/reference/cartographer/mapping/2d/grid_2d.h:113-116, value semantics of
mapping/probability_values.h:85-88) plus fp32 (x, y, 0) point clouds in the sensor
frame (TrajectoryNode::Data::filtered_gravity_aligned_point_cloud,
mapping/trajectory_node.h:54).

World: rectilinear "office" occupancy at `res` m/cell: outer walls, K = area/25 m^2
random axis-aligned wall segments (length U[2,8] m, thickness 2 cells), doorways cut
with p = 0.3; wall cells get probability U[0.65,0.9], observed free space
U[0.12,0.35], the outside of the outer wall and a 15 % blocky mask stay unknown (0).
Scans: 360 beams, 1 degree spacing, max range 10 m, hits only, sigma = 0.01 m range
noise.
"""
import math

import numpy as np


def probability_to_value(p):
    """CorrespondenceCostToValue(1 - p) in fp32 (probability_values.h:32-46,85-88)."""
    p = np.asarray(p, dtype=np.float32)
    lo = np.float32(1.0) - (np.float32(1.0) - np.float32(0.1))  # kMinCorrespondenceCost
    hi = np.float32(1.0) - np.float32(0.1)                      # kMaxCorrespondenceCost
    c = np.clip(np.float32(1.0) - p, lo, hi).astype(np.float32)
    scaled = ((c - lo) * (np.float32(32766.0) / (hi - lo))).astype(np.float32)
    return (np.floor(scaled.astype(np.float64) + 0.5).astype(np.int64) + 1).astype(np.uint16)


class World:
    """A synthetic submap: cells (uint16 [ny, nx]), MapLimits and the wall mask."""

    def __init__(self, nx, ny, seed, res=0.05, unknown_fraction=0.15):
        rng = np.random.default_rng(seed)
        self.nx, self.ny, self.res = nx, ny, res
        # Submap centred on the origin: MapLimits.max = (ny*res/2, nx*res/2) because
        # cell.x runs along world -y and cell.y along world -x (map_limits.h:69-76).
        self.max_x = ny * res / 2.0
        self.max_y = nx * res / 2.0
        occ = np.zeros((ny, nx), dtype=bool)
        m = 6  # outer wall inset, cells
        occ[m:m + 2, m:nx - m] = True
        occ[ny - m - 2:ny - m, m:nx - m] = True
        occ[m:ny - m, m:m + 2] = True
        occ[m:ny - m, nx - m - 2:nx - m] = True
        area = nx * ny * res * res
        for _ in range(max(1, int(area / 25.0))):
            length = int(rng.uniform(2.0, 8.0) / res)
            horizontal = rng.random() < 0.5
            if horizontal:
                y0 = int(rng.integers(m + 2, max(m + 3, ny - m - 4)))
                x0 = int(rng.integers(m + 2, max(m + 3, nx - m - 2 - min(length, nx - 2 * m - 6))))
                x1 = min(x0 + length, nx - m - 2)
                occ[y0:y0 + 2, x0:x1] = True
                if rng.random() < 0.3 and x1 - x0 > 30:
                    d = int(rng.integers(x0 + 5, x1 - 25))
                    occ[y0:y0 + 2, d:d + 20] = False
            else:
                x0 = int(rng.integers(m + 2, max(m + 3, nx - m - 4)))
                y0 = int(rng.integers(m + 2, max(m + 3, ny - m - 2 - min(length, ny - 2 * m - 6))))
                y1 = min(y0 + length, ny - m - 2)
                occ[y0:y1, x0:x0 + 2] = True
                if rng.random() < 0.3 and y1 - y0 > 30:
                    d = int(rng.integers(y0 + 5, y1 - 25))
                    occ[d:d + 20, x0:x0 + 2] = False
        self.occ = occ
        prob = np.zeros((ny, nx), dtype=np.float32)
        inside = np.zeros((ny, nx), dtype=bool)
        inside[m:ny - m, m:nx - m] = True
        free = inside & ~occ
        prob[free] = rng.uniform(0.12, 0.35, size=int(free.sum())).astype(np.float32)
        prob[occ] = rng.uniform(0.65, 0.9, size=int(occ.sum())).astype(np.float32)
        known = inside.copy()
        # 15 % blocky unobserved mask (8x8-cell blocks).
        by, bx = (ny + 7) // 8, (nx + 7) // 8
        blocks = rng.random((by, bx)) < unknown_fraction
        mask = np.kron(blocks, np.ones((8, 8), dtype=bool))[:ny, :nx]
        known &= ~mask
        cells = np.zeros((ny, nx), dtype=np.uint16)
        cells[known] = probability_to_value(prob[known])
        self.cells = np.ascontiguousarray(cells)
        self.free = free

    # world <-> continuous cell coordinates (cell centre of index i is at i + 0.5)
    def world_to_cell(self, px, py):
        return (self.max_y - py) / self.res, (self.max_x - px) / self.res  # (cx, cy)

    def cell_to_world(self, cx, cy):
        return self.max_x - cy * self.res, self.max_y - cx * self.res  # (px, py)

    def random_free_pose(self, rng):
        ys, xs = np.nonzero(self.free)
        k = int(rng.integers(0, len(xs)))
        px, py = self.cell_to_world(xs[k] + 0.5, ys[k] + 0.5)
        return px, py, float(rng.uniform(-math.pi, math.pi))

    def scan(self, pose, rng, beams=360, max_range=10.0, sigma=0.01):
        """Ray-casts `beams` rays from world pose (x, y, yaw); returns fp32 [n, 3]
        hit points in the sensor frame."""
        px, py, yaw = pose
        ang = yaw + np.deg2rad(np.arange(beams, dtype=np.float64) * (360.0 / beams))
        step = 0.5 * self.res
        n_steps = int(max_range / step)
        r = (np.arange(1, n_steps + 1, dtype=np.float64) * step)[None, :]
        wx = px + r * np.cos(ang)[:, None]
        wy = py + r * np.sin(ang)[:, None]
        cx = np.floor((self.max_y - wy) / self.res).astype(np.int64)
        cy = np.floor((self.max_x - wx) / self.res).astype(np.int64)
        inb = (cx >= 0) & (cy >= 0) & (cx < self.nx) & (cy < self.ny)
        hit = np.zeros_like(inb)
        hit[inb] = self.occ[cy[inb], cx[inb]]
        any_hit = hit.any(axis=1)
        first = hit.argmax(axis=1)
        rng_hit = (first + 1) * step + rng.normal(0.0, sigma, size=beams)
        rel = ang - yaw
        pts = np.stack([rng_hit * np.cos(rel), rng_hit * np.sin(rel), np.zeros(beams)], axis=1)
        return np.ascontiguousarray(pts[any_hit].astype(np.float32))


def make_scan(world, seed, **kw):
    rng = np.random.default_rng(seed)
    pose = world.random_free_pose(rng)
    pts = world.scan(pose, rng, **kw)
    return pose, pts
