"""Synthetic 3D worlds for the FastCorrelativeScanMatcher3D parity tests and the C5
bench: an axis-aligned room with boxes, voxelised into a high- and a low-resolution
HybridGrid voxel list, plus node clouds and rotational histograms.  Input data only;
nothing here is part of the product or of the oracle."""
import numpy as np


def probability_to_value(p):
    """probability_values.h:32-46 on fp32 (input generation only)."""
    p = np.clip(np.asarray(p, np.float32), np.float32(0.1), np.float32(0.9))
    scale = np.float32(32766.0) / (np.float32(0.9) - np.float32(0.1))
    return (np.round((p - np.float32(0.1)) * scale).astype(np.int32) + 1).astype(np.uint16)


def yaw_quat(theta):
    return np.array([np.cos(0.5 * theta), 0.0, 0.0, np.sin(0.5 * theta)])


def quat_mul(a, b):
    aw, ax, ay, az = a
    bw, bx, by, bz = b
    return np.array([aw * bw - ax * bx - ay * by - az * bz, aw * bx + ax * bw + ay * bz - az * by,
                     aw * by + ay * bw + az * bx - ax * bz, aw * bz + az * bw + ax * by - ay * bx])


def quat_rotate(q, pts):
    w, x, y, z = q
    r = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                  [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                  [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])
    return pts @ r.T


def quat_conj(q):
    return np.array([q[0], -q[1], -q[2], -q[3]])


def pose7(t, q):
    return np.concatenate([np.asarray(t, np.float64), np.asarray(q, np.float64)])


class World3D:
    def __init__(self, seed, size_xy=12.0, height=3.0, res=0.1, low_res=0.45, n_boxes=5,
                 samples_per_m2=400.0):
        self.rng = np.random.default_rng(seed)
        self.res, self.low_res = np.float32(res), np.float32(low_res)
        self.size_xy, self.height = size_xy, height
        rng = self.rng
        h = size_xy / 2.0
        pts, tang = [], []

        def plane(origin, u, v, tangent):
            area = np.linalg.norm(u) * np.linalg.norm(v)
            n = max(16, int(area * samples_per_m2))
            a, b = rng.random(n), rng.random(n)
            pts.append(origin + a[:, None] * u + b[:, None] * v)
            tang.append(np.full(n, tangent))

        z0 = -1.0
        # walls (tangent direction in the xy plane), floor and ceiling (no xy tangent)
        plane(np.array([-h, -h, z0]), np.array([2 * h, 0, 0.0]), np.array([0, 0, height]), 0.0)
        plane(np.array([-h, h, z0]), np.array([2 * h, 0, 0.0]), np.array([0, 0, height]), 0.0)
        plane(np.array([-h, -h, z0]), np.array([0, 2 * h, 0.0]), np.array([0, 0, height]), np.pi / 2)
        plane(np.array([h, -h, z0]), np.array([0, 2 * h, 0.0]), np.array([0, 0, height]), np.pi / 2)
        plane(np.array([-h, -h, z0]), np.array([2 * h, 0, 0.0]), np.array([0, 2 * h, 0.0]), np.nan)
        plane(np.array([-h, -h, z0 + height]), np.array([2 * h, 0, 0.0]), np.array([0, 2 * h, 0.0]), np.nan)
        for _ in range(n_boxes):
            c = rng.uniform(-h * 0.8, h * 0.8, 2)
            s = rng.uniform(0.4, 1.6, 2)
            bh = rng.uniform(0.5, height * 0.8)
            x0, x1, y0, y1 = c[0] - s[0] / 2, c[0] + s[0] / 2, c[1] - s[1] / 2, c[1] + s[1] / 2
            plane(np.array([x0, y0, z0]), np.array([x1 - x0, 0, 0.0]), np.array([0, 0, bh]), 0.0)
            plane(np.array([x0, y1, z0]), np.array([x1 - x0, 0, 0.0]), np.array([0, 0, bh]), 0.0)
            plane(np.array([x0, y0, z0]), np.array([0, y1 - y0, 0.0]), np.array([0, 0, bh]), np.pi / 2)
            plane(np.array([x1, y0, z0]), np.array([0, y1 - y0, 0.0]), np.array([0, 0, bh]), np.pi / 2)
            plane(np.array([x0, y0, z0 + bh]), np.array([x1 - x0, 0, 0.0]), np.array([0, y1 - y0, 0.0]), np.nan)
        self.points = np.concatenate(pts)
        self.tangent = np.concatenate(tang)

    def voxels(self, res, p_lo=0.55, p_hi=0.9, free_fraction=0.5):
        """Voxel list (xyz int32, uint16 values): hit voxels with random occupied
        probabilities plus a shell of low-probability (free-space) voxels."""
        idx = np.round(self.points.astype(np.float32) / np.float32(res)).astype(np.int32)
        occ = np.unique(idx, axis=0)
        prob = self.rng.uniform(p_lo, p_hi, len(occ))
        shell = np.unique(np.concatenate([occ + d for d in ([1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0],
                                                          [0, 0, 1], [0, 0, -1])]), axis=0)
        keep = self.rng.random(len(shell)) < free_fraction
        shell = shell[keep]
        # drop shell voxels that coincide with occupied ones
        key = lambda a: (a[:, 0].astype(np.int64) << 42) + (a[:, 1].astype(np.int64) << 21) + a[:, 2]
        shell = shell[~np.isin(key(shell + (1 << 20)), key(occ + (1 << 20)))]
        sprob = self.rng.uniform(0.12, 0.4, len(shell))
        xyz = np.concatenate([occ, shell]).astype(np.int32)
        values = probability_to_value(np.concatenate([prob, sprob]))
        return np.ascontiguousarray(xyz), np.ascontiguousarray(values)

    def histogram(self, tangent_world, yaw_frame, buckets):
        """Bucketised surface-tangent angles (mod pi) seen from a frame with the given yaw."""
        t = tangent_world[~np.isnan(tangent_world)]
        a = np.mod(t - yaw_frame, np.pi) / np.pi * buckets
        lo = np.floor(a).astype(int) % buckets
        frac = a - np.floor(a)
        hist = np.zeros(buckets)
        np.add.at(hist, lo, 1.0 - frac)
        np.add.at(hist, (lo + 1) % buckets, frac)
        return hist.astype(np.float32)

    def submap_histogram(self, buckets=120):
        return self.histogram(self.tangent, 0.0, buckets)

    def node(self, position, yaw, max_range=6.0, n_high=300, n_low=80, noise=0.01, buckets=120,
             gravity_yaw=0.0, tilt=(0.0, 0.0)):
        """Returns (high points, low points, histogram, gravity quaternion) of a node at
        the given true pose in the world (= submap) frame."""
        position = np.asarray(position, np.float64)
        d = np.linalg.norm(self.points - position, axis=1)
        near = np.nonzero(d < max_range)[0]
        pick = self.rng.choice(near, size=min(n_high, len(near)), replace=False)
        q_node = yaw_quat(yaw)
        if tilt != (0.0, 0.0):
            q_node = quat_mul(q_node, np.array([np.cos(0.5 * tilt[0]), np.sin(0.5 * tilt[0]), 0.0, 0.0]))
            q_node = quat_mul(q_node, np.array([np.cos(0.5 * tilt[1]), 0.0, np.sin(0.5 * tilt[1]), 0.0]))
        world = self.points[pick] + self.rng.normal(0.0, noise, (len(pick), 3))
        local = quat_rotate(quat_conj(q_node), world - position)
        low_pick = self.rng.choice(near, size=min(n_low, len(near)), replace=False)
        low_world = self.points[low_pick] + self.rng.normal(0.0, noise, (len(low_pick), 3))
        low_local = quat_rotate(quat_conj(q_node), low_world - position)
        gravity = yaw_quat(gravity_yaw)
        hist = self.histogram(self.tangent[pick], yaw - gravity_yaw, buckets)
        return (local.astype(np.float32), low_local.astype(np.float32), hist, gravity,
                pose7(position, q_node))
