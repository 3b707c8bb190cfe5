"""Inputs of the full-size parity tests (BASELINE.json configs C3, C4, C5), regenerated from
fixed seeds; shared by tests/test_gpu_fullsize.py and tools/make_golden_fullsize.py."""
import numpy as np

import synth
import synth3d

LINEAR, ANGULAR = 7.0, float(np.deg2rad(30.0))
C3_MIN_SCORE = 0.6
C5_OPT = [8, 3, 0.77, 0.55, 5.0, 1.0, float(np.radians(15.0))]  # pose_graph.lua 3D defaults
C5_MIN_SCORE = 0.55


def c4_world():
    return synth.World(2000, 2000, seed=4000)


def c3_case():
    """One 1000 x 1000 submap; scans 0, 1 from inside it, scans 2, 3 from another world."""
    w = synth.World(1000, 1000, seed=3000)
    decoy = synth.World(1000, 1000, seed=3999)
    clouds = [synth.make_scan(w, 3000)[1], synth.make_scan(w, 3006)[1],
              synth.make_scan(decoy, 3001)[1], synth.make_scan(decoy, 3002)[1]]
    return w, clouds


def c5_case():
    """40 x 40 x 6 m room at 0.1 / 0.45 m; nodes = (high cloud, low cloud, histogram, gravity,
    initial pose7 as float32, full_submap)."""
    w = synth3d.World3D(9, size_xy=40.0, height=6.0, n_boxes=40, samples_per_m2=300.0)
    hi_xyz, hi_val = w.voxels(w.res)
    lo_xyz, lo_val = w.voxels(w.low_res)
    rng = np.random.default_rng(10)
    nodes = []
    for k in range(3):
        full = k == 2
        pos = np.append(rng.uniform(-18, 18, 2), rng.uniform(-0.3, 1.0))
        yaw = rng.uniform(-np.pi, np.pi)
        hi, lo, hist, grav, true = w.node(pos, yaw, max_range=20.0, n_high=1500, n_low=300,
                                          gravity_yaw=rng.uniform(-0.3, 0.3))
        init = true.copy()
        if full:
            init[:3] = 0.0
        else:
            init[:3] += rng.uniform(-3.0, 3.0, 3) * [1, 1, 0.2]
            init[3:] = synth3d.quat_mul(synth3d.yaw_quat(rng.uniform(-0.15, 0.15)), true[3:])
        nodes.append((hi, lo, hist, grav, init.astype(np.float32), full))
    return w, (hi_xyz, hi_val, lo_xyz, lo_val), nodes
