"""Loop-closure searches over a saved mapping state (.pbstream), 2D: every node against every
finished submap it was not inserted into, through ConstraintBuilder2D on the device.

Which search a (node, submap) pair gets follows PoseGraph2D::ComputeConstraint
(mapping/internal/2d/pose_graph_2d.cc:261-310): same trajectory -> MaybeAddConstraint with
initial_relative_pose = submap global pose^-1 * node global_pose_2d; different trajectories ->
MaybeAddGlobalConstraint (the connectivity-time and sampler gates of a live pose graph have no
counterpart in a saved state: every cross-trajectory pair is searched).  Poses are rebuilt the way
AddSubmapFromProto / AddNodeFromProto do (pose_graph_2d.cc:690-712,740-761): submap 2D pose =
Project2D(global submap pose), node global_pose_2d = Project2D(global pose * gravity_alignment^-1).

    python tools/pbstream_job.py map.pbstream [--check-oracle N] [--devices 0,1]

prints one JSON line (searches, found, seconds, searches/s).  --check-oracle compares the first N
searches with the CPU oracle (test infrastructure; the product path never needs it)."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def _qmul(a, b):
    aw, ax, ay, az = a
    bw, bx, by, bz = b
    return np.array([aw * bw - ax * bx - ay * by - az * bz, aw * bx + ax * bw + ay * bz - az * by,
                     aw * by - ax * bz + ay * bw + az * bx, aw * bz + ax * by - ay * bx + az * bw])


def _yaw(q):
    """transform::GetYaw (transform.h:57-61): the heading of q * UnitX."""
    w, x, y, z = q
    return float(np.arctan2(2.0 * (x * y + w * z), 1.0 - 2.0 * (y * y + z * z)))


def project2d(pose7, right_rotation=None):
    """transform::Project2D(pose * Rotation(right_rotation)) -> (x, y, yaw)."""
    q = np.asarray(pose7[3:], np.float64)
    if right_rotation is not None:
        q = _qmul(q, right_rotation)
    return np.array([pose7[0], pose7[1], _yaw(q)])


def relative2d(a, b):
    """a^-1 * b for Rigid2d (x, y, yaw) (rigid_transform.h:86-97,105-113)."""
    c, s = np.cos(-a[2]), np.sin(-a[2])
    dx, dy = b[0] - a[0], b[1] - a[1]
    return np.array([c * dx - s * dy, s * dx + c * dy, b[2] - a[2]])


def job_from_state(state):
    """-> (grids, submap local poses [n, 3], search_submap, search_full, rel_poses [m, 3], clouds, pairs)."""
    intra = {(c["node_id"], c["submap_id"]) for c in state.constraints if c["tag"] == "INTRA_SUBMAP"}
    submaps = [s for s in state.submaps if not s["is_3d"] and s["finished"] and s["grid"] is not None]
    grids = [s["grid"] for s in submaps]
    # ComputeSubmapPose (constraints/constraint_builder_2d.cc:55-57) = Project2D(submap.local_pose()): the
    # frame the submap's grid is expressed in
    local_poses = np.array([project2d(s["local_pose"]) for s in submaps]).reshape(-1, 3)
    global_poses = [project2d(s["global_pose"]) for s in submaps]
    sub, full, rel, clouds, pairs = [], [], [], [], []
    for n in state.nodes:
        if len(n["filtered_gravity_aligned_point_cloud"]) == 0:
            continue
        g = n["gravity_alignment"]
        node_global_2d = project2d(n["global_pose"], right_rotation=np.array([g[0], -g[1], -g[2], -g[3]]))
        node_id = (n["trajectory_id"], n["node_index"])
        for k, s in enumerate(submaps):
            submap_id = (s["trajectory_id"], s["submap_index"])
            if (node_id, submap_id) in intra:
                continue
            sub.append(k)
            full.append(0 if node_id[0] == submap_id[0] else 1)
            rel.append(relative2d(global_poses[k], node_global_2d))
            clouds.append(n["filtered_gravity_aligned_point_cloud"])
            pairs.append((node_id, submap_id))
    return grids, local_poses, sub, full, np.array(rel).reshape(-1, 3), clouds, pairs


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("pbstream")
    ap.add_argument("--devices", default="0")
    ap.add_argument("--check-oracle", type=int, default=0)
    ap.add_argument("--min-score", type=float, default=0.55)
    ap.add_argument("--global-min-score", type=float, default=0.6)
    ap.add_argument("--linear-window", type=float, default=7.0)
    ap.add_argument("--angular-window", type=float, default=float(np.deg2rad(30.0)))
    ap.add_argument("--depth", type=int, default=7)
    a = ap.parse_args()
    import conftest
    csm = conftest.load_package()
    state = csm.MappingState(a.pbstream)
    grids, poses, sub, full, rel, clouds, pairs = job_from_state(state)
    kw = dict(min_score=a.min_score, global_min_score=a.global_min_score, linear_window=a.linear_window,
              angular_window=a.angular_window, depth=a.depth, max_constraint_distance=15.0)
    devices = tuple(int(d) for d in a.devices.split(","))
    csm.constraint_builder_run(grids, poses, sub[:8], full[:8], rel[:8], clouds[:8], devices=devices, **kw)  # warm-up
    # a node's searches enter as one MaybeAddConstraints burst (whole-node batches)
    secs, found, tf, stats = csm.constraint_builder_run(grids, poses, sub, full, rel, clouds, devices=devices,
                                                        burst=max(1, len(grids)), **kw)
    out = {"pbstream": os.path.basename(a.pbstream), "submaps": len(grids), "nodes": len(state.nodes),
           "searches": int(stats[0]), "found": int(found.sum()), "seconds": secs,
           "searches_per_s": stats[0] / secs if secs > 0 else None, "candidates_scored": int(stats[2])}
    if a.check_oracle:
        import oracle_binding as ob
        k = min(a.check_oracle, len(clouds))
        ogrids = [ob.OracleGrid(g.cells(), g.res, g.max_x, g.max_y) for g in grids]
        _, o_out, _ = ob.constraint_batch(os.cpu_count() or 1, ogrids, poses, sub[:k], full[:k], rel[:k], clouds[:k], **kw)
        bad = [i for i in range(k) if found[i] != o_out[i].found or
               (found[i] and (abs(tf[i][0] - o_out[i].pose_x) > 1e-9 or abs(tf[i][1] - o_out[i].pose_y) > 1e-9))]
        out["oracle_checked"], out["oracle_mismatches"] = k, len(bad)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
