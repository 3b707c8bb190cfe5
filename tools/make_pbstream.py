"""Writes a synthetic saved mapping state (.pbstream) of C3-shaped submaps with Python's gzip and the
protobuf runtime (tests/proto_fixtures.py) -- no product code -- for tools/pbstream_job.py:

    python tools/make_pbstream.py out.pbstream --submaps 8 --nodes-per-submap 25 --size 1000

Submaps alternate between two trajectories, so a node is searched locally against the submaps of its
own trajectory and globally (MatchFullSubmap) against the others."""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import proto_fixtures as pf  # noqa: E402
import synth  # noqa: E402


def pose7(x, y, yaw):
    return (x, y, 0.0, np.cos(yaw / 2), 0.0, 0.0, np.sin(yaw / 2))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("out")
    ap.add_argument("--submaps", type=int, default=8)
    ap.add_argument("--nodes-per-submap", type=int, default=25)
    ap.add_argument("--size", type=int, default=1000)
    ap.add_argument("--seed", type=int, default=3)
    a = ap.parse_args()
    rng = np.random.default_rng(a.seed)
    submaps, nodes, constraints = [], [], []
    for k in range(a.submaps):
        w = synth.World(a.size, a.size, seed=3000 + k)
        local = (rng.uniform(-2, 2), rng.uniform(-2, 2), rng.uniform(-1.5, 1.5))
        submaps.append({"id": (k % 2, k // 2), "grid_bytes": pf.grid2d_bytes(w.cells, w.res, w.max_x, w.max_y),
                        "local_pose": pose7(*local), "global_pose": pose7(*local), "num_range_data": 90,
                        "finished": True})
        for j in range(a.nodes_per_submap):
            pose, pts = synth.make_scan(w, 3000 * 1000 + 100 * k + j)
            c, s = np.cos(-local[2]), np.sin(-local[2])
            dx, dy = pose[0] - local[0], pose[1] - local[1]
            rel = (c * dx - s * dy + rng.uniform(-0.5, 0.5), s * dx + c * dy + rng.uniform(-0.5, 0.5),
                   pose[2] - local[2] + rng.uniform(-0.2, 0.2))
            cg, sg = np.cos(local[2]), np.sin(local[2])
            glob = (local[0] + cg * rel[0] - sg * rel[1], local[1] + sg * rel[0] + cg * rel[1], local[2] + rel[2])
            node = {"id": (k % 2, len(nodes)), "timestamp": 1000 + len(nodes), "gravity_alignment": (1, 0, 0, 0),
                    "local_pose": pose7(*pose), "global_pose": pose7(*glob), "filtered": pts}
            nodes.append(node)
            constraints.append({"submap_id": submaps[k]["id"], "node_id": node["id"], "relative_pose": pose7(*rel),
                                "translation_weight": 5e2, "rotation_weight": 1.6e3, "tag": 0})
    blob = pf.pbstream_bytes(pf.mapping_state_messages(submaps, nodes, constraints))
    with open(a.out, "wb") as f:
        f.write(blob)
    print("%s: %d submaps of %d^2, %d nodes, %.1f MB" % (a.out, len(submaps), a.size, len(nodes), len(blob) / 1e6))


if __name__ == "__main__":
    main()
