#!/usr/bin/env python
"""Writes tests/golden/fullsize_golden.json: what the CPU oracle returns at BASELINE.json's
full sizes, so the GPU tests can hold the CUDA path to it without paying the oracle's minutes
on the GPU box (a C5 MatchFullSubmap alone takes the restatement ~90 s on one core).

  C4  SHA-1 of every level of the 2000 x 2000, depth-7 stack (PrecomputationGrid2D with the
      literal SlidingWindowMaximum, FCSM2D.cc:41-161)
  C3  MatchFullSubmap winners on a 1000 x 1000 submap (2 scans from inside, 2 from elsewhere)
  C5  2 Match + 1 MatchFullSubmap on the 40 x 40 x 6 m room at 0.1 / 0.45 m, depth 8,
      full_resolution_depth 3

Inputs are regenerated from the seeds by tests/fullsize_cases.py (the same code the tests
run); their hashes are stored so that a drifting generator fails loudly instead of silently
comparing different inputs.  Run from the repository root: python tools/make_golden_fullsize.py
"""
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import fullsize_cases as fc  # noqa: E402
import golden_util as gu  # noqa: E402
import oracle_binding as ob  # noqa: E402


def main():
    doc = {"generator": "tools/make_golden_fullsize.py", "oracle": "oracle/csm_oracle.h, csm_oracle3d.h"}
    t0 = time.time()
    w = fc.c4_world()
    om = ob.OracleFCSM2D(ob.OracleGrid(w.cells, w.res, w.max_x, w.max_y), 7.0, 0.5, 7)
    doc["c4"] = {"cells_sha1": gu.sha(w.cells),
                 "level_sha1": [gu.sha(om.level(i)) for i in range(7)],
                 "level_shape": [list(om.level(i).shape) for i in range(7)]}
    print("C4 %.1f s" % (time.time() - t0))

    t0 = time.time()
    w, clouds = fc.c3_case()
    om = ob.OracleFCSM2D(ob.OracleGrid(w.cells, w.res, w.max_x, w.max_y), fc.LINEAR, fc.ANGULAR, 7)
    with ThreadPoolExecutor(len(clouds)) as ex:
        res = list(ex.map(lambda c: om.match(c, fc.C3_MIN_SCORE), clouds))
    doc["c3"] = {"cells_sha1": gu.sha(w.cells), "clouds_sha1": [gu.sha(c) for c in clouds],
                 "results": [{"found": int(r.found), "scan_index": r.scan_index,
                              "x_index_offset": r.x_index_offset, "y_index_offset": r.y_index_offset,
                              "score_bits": gu.score_bits(r.score), "pose": [r.pose_x, r.pose_y, r.pose_theta]}
                             for r in res]}
    print("C3 %.1f s, found %s" % (time.time() - t0, [int(r.found) for r in res]))

    t0 = time.time()
    w3, vox, nodes = fc.c5_case()
    hi_xyz, hi_val, lo_xyz, lo_val = vox
    om3 = ob.OracleFCSM3D(ob.OracleGrid3D(0.1, hi_xyz, hi_val), ob.OracleGrid3D(0.45, lo_xyz, lo_val),
                          w3.submap_histogram(), np.array(fc.C5_OPT))

    def one(nd):
        hi, lo, hist, grav, init, full = nd
        return om3.match(hi, lo, hist, fc.C5_MIN_SCORE, init, ob.IDENTITY_POSE7, grav, bool(full))

    with ThreadPoolExecutor(len(nodes)) as ex:
        res3 = list(ex.map(one, nodes))
    doc["c5"] = {"voxels_sha1": gu.sha(hi_xyz) + gu.sha(hi_val), "num_voxels": int(len(hi_val)),
                 "clouds_sha1": [gu.sha(nd[0]) for nd in nodes],
                 "results": [{"found": int(bool(r.found)), "scan_index": r.scan_index, "offset": list(r.offset),
                              "score_bits": gu.score_bits(r.score),
                              "low_resolution_score_bits": gu.score_bits(r.low_resolution_score),
                              "rotational_score_bits": gu.score_bits(r.rotational_score),
                              "pose": [float(x) for x in r.pose]} for r in res3]}
    print("C5 %.1f s, found %s" % (time.time() - t0, [int(bool(r.found)) for r in res3]))
    path = os.path.join(ROOT, "tests", "golden", "fullsize_golden.json")
    with open(path, "w") as f:
        json.dump(doc, f, indent=1)
    print("wrote", path)


if __name__ == "__main__":
    main()
