#!/usr/bin/env python
"""Builds one 2000 x 2000 depth-7 stack a few times (for ncu: per-launch durations of the build kernels)."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from __graft_entry__ import _load_package  # noqa: E402
import synth  # noqa: E402

import torch  # noqa: E402

size = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
pkg = _load_package()
ctx = pkg.Context(0)
w = synth.World(size, size, seed=4000 + size)
g = pkg.Grid(w.cells, w.res, w.max_x, w.max_y)
m = pkg.Matcher(ctx, g, 7.0, 0.5, 7)
cells = torch.from_numpy(w.cells.view(np.int16)).cuda()
for _ in range(3):
    pkg.host().csmh_matcher_rebuild_dev(ctx.h, m.h, C.c_void_p(cells.data_ptr()))
torch.cuda.synchronize()
