"""CPU-only checks: the oracle against the reference's own known-answer tests, and
the C-ABI library's exported surface (no compute calls without a GPU)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import oracle_binding as ob
import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_ported_reference_tests_pass():
    ob.build_oracle()
    out = subprocess.run([os.path.join(ROOT, "oracle", "oracle_tests")], capture_output=True,
                         text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "0 failures" in out.stdout


def test_oracle_finds_true_pose_in_synthetic_world():
    w = synth.World(200, 200, seed=11)
    g = ob.OracleGrid(w.cells, w.res, w.max_x, w.max_y)
    m = ob.OracleFCSM2D(g, 7.0, np.deg2rad(30.0), 5)
    pose, pts = synth.make_scan(w, 5, max_range=6.0)
    r = m.match(pts, 0.5)
    assert r.found == 1
    assert abs(r.pose_x - pose[0]) < 0.15 and abs(r.pose_y - pose[1]) < 0.15


def test_abi_exports_every_declared_symbol(csm):
    header = open(os.path.join(ROOT, "include", "csm_b200.h")).read()
    declared = set(re.findall(r"\b(csm_[a-z0-9_]+)\s*\(", header))
    lib = C.CDLL(csm.LIB_ABI)
    missing = [s for s in sorted(declared) if not hasattr(lib, s)]
    assert not missing, missing
    assert set(csm.ABI_SYMBOLS) <= declared


def test_no_device_means_loud_failure(csm):
    lib = csm.abi()
    if lib.csm_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(RuntimeError):
        csm.Context(0)
    ctx = C.c_void_p()
    assert lib.csm_ctx_create(0, C.byref(ctx)) != 0
    assert b"no CPU fallback" in lib.csm_last_error()
