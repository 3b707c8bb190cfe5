"""The drop-in boundary as C++ sees it: tests/cpp/boundary_tests.cc is compiled
against the REPLACEMENT headers under cartographer-1_b200/host and linked with libcsm_host.so /
libcsm_b200.so; it runs the reference's own unit-test inputs and expectations for the public
symbols of those headers (PrecomputationGrid2D, PrecomputationGridStack2D / 3D,
CreateFastCorrelativeScanMatcherOptions2D / 3D, GenerateRotatedScans, DiscretizeScans,
RealTimeCorrelativeScanMatcher2D::ScoreCandidates, Match / MatchFullSubmap) on the device."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "cartographer-1_b200")


@pytest.mark.gpu
def test_reference_unit_tests_against_the_replacement_headers(csm, tmp_path):
    csm.host()  # the libraries must be built (no fallback)
    exe = str(tmp_path / "boundary_tests")
    subprocess.check_call(
        ["g++", "-O1", "-std=c++17", "-ffp-contract=off", "-Wall", "-I" + os.path.join(ROOT, "include"),
         "-I" + os.path.join(PKG, "host"), "-o", exe, os.path.join(ROOT, "tests", "cpp", "boundary_tests.cc"),
         "-L" + PKG, "-lcsm_host", "-lcsm_b200", "-Wl,-rpath," + PKG])
    run = subprocess.run([exe], capture_output=True, text=True, timeout=600)
    assert run.returncode == 0, run.stdout + run.stderr
    assert "0 failures" in run.stdout, run.stdout
