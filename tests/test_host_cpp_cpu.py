"""The host-side pieces next to the path as C++ sees them: tests/cpp/io_tests.cc is compiled against
the REPLACEMENT headers under cartographer-1_b200/host, linked with libcsm_host.so, and runs the bodies
of the reference's own unit tests (io/proto_stream_test.cc, io/proto_stream_deserializer_test.cc,
sensor/compressed_point_cloud_test.cc, sensor/internal/voxel_filter_test.cc).  No GPU needed; the
compressing constructor is also held word for word to the reference's own source in oracle/_ref."""
import os
import subprocess

import numpy as np
import pytest

import ref_binding as rb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "cartographer-1_b200")


def test_reference_io_and_sensor_unit_tests_against_the_replacement_headers(csm, tmp_path):
    csm.host()  # the libraries must be built
    exe = str(tmp_path / "io_tests")
    subprocess.check_call(
        ["g++", "-O1", "-std=c++17", "-ffp-contract=off", "-Wall", "-I" + os.path.join(ROOT, "include"),
         "-I" + os.path.join(PKG, "host"), "-o", exe, os.path.join(ROOT, "tests", "cpp", "io_tests.cc"),
         "-L" + PKG, "-lcsm_host", "-lcsm_b200", "-Wl,-rpath," + PKG])
    run = subprocess.run([exe, str(tmp_path)], capture_output=True, text=True, timeout=300)
    assert run.returncode == 0, run.stdout + run.stderr
    assert "0 failures" in run.stdout, run.stdout


@pytest.mark.skipif(not rb.available(), reason="oracle/_ref is not built and /root/reference is absent")
@pytest.mark.parametrize("n,spread", [(0, 1.0), (1, 0.3), (500, 0.9), (3000, 30.0), (2000, 70.0), (400, 4000.0),
                                      (5000, 8000.0)])
def test_compressed_point_cloud_words_equal_the_reference_encoder(csm, n, spread):
    """Same blocks, same block order (the reference walks a HybridGridBase that grows with the cloud's
    extent), same packed points: the words a node record would carry."""
    rng = np.random.default_rng(300 + n)
    pts = rng.uniform(-spread, spread, (n, 3)).astype(np.float32)
    assert csm.compress_point_cloud(pts).tobytes() == rb.compress_point_cloud(pts).tobytes()
