"""The C++ host-side mirror of the reference API (host/pcseg_host.h, namespace
mobility) above the C ABI.  CPU: the test binary compiles and links against the
library.  GPU: tests/host/clustering_host_test.cc runs the known-answer test of
ps/clustering_test.cc:115-151 and the other entry points through the shim."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "tests", "host", "clustering_host_test")


def test_host_shim_compiles_and_links():
    subprocess.check_call(["make", "-C", ROOT, "host_test"])
    assert os.path.exists(BIN)
    out = subprocess.run(["ldd", BIN], capture_output=True, text=True).stdout
    assert "libpcseg_b200.so" in out and "not found" not in out.split("libpcseg_b200.so")[1].split("\n")[0]


def test_shim_compiles_beside_the_reference_headers(tmp_path):
    """the mirror lives in mobility::b200, so one translation unit can include it together with
    the reference's planar_region.h / cluster_region.h / algorithms.h (built against the oracle's shims)."""
    ref = "/root/reference"
    if not os.path.isdir(os.path.join(ref, "pointcloud_segmentation")):
        pytest.skip("reference tree not present on this machine")
    src = os.path.join(ROOT, "tests", "host", "include_beside_reference.cc")
    subprocess.check_call([os.environ.get("CXX", "g++"), "-std=c++17", "-fsyntax-only", "-Wno-return-type", "-Wno-unused",
                           "-I" + os.path.join(ROOT, "include"),
                           "-I" + os.path.join(ROOT, "x-edr-pointcloud-segmentation_b200", "host"),
                           "-I" + os.path.join(ROOT, "oracle", "_ref", "patched"),
                           "-I" + os.path.join(ROOT, "oracle", "shims"), "-I" + ref, src])


@pytest.mark.gpu
def test_host_shim_known_answers():
    assert os.path.exists(BIN), "run `make host_test` (build() does) before the GPU tests"
    r = subprocess.run([BIN], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "PASS clustering_host_test" in r.stdout
