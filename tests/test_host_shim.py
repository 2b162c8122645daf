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


@pytest.mark.gpu
def test_host_shim_known_answers():
    assert os.path.exists(BIN), "run `make host_test` (build() does) before the GPU tests"
    r = subprocess.run([BIN], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "PASS clustering_host_test" in r.stdout
