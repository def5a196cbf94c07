"""Runs the C++ host-mirror tests (tests/cpp/cluster_test.cpp): the reference's own
rpc_test.go / registry_test.go cases restated against ptype_b200/host/cluster.hpp."""
import subprocess
from pathlib import Path

import pytest

CPP = Path(__file__).parent / "cpp"


def _build():
    subprocess.run(["make", "-C", str(CPP)], check=True, capture_output=True)
    return CPP / "cluster_test"


def test_registry_mirror_cpu():
    exe = _build()
    out = subprocess.run([str(exe), "cpu"], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "6 tests, 0 failed" in out.stdout


@pytest.mark.gpu
def test_client_and_balancer_mirror_gpu():
    exe = _build()
    out = subprocess.run([str(exe), "gpu"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "23 tests, 0 failed" in out.stdout
