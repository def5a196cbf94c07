import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on a B200)")
    # The package refuses to import without its CUDA library (no CPU fallback).  nvcc
    # cross-compiles it without a GPU, so build it here if a fresh checkout has not yet.
    lib = ROOT / "ptype_b200" / "libptype_b200.so"
    if not lib.exists():
        import subprocess
        subprocess.run(["make", "-C", str(ROOT / "ptype_b200" / "csrc")], check=True)


@pytest.fixture(scope="session")
def oracle():
    from oracle import binding
    binding.build()
    return binding


@pytest.fixture(scope="session")
def ctx():
    """One device context on cuda:0 for the whole GPU test session."""
    import ptype_b200 as p
    c = p.Context([0])
    yield c
    c.close()
