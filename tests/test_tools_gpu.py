"""tools/ptype_b200_bench: the C ABI driven from plain C++ (no Python in the loop).  Its reply
checksum for the config-2 stream equals the oracle's."""
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mesh", [0, 1])
def test_cpp_driver_checksum_matches_oracle(oracle, mesh):
    import workloads as w
    subprocess.run(["make", "-C", str(ROOT / "tools")], check=True, capture_output=True)
    n = 300_000
    out = subprocess.run([str(ROOT / "tools" / "ptype_b200_bench"), str(n), "50", "3", str(mesh)],
                         capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    fields = out.stdout.split()
    got = dict(zip(fields[0::2], fields[1::2]))
    kv, snap = oracle.snapshot_from_spec(w.calculator_registry(mesh=bool(mesh)))
    want = oracle.reply_checksum(snap.dispatch(w.calculator_calls(n), threads=4))
    assert int(got["checksum"]) == want
    assert float(got["calls_per_sec"]) > 1e9


@pytest.mark.parametrize("mesh", [0, 1])
def test_cpp_mirror_end_to_end(mesh):
    """The `mirror` mode drives cluster::Client::CallBatch (C++ mirror of the reference's Client)
    on page-locked batches and through plain arrays; it checks every value itself."""
    subprocess.run(["make", "-C", str(ROOT / "tools")], check=True, capture_output=True)
    out = subprocess.run([str(ROOT / "tools" / "ptype_b200_bench"), "mirror", "200000", "10", "0", str(mesh)],
                         capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    fields = out.stdout.split()
    got = dict(zip(fields[0::2], fields[1::2]))
    assert got["parity"] == "ok" and float(got["calls"]) == 2_000_000 and float(got["seconds"]) > 0
    assert float(got["array_calls"]) == 2_000_000 and float(got["array_seconds"]) > 0
