"""The driver's bench contract, exercised on the CPU: the reference arm prints one JSON line with
the agreed keys, and the `ours` arm refuses to run without a GPU (no CPU fallback)."""
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def test_reference_arm_json_line():
    out = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--steps", "2",
                          "--warmup", "1", "--batch", "65536"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["metric"] == "rpc_calls_per_sec_batched_dispatch"
    assert line["unit"] == "calls/s" and line["higher_is_better"] is True and line["value"] > 0
    assert line["steps"] == 2 and line["warmup"] == 1 and line["vs_baseline"] is None
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["cpu_baseline"]["value"] == line["value"]
    assert line["e2e"] == {"value": line["value"], "unit": "calls/s", "h2d_bytes_per_step": 0,
                           "d2h_bytes_per_step": 0}
    assert "no Go toolchain" in line["note"]
    # the CPU arm depends on nothing of the product: only the oracle's library is mapped
    assert line["in_tree_libs_loaded"] == ["oracle/_build/libptype_oracle.so"]
    assert set(line["config"]) == {"workload", "batch_calls_per_gpu", "resident_batches", "l2"}
    # config 1: the reference's real path restated over loopback TCP
    assert line["loopback_1_caller"]["value"] > 0 and line["loopback_all_threads"]["value"] > 0
    assert 30 < line["loopback_1_caller"]["bytes_per_call_out"] < 120


def test_ours_arm_needs_a_gpu():
    import torch
    if torch.cuda.is_available():
        return
    out = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--steps", "1", "--warmup", "1"],
                         capture_output=True, text=True, timeout=300)
    assert out.returncode != 0 and "no CPU fallback" in (out.stderr + out.stdout)
