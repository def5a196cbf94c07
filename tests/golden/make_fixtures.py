"""Regenerates tests/golden/dispatch_fixtures.npz: small request batches for three
registries and the oracle's replies to them.  The reference is Go and cannot run in
this image (no Go toolchain), so these are ORACLE outputs frozen at a commit where
the oracle passed every reference-derived vector in reference_vectors.json — they
guard the oracle and the CUDA path against drifting together.

    python tests/golden/make_fixtures.py
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

from oracle import binding as ob          # noqa: E402
import workloads as w          # noqa: E402

out = {}
cases = {
    "calc3": (w.calculator_registry(), lambda: np.concatenate(
        [w.calculator_calls(2048), w.edge_case_calls(["calculator"], n_random=1024)])),
    "calcmesh": (w.calculator_registry(mesh=True), lambda: np.concatenate(
        [w.calculator_calls(2048, seq_base=(1 << 32) - 1000), w.edge_case_calls(["calculator"], n_random=512)])),
    "optimus": (w.optimus_registry(), lambda: np.concatenate(
        [w.optimus_calls(2048, t_lo=50, t_hi=5000), w.edge_case_calls(["prime_worker"], n_random=512)])),
}
for name, (spec, gen) in cases.items():
    kv, snap = ob.snapshot_from_spec(spec)
    req = gen()
    rep = snap.dispatch(req, check_limit=1 << 16)
    out[f"{name}_req"] = req.view(np.uint8)
    out[f"{name}_rep"] = rep.view(np.uint8)
    print(name, len(req), np.bincount(rep["status"], minlength=8))
np.savez_compressed(Path(__file__).parent / "dispatch_fixtures.npz", **out)
