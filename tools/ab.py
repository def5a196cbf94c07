"""A/B runner for bench.py: each argument is `ENV=VAL,ENV2=VAL2:bench args`; prints one summary line per
run (us per step, roofline fraction, parity) so that several configurations fit in one GPU session."""
import json
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent

for spec in sys.argv[1:]:
    envs, _, rest = spec.partition(":")
    env = dict(os.environ)
    for kv in filter(None, envs.split(",")):
        k, _, v = kv.partition("=")
        env[k] = v
    cmd = [sys.executable, str(ROOT / "bench.py")] + rest.split()
    out = subprocess.run(cmd, env=env, capture_output=True, text=True)
    try:
        d = json.loads(out.stdout.strip().splitlines()[-1])
        ex = {k: round(v["ms_per_step"] * 1e3, 3) for k, v in d.get("extras", {}).items()}
        print(f"{spec:60s} us/step {d['ms_per_step'] * 1e3:8.3f} frac {d['roofline']['frac']:.3f} "
              f"parity {d['parity']['checksum_ok']} lone {d['per_launch'] and round(d['per_launch']['median_us'], 2)} "
              f"extras {ex}", flush=True)
    except Exception as e:  # noqa: BLE001
        print(f"{spec}: FAILED rc={out.returncode} {e} {out.stderr[-400:]}", flush=True)
