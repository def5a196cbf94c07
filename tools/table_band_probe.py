"""Times dispatch_batches for registries whose slot table falls between the launch shapes
(a few tens of KiB up to the shared-memory budget): microseconds per 2^20 calls, 16 batches of
2^22 calls per launch, wall clock over 5 launches.  PTYPE_B200_BIG_CTA_KB moves the table size
from which one 1024-thread CTA per SM is launched instead of several 256-thread ones.

    python tools/table_band_probe.py [n_services ...]
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ptype_b200 as p  # noqa: E402
import workloads as w  # noqa: E402


def main():
    sizes = [int(a) for a in sys.argv[1:]] or [700, 1000, 1400, 1600, 2000, 3100, 4500, 6000]
    ctx = p.Context()
    n = 1 << 22
    for n_services in sizes:
        spec = w.stress_registry(n_services, 4)
        image = spec.image()
        ctx.publish(image)
        reqs = w.stress_calls(n, n_services, seed=3)
        d_req = ctx.device_buffer(n * 48).upload(reqs)
        d_rep = ctx.device_buffer(n * 16)
        batches = ctx.make_batches([(d_req, d_rep, n)] * 16)
        for _ in range(3):
            ctx.dispatch_batches(batches, sync=True)
        t0 = time.perf_counter()
        for _ in range(5):
            ctx.dispatch_batches(batches, sync=False)
        ctx.sync()
        us = (time.perf_counter() - t0) / (5 * 16 * 4) * 1e6
        print(f"{n_services:6d} services, table {image.header.n_slots * 16 / 1024:6.1f} KiB: "
              f"{us:6.2f} us per 2^20 calls = {64 * (1 << 20) / us / 1e3:7.1f} GB/s", flush=True)
        d_req.free()
        d_rep.free()


if __name__ == "__main__":
    main()
