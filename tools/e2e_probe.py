"""Where the time of one host-buffer batch goes (ptype_b200_call_batch_values over pinned memory).

Runs the C2 client batch (2^20 calls) through the public call with and without the per-call method
byte; run it under different PTYPE_B200_ZC_CTAS / PTYPE_B200_ZERO_COPY settings to compare grids and
the chunked copy-engine pipeline.  Prints one line per variant: microseconds per batch, calls/s.
"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ptype_b200 as p  # noqa: E402
import workloads as wl  # noqa: E402


def main() -> None:
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 200
    spec = wl.calculator_registry()
    image = spec.image(epoch=1)
    ctx = p.Context([0])
    ctx.publish(image)
    stream = wl.calculator_calls(n, seed=0x5EED0002, salt=image.salt)
    key, base, methods, sel, cargs = wl.as_client_batch(stream)
    pin_args = p.PinnedArray(cargs.size, np.int64)
    pin_args.array[:] = cargs.reshape(-1)
    pa = pin_args.array.reshape(cargs.shape)
    pin_sel = p.PinnedArray(n, np.uint8)
    pin_sel.array[:] = sel
    pin_val = p.PinnedArray(n, np.int64)
    pin_rep = p.PinnedArray(n, p.REPLY_DTYPE)
    tag = "ZC_CTAS=%s ZERO_COPY=%s" % (os.environ.get("PTYPE_B200_ZC_CTAS", "-"),
                                        os.environ.get("PTYPE_B200_ZERO_COPY", "-"))

    def timed(label, fn):
        for _ in range(5):
            fn()
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        dt = (time.perf_counter() - t0) / reps
        print(f"{tag:28s} {label:44s} {dt * 1e6:8.1f} us/batch  {n / dt / 1e9:6.3f} G calls/s", flush=True)

    timed("values, method byte per call",
          lambda: ctx.call_batch_values(key, base, methods, pa, pin_sel.array, pin_val.array, failed_cap=1024))
    timed("values, one method (no method bytes)",
          lambda: ctx.call_batch_values(key, base, methods[:1], pa, None, pin_val.array, failed_cap=1024))
    timed("16-byte reply records, method byte per call",
          lambda: ctx.call_batch(key, base, methods, pa, pin_sel.array, pin_rep.array))
    timed("values, 1024 calls (fixed cost of a call)",
          lambda: ctx.call_batch_values(key, base, methods, pa[:1024], pin_sel.array[:1024], pin_val.array[:1024],
                                        failed_cap=1024))


if __name__ == "__main__":
    main()
