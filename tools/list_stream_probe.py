"""Lists of `per_list` batches submitted back to back (asynchronously, one C call per list): microseconds per
2^20-call batch once the GPU never runs dry.  PTYPE_B200_TAIL_PERCENT=0 switches the tail launch off.

    python tools/list_stream_probe.py [per_list=20] [lists=200]
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ptype_b200 as p  # noqa: E402
import workloads as w  # noqa: E402


def main():
    per_list = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    lists = int(sys.argv[2]) if len(sys.argv) > 2 else 200
    n, nbuf = 1 << 20, 8
    ctx = p.Context([0])
    ctx.publish(w.calculator_registry().image())
    bufs = []
    for b in range(nbuf):
        reqs = w.calculator_calls(n, seed=100 + b)
        bufs.append((ctx.device_buffer(n * 48).upload(reqs), ctx.device_buffer(n * 16), n))
    batches = ctx.make_batches(bufs[i % nbuf] for i in range(per_list))
    for _ in range(5):
        ctx.dispatch_batches(batches, sync=False)
    ctx.sync()
    best = None
    for _ in range(3):
        t0 = time.perf_counter()
        for _ in range(lists):
            ctx.dispatch_batches(batches, sync=False)
        ctx.sync()
        us = (time.perf_counter() - t0) / (lists * per_list) * 1e6
        best = us if best is None else min(best, us)
    print(f"lists of {per_list} batches back to back: {best:.3f} us per batch "
          f"(PTYPE_B200_TAIL_PERCENT={os.environ.get('PTYPE_B200_TAIL_PERCENT', 'default')})", flush=True)


if __name__ == "__main__":
    main()
