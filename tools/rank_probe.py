"""Time ptype_b200_assign_seq (three kernels, device-resident records) on the C2 and C3 shapes."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ptype_b200 as p  # noqa: E402
import workloads as wl  # noqa: E402


def main() -> None:
    only = sys.argv[1] if len(sys.argv) > 1 else None
    for name, spec, make, n in (("c2", wl.calculator_registry(), wl.calculator_calls, 1 << 20),
                                ("c3", wl.stress_registry(), wl.stress_calls, 1 << 24)):
        if only and name != only:
            continue
        image = spec.image(epoch=1)
        with p.Context([0]) as ctx:
            ctx.publish(image)
            reqs = make(n, salt=image.salt)
            d_req = ctx.device_buffer(n * 48).upload(reqs)
            d_rep = ctx.device_buffer(n * 16)
            for _ in range(3):
                ctx.assign_seq(d_req, n)
                ctx.dispatch_device(d_req, n, d_rep)
            reps = 20
            t0 = time.perf_counter()
            for _ in range(reps):
                ctx.assign_seq(d_req, n)
            rank_s = (time.perf_counter() - t0) / reps
            t0 = time.perf_counter()
            for _ in range(reps):
                ctx.dispatch_device(d_req, n, d_rep)
            disp_s = (time.perf_counter() - t0) / reps
            d_seq = ctx.device_buffer(n * 8)
            for _ in range(3):
                ctx.assign_seq_array(d_req, n, d_seq)
                ctx.dispatch_seq(d_req, d_seq, n, d_rep)
            t0 = time.perf_counter()
            for _ in range(reps):
                ctx.assign_seq_array(d_req, n, d_seq)
            arr_s = (time.perf_counter() - t0) / reps
            t0 = time.perf_counter()
            for _ in range(reps):
                ctx.assign_seq_array(d_req, n, d_seq, sync=False)
                ctx.dispatch_seq(d_req, d_seq, n, d_rep)
            both_s = (time.perf_counter() - t0) / reps
            print(f"{name}: assign_seq_array {arr_s * 1e6:9.1f} us   assign_seq_array + dispatch_seq {both_s * 1e6:9.1f} us",
                  flush=True)
            print(f"{name}: n={n} slots={image.header.n_slots} assign_seq {rank_s * 1e6:9.1f} us "
                  f"({n / rank_s / 1e9:6.2f} G calls/s)   dispatch (one launch, synchronous) {disp_s * 1e6:9.1f} us",
                  flush=True)


if __name__ == "__main__":
    main()
