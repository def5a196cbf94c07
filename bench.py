#!/usr/bin/env python
"""bench.py — batched RPC dispatch throughput (BASELINE.json metric) on N B200s.

A step = one pass of the fused dispatch kernel over one batch of the named workload.
Default workload (N=1 and up): BASELINE config 2 — 2^20 mixed Add/Multiply calls,
1 service x 16 registered nodes, MaxConnections 3.  For N>1 the stream is sharded by
request-id hash (mix64(id) % N), a full registry replica per GPU installed by one
ncclBroadcast (inside the library); the data path has no collective ("scaling": "weak").
Barriers and the max-over-ranks reduction of the device times use a gloo group.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload c2|c2mesh|c3|c4]

`value`     calls/s with batches resident in HBM (CUDA events on the library's stream).
`e2e`       calls/s through the C ABI with pinned HOST buffers, H2D/D2H inside the timing.
`roofline`  64 algorithmic bytes per call / average launch duration vs measured HBM peak.
`cpu_baseline` the CPU oracle (a C port of the reference's path; Go itself cannot run on
            this image) on the box's host cores, bounded sample.
--impl reference times that same CPU port as the reference arm.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

BYTES_PER_CALL = 64  # 48 B request + 16 B reply
BATCH = 1 << 20
# calls per step (per GPU) and resident batches rotated, per workload: C3 is quoted on 16M calls
BATCH_OF = {"c2": 1 << 20, "c2mesh": 1 << 20, "c3": 1 << 24, "c4": 1 << 20, "c5": 1 << 20}
BUFFERS_OF = {"c2": 8, "c2mesh": 8, "c3": 2, "c4": 8, "c5": 8}


def measured_peak():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


def make_workload(name: str, n: int, seed: int, salt: int = 0):
    import ptype_b200.workloads as w
    if name == "c2":
        return w.calculator_registry(), w.calculator_calls(n, seed=seed, salt=salt)
    if name == "c2mesh":
        return w.calculator_registry(mesh=True), w.calculator_calls(n, seed=seed, salt=salt)
    if name in ("c3", "c5"):
        return w.stress_registry(10_000, 64), w.stress_calls(n, 10_000, seed=seed, salt=salt)
    if name == "c4":
        return w.optimus_registry(), w.optimus_calls(n, seed=seed, salt=salt)[:n]
    raise SystemExit(f"unknown workload {name}")


WORKLOAD_TEXT = {
    "c2": "calculator batch of 2^20 mixed Add/Multiply calls, 1 service x 16 nodes, MaxConnections 3",
    "c2mesh": "calculator batch of 2^20 mixed Add/Multiply calls, 1 service x 16 nodes, mesh",
    "c3": "registry lookup stress: 10k services x 64 nodes, batches of 2^24 routed calls (1% unregistered)",
    "c4": "example/optimus splitWork stream replayed as batches of 2^20 Prime.Check calls",
    "c5": "sharded stream over the 10k x 64 registry, 2^20 routed calls per GPU per step, live registry "
          "snapshot updates (one node leaves or joins) broadcast to every GPU between batches",
}


class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons of one GPU through NVML while the bench runs."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self._stop_evt = index, [], threading.Event()
        self.max_mhz, self.ok = None, False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        while not self._stop_evt.is_set():
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                try:
                    reasons = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    reasons = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.samples.append((time.perf_counter(), mhz, reasons))
            except Exception:
                pass
            time.sleep(0.002)

    def stop(self):
        self._stop_evt.set()

    def summary(self, windows):
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["nvml_unavailable"]}
        inside = [s for s in self.samples if any(a <= s[0] <= b for a, b in windows)]
        use = inside or self.samples
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksThrottleReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake",
        }
        seen = set()
        for _, _, r in use:
            for bit, nm in names.items():
                if r & bit:
                    seen.add(nm)
        return {"sm_mhz": float(np.median([s[1] for s in use])), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(seen), "samples": len(use), "samples_in_timed_region": len(inside)}


_ORIGINAL_AFFINITY: set = set()


def restore_affinity() -> None:
    """Give the CPU baseline every host core back."""
    if _ORIGINAL_AFFINITY:
        try:
            os.sched_setaffinity(0, _ORIGINAL_AFFINITY)
        except Exception:
            pass


def bind_to_gpu_numa_node(index: int) -> None:
    """Run this rank on the CPUs next to its GPU (NVML's affinity mask) so that pinned buffers are
    first-touched on the local NUMA node; host<->device traffic of 8 ranks otherwise crosses sockets."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = {w * 64 + b for w, word in enumerate(words) for b in range(64) if (word >> b) & 1}
        allowed = os.sched_getaffinity(0)
        if cpus & allowed:
            _ORIGINAL_AFFINITY.update(allowed)
            os.sched_setaffinity(0, cpus & allowed)
    except Exception:
        pass


def cpu_port_rate(spec, reqs, threads: int, min_seconds: float, check_limit: int = 1 << 16):
    """calls/s of the CPU oracle (C port of the reference path) over repeated passes."""
    from oracle import binding as ob
    kv, snap = ob.snapshot_from_spec(spec)
    rep = np.empty(len(reqs), dtype=ob.REPLY_DTYPE)
    snap.dispatch_into(reqs, rep, check_limit, threads)  # warm
    passes, t0 = 0, time.perf_counter()
    while True:
        snap.dispatch_into(reqs, rep, check_limit, threads)
        passes += 1
        el = time.perf_counter() - t0
        if el >= min_seconds:
            break
    return passes * len(reqs) / el, passes, el, rep


def loopback_rate(reqs, callers: int, calls_per_caller: int, k: int, receivers: int):
    """BASELINE config 1: the reference's REAL path restated in C (oracle/loopback.c) - sequential
    synchronous Client.Call()s over TCP 127.0.0.1 with the net/rpc handshake and gob framing,
    `callers` independent Clients (threads), each holding k connections to its own servers."""
    from oracle import binding as ob
    import threading as th
    ob.lib()
    results = [None] * callers

    def one(i):
        part = np.ascontiguousarray(reqs[i * calls_per_caller:(i + 1) * calls_per_caller])
        results[i] = ob.loopback_run(part, k=k, receivers=receivers)[1]

    t0 = time.perf_counter()
    ts = [th.Thread(target=one, args=(i,)) for i in range(callers)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    el = time.perf_counter() - t0
    calls = sum(r.calls for r in results)
    return {"value": calls / el, "unit": "calls/s", "callers": callers, "connections_per_caller": k,
            "calls": int(calls), "seconds": el,
            "bytes_per_call_out": results[0].bytes_client_to_server / max(results[0].calls, 1),
            "bytes_per_call_in": results[0].bytes_server_to_client / max(results[0].calls, 1),
            "what": "C restatement of Client.Call -> net/rpc -> TCP loopback -> method (gob framing, one call in "
                    "flight per Client); the Go runtime itself is not restated"}


def run_reference(args, rank: int):
    """Reference arm: the reference's own path restated on the CPU (oracle port), all host
    threads, same config/metric.  Go itself cannot run here (no toolchain on this image)."""
    if rank != 0:
        return
    spec, reqs = make_workload(args.workload, args.batch or BATCH_OF[args.workload], 0x5EED0002)
    from oracle import binding as ob
    kv, snap = ob.snapshot_from_spec(spec)
    threads = os.cpu_count() or 1
    rep = np.empty(len(reqs), dtype=ob.REPLY_DTYPE)
    for _ in range(args.warmup):
        snap.dispatch_into(reqs, rep, 1 << 16, threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        snap.dispatch_into(reqs, rep, 1 << 16, threads)
    el = time.perf_counter() - t0
    value = args.steps * len(reqs) / el
    k = len(snap.selected("prime_worker" if args.workload == "c4" else "calculator") or [0]) \
        if args.workload != "c3" else 3
    lb1 = loopback_rate(reqs, 1, min(len(reqs), 30_000), k, 2 if args.workload == "c4" else 1)
    lbn = loopback_rate(reqs, threads, min(len(reqs) // threads, 20_000), k, 2 if args.workload == "c4" else 1)
    line = {
        "impl": "reference", "metric": "rpc_calls_per_sec_batched_dispatch", "value": value,
        "unit": "calls/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": el / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int64", "data": "synthetic",
        "config": {"workload": WORKLOAD_TEXT[args.workload], "batch_calls": len(reqs),
                   "note": "Go reference path not run - no Go toolchain on this image; C port of "
                           "cluster/rpc.go + registry.go + example method bodies, in-process"},
        "cpu_baseline": {"value": value, "unit": "calls/s", "cores": threads, "kind": "port",
                         "sample": f"{args.steps} passes over one {len(reqs)}-call batch",
                         "what": "in-process C port of the path's algorithm, no transport: the fastest CPU "
                                 "form, used for `value`"},
        "loopback_1_caller": lb1, "loopback_all_threads": lbn,
        "e2e": {"value": value, "unit": "calls/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=list(WORKLOAD_TEXT))
    ap.add_argument("--buffers", type=int, default=0, help="distinct resident batches rotated (0 = per workload)")
    ap.add_argument("--batch", type=int, default=0, help="calls per step per GPU (0 = per workload)")
    ap.add_argument("--e2e-steps", type=int, default=0, help="0 = auto (about 1 s)")
    ap.add_argument("--cpu-seconds", type=float, default=8.0)
    ap.add_argument("--flags", type=int, default=0, help="extra dispatch flags (4 = probe table in L2)")
    ap.add_argument("--update-every", type=int, default=64, help="c5: steps between snapshot updates")
    ap.add_argument("--update-mode", default="image", choices=["image", "delta"],
                    help="c5: 'image' rebuilds the snapshot on rank 0 and broadcasts it (NCCL); 'delta' has "
                         "every rank apply the same watch event to its table on the device")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    import torch.distributed as dist
    import ptype_b200 as p
    from ptype_b200.sharding import shard_indices

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the dispatch path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    bind_to_gpu_numa_node(local_rank)
    multi = world > 1
    if multi:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # Host-side plumbing (barriers, the max over ranks of the device times, the NCCL id) goes
        # over gloo.  The path's one collective, the registry-image broadcast, is NCCL inside the
        # library.  A torch NCCL process group is deliberately NOT created: with it (peer access
        # enabled by torch's NCCL init) the same HBM-bound kernel ran 12.1 us per step instead of
        # 9.5 us on every rank (measured at N=2; NCCL_P2P_DISABLE=1 restored 9.7 us).
        # PTYPE_BENCH_TORCH_NCCL=1 reproduces that.
        if os.environ.get("PTYPE_BENCH_TORCH_NCCL") == "1":
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        else:
            dist.init_process_group("gloo")

    sampler = ClockSampler(local_rank)
    sampler.start()

    ctx = p.Context([local_rank])
    # ---- registry snapshot: built on rank 0, replicated to every GPU over NCCL
    spec, _ = make_workload(args.workload, 32, 1)
    live = args.workload == "c5"
    builder = None
    if live and rank == 0:   # a builder kept alive: each update changes one key and re-emits
        builder = p.ImageBuilder(spec.local_addr, spec.max_connections)
        for svc, node, addr, port in spec.nodes:
            builder.register(svc, node, addr, port)
        for svc, recv in spec.expose:
            builder.expose(svc, recv)
        builder.set_device_only(True)
    first_image = (builder.finish(1) if live else spec.image(epoch=1)) if rank == 0 or not multi else None
    torch_nccl = os.environ.get("PTYPE_BENCH_TORCH_NCCL") == "1"
    if multi:
        ids = [p.Context.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.comm_init(ids[0], rank, world)
        image = ctx.broadcast_image(first_image, root=0)
    else:
        image = first_image
        ctx.publish(image)
    live_nodes = {(x[0], x[1]): x for x in spec.nodes} if live else {}
    live_rng = np.random.default_rng(0x5EED0005)
    update_ms = []

    def live_update(epoch: int):
        """One registry change on rank 0 (a random node leaves, or a departed one returns), a new
        device-only image, and its installation on every GPU: ncclBroadcast, or publish on 1 GPU."""
        t0 = time.perf_counter()
        img = None
        # every rank draws the same event (same seed), like every host seeing the same etcd watch
        svc = f"svc_{int(live_rng.integers(0, 10_000)):05d}"
        node = f"node_{int(live_rng.integers(0, 64)):02d}"
        leaving = (svc, node) in live_nodes
        if leaving:
            del live_nodes[(svc, node)]
        else:
            live_nodes[(svc, node)] = (svc, node, "10.99.0.1", 9000)
        if args.update_mode == "delta":
            ctx.update_service(svc, -1 if leaving else +1)
            update_ms.append((time.perf_counter() - t0) * 1e3)
            return
        if rank == 0:
            if leaving:
                builder.unregister(svc, node)
            else:
                builder.register(svc, node, "10.99.0.1", 9000)
            img = builder.finish(epoch)
        if multi:
            ctx.broadcast_image(img, root=0)
        else:
            ctx.publish(img)
        update_ms.append((time.perf_counter() - t0) * 1e3)

    # ---- per-rank shard of the step's global stream, `buffers` distinct steps resident in HBM
    nbuf = args.buffers or BUFFERS_OF[args.workload]
    batch = args.batch or BATCH_OF[args.workload]
    d_reqs, d_reps, counts, host_batches = [], [], [], []
    for b in range(nbuf):
        _, stream = make_workload(args.workload, batch * world, 0x5EED0002 + b, image.salt)
        mine = stream[shard_indices(len(stream), world, rank)] if multi else stream
        host_batches.append(mine)
        counts.append(len(mine))
        d_reqs.append(ctx.device_buffer(len(mine) * 48).upload(mine))
        d_reps.append(ctx.device_buffer(len(mine) * 16))
    stream_ptr = ctx.stream()
    ext = torch.cuda.ExternalStream(stream_ptr, device=torch.device("cuda", local_rank))

    def step(i: int):
        b = i % nbuf
        ctx.dispatch_device(d_reqs[b], counts[b], d_reps[b], flags=args.flags, sync=False)

    # the K timed steps are submitted by one C call (ptype_b200_dispatch_batches): K kernel
    # launches back to back, no Python between them
    timed = ctx.make_batches((d_reqs[i % nbuf], d_reps[i % nbuf], counts[i % nbuf])
                             for i in range(args.steps))

    def barrier():
        if multi:
            dist.barrier()
        torch.cuda.synchronize()
        ctx.sync()

    for i in range(max(args.warmup, 3)):
        step(i)
    if live:   # the update path warms up too (first-use costs of the rebuild and the collective)
        live_update(2)
        step(0)
        update_ms.clear()
    barrier()
    launches0 = ctx.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    w0 = time.perf_counter()
    ev0.record(ext)
    if live:
        every = max(args.update_every, 1)
        for lo in range(0, args.steps, every):
            seg = ctx.make_batches((d_reqs[i % nbuf], d_reps[i % nbuf], counts[i % nbuf])
                                   for i in range(lo, min(lo + every, args.steps)))
            ctx.dispatch_batches(seg, flags=args.flags, sync=False)
            if lo + every < args.steps:
                live_update(3 + lo // every)
    else:
        ctx.dispatch_batches(timed, flags=args.flags, sync=False)
    host_enqueue_s = time.perf_counter() - w0
    ev1.record(ext)
    barrier()
    w1 = time.perf_counter()
    launches = ctx.launch_count() - launches0
    ms = ev0.elapsed_time(ev1)
    calls_mine = sum(counts[i % nbuf] for i in range(args.steps))
    if multi:
        rdev = "cuda" if torch_nccl else "cpu"
        t = torch.tensor([ms], device=rdev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_max = float(t.item())
        c = torch.tensor([calls_mine], device=rdev, dtype=torch.float64)
        dist.all_reduce(c, op=dist.ReduceOp.SUM)
        calls_all = float(c.item())
    else:
        ms_max, calls_all = ms, float(calls_mine)
    value = calls_all / (ms_max * 1e-3)

    # ---- CUDA events around >= 50 individual launches, median and best.  Outside
    # the timed region; the events between launches keep consecutive kernels from overlapping, so
    # this is the duration of one launch on its own, not the pipelined step time above.
    per_launch = None
    if not live:
        pairs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(64)]
        for i, (a, b) in enumerate(pairs):
            a.record(ext)
            step(i)
            b.record(ext)
        barrier()
        us = sorted(a.elapsed_time(b) * 1e3 for a, b in pairs)
        per_launch = {"launches": len(us), "median_us": us[len(us) // 2], "best_us": us[0],
                      "what": "one dispatch launch per batch, CUDA events around each (rank 0's GPU)"}

    # ---- e2e: the C-ABI call a host makes, pinned HOST buffers, copies inside the timing
    hb = host_batches[0]
    pin_req = p.PinnedArray(len(hb), p.REQ_DTYPE)
    pin_rep = p.PinnedArray(len(hb), p.REPLY_DTYPE)
    pin_req.array[:] = hb
    for _ in range(3):
        ctx.dispatch(pin_req.array, pin_rep.array, flags=args.flags)
    t0 = time.perf_counter()
    ctx.dispatch(pin_req.array, pin_rep.array, flags=args.flags)
    one = time.perf_counter() - t0
    e2e_steps = args.e2e_steps or max(5, min(2000, int(1.0 / max(one, 1e-6))))
    barrier()
    e0 = time.perf_counter()
    for _ in range(e2e_steps):
        ctx.dispatch(pin_req.array, pin_rep.array, flags=args.flags)
    barrier()
    e1 = time.perf_counter()
    e2e_s = e1 - e0
    if multi:
        t = torch.tensor([e2e_s], device=rdev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
        c = torch.tensor([float(len(hb) * e2e_steps)], device=rdev, dtype=torch.float64)
        dist.all_reduce(c, op=dist.ReduceOp.SUM)
        e2e_calls = float(c.item())
    else:
        e2e_calls = float(len(hb) * e2e_steps)
    e2e_value = e2e_calls / e2e_s
    records_ok = pin_rep.array.tobytes()

    # ---- e2e of ONE Client's batch in compact form (ptype_b200_call_batch): what
    # (*Client).CallBatch submits - arguments plus a method byte per call instead of 48-byte
    # records.  Only for the single-service workloads.
    # At N > 1 every rank is a host process with its own Client (its own counter, rpc.go:129),
    # submitting its own batch to its own GPU over its own PCIe link: weak scaling, like `value`.
    e2e_compact = None
    cb = hb   # the Client batch that is timed
    if args.workload not in ("c3", "c5"):
        import ptype_b200.workloads as wl
        if multi:
            cb = make_workload(args.workload, batch, 0x5EED0100 + rank, image.salt)[1]
            pin_rep.free()
            pin_rep = p.PinnedArray(len(cb), p.REPLY_DTYPE)
        key, base, methods, sel, cargs = wl.as_client_batch(cb)
        pin_args = p.PinnedArray(cargs.size, np.int64)
        pin_args.array[:] = cargs.reshape(-1)
        pa = pin_args.array.reshape(cargs.shape)
        pin_sel = None
        if sel is not None:
            pin_sel = p.PinnedArray(len(sel), np.uint8)
            pin_sel.array[:] = sel
        ps = pin_sel.array if pin_sel is not None else None
        for _ in range(3):
            ctx.call_batch(key, base, methods, pa, ps, pin_rep.array, flags=args.flags & 4)
        compact_ok = pin_rep.array.tobytes()
        barrier()
        c0 = time.perf_counter()
        for _ in range(e2e_steps):
            ctx.call_batch(key, base, methods, pa, ps, pin_rep.array, flags=args.flags & 4)
        barrier()
        c1 = time.perf_counter()
        cs = c1 - c0
        cb_calls = float(len(cb) * e2e_steps)
        if multi:
            t = torch.tensor([cs], device=rdev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            cs = float(t.item())
            c = torch.tensor([cb_calls], device=rdev, dtype=torch.float64)
            dist.all_reduce(c, op=dist.ReduceOp.SUM)
            cb_calls = float(c.item())
        # ... and with the replies CallBatch's caller reads: 8-byte values + the failed list
        pin_val = p.PinnedArray(len(cb), np.int64)
        for _ in range(3):
            _, vfailed, vcount = ctx.call_batch_values(key, base, methods, pa, ps, pin_val.array,
                                                       failed_cap=1024, flags=args.flags & 4)
        values_ok = pin_val.array.copy()
        barrier()
        v0 = time.perf_counter()
        for _ in range(e2e_steps):
            ctx.call_batch_values(key, base, methods, pa, ps, pin_val.array, failed_cap=1024,
                                  flags=args.flags & 4)
        barrier()
        vs = time.perf_counter() - v0
        if multi:
            t = torch.tensor([vs], device=rdev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            vs = float(t.item())
        e2e_values = {"value": cb_calls / vs, "unit": "calls/s",
                      "h2d_bytes_per_step": int(cargs.nbytes + (len(sel) if sel is not None else 0)),
                      "d2h_bytes_per_step": int(len(cb) * 8), "steps": e2e_steps, "failed_calls": int(vcount),
                      "api": "ptype_b200_call_batch_values(args + method byte in, 8-byte values + failed list "
                             "out, pinned host buffers)"}
        e2e_compact = {"value": cb_calls / cs, "unit": "calls/s",
                       "h2d_bytes_per_step": int(cargs.nbytes + (len(sel) if sel is not None else 0)),
                       "d2h_bytes_per_step": int(len(cb) * 16), "steps": e2e_steps,
                       "api": "ptype_b200_call_batch(one Client's batch: args + method byte per call, "
                              "pinned host buffers)"}

    sampler.stop()
    sampler.join(timeout=1.0)
    clocks = sampler.summary([(w0, w1), (e0, e1)])
    e2e_records = {"value": e2e_value, "unit": "calls/s", "h2d_bytes_per_step": int(len(hb) * 48),
                   "d2h_bytes_per_step": int(len(hb) * 16), "steps": e2e_steps,
                   "api": "ptype_b200_dispatch(host pinned 48-byte records -> host pinned replies)"}

    # ---- parity spot check of what was just timed, and the CPU baseline (rank 0, N=1 only)
    cpu_baseline = None
    if rank == 0:
        from oracle import binding as ob
        if live:   # the registry as it stands after the last update
            import ptype_b200.workloads as wl
            spec = wl.RegistrySpec(sorted(live_nodes.values()), spec.expose, [], spec.local_addr,
                                   spec.max_connections)
            ctx.dispatch_device(d_reqs[0], counts[0], d_reps[0], flags=args.flags)
        kv, snap = ob.snapshot_from_spec(spec, image.salt)
        want = snap.dispatch(hb, check_limit=1 << 16, threads=os.cpu_count() or 1)
        got = d_reps[0].download(p.REPLY_DTYPE, counts[0])
        if got.tobytes() != want.tobytes() or records_ok != want.tobytes():
            raise SystemExit("bench.py: CUDA replies differ from the oracle - number withheld")
        want_cb = want if cb is hb else snap.dispatch(cb, check_limit=1 << 16, threads=os.cpu_count() or 1)
        if e2e_compact is not None and compact_ok != want_cb.tobytes():
            raise SystemExit("bench.py: call_batch replies differ from the oracle - number withheld")
        if e2e_compact is not None and (not np.array_equal(values_ok, want_cb["value"]) or
                                        vcount != int((want_cb["status"] != 0).sum())):
            raise SystemExit("bench.py: call_batch_values differs from the oracle - number withheld")
        if not multi:
            restore_affinity()
            threads = os.cpu_count() or 1
            rate, passes, el, _ = cpu_port_rate(spec, hb, threads, args.cpu_seconds)
            rate1, passes1, el1, _ = cpu_port_rate(spec, hb, 1, min(2.0, args.cpu_seconds))
            lb = loopback_rate(hb, 1, min(len(hb), 20_000), 3 if args.workload != "c4" else 2,
                               2 if args.workload == "c4" else 1) if args.workload not in ("c3", "c5") else None
            cpu_baseline = {"value": rate, "unit": "calls/s", "cores": threads, "kind": "port",
                            "loopback_config1": lb,
                            "sample": f"{passes} passes over one {len(hb)}-call batch of the same workload "
                                      f"in {el:.1f} s, {threads} threads",
                            "single_thread_value": rate1,
                            "note": "Go reference path not run - no Go toolchain on this image"}

    if rank == 0:
        peak, peak_kind = measured_peak()
        launch_s = (ms_max * 1e-3) / args.steps
        per_launch_calls = calls_all / world / args.steps
        achieved = per_launch_calls * BYTES_PER_CALL / launch_s / 1e9
        traffic, traffic_note = None, None
        tp = ROOT / "profiles" / "dispatch_traffic.json"
        if tp.exists():
            try:
                t = json.loads(tp.read_text()).get(args.workload + ("_one_launch_per_batch" if args.flags & 32 else ""))
                if isinstance(t, dict):   # measured on a launch of `batches_per_launch` batches
                    traffic = t["per_launch_bytes"]
                    traffic_note = (f"ncu: one launch = {t['batches_per_launch']} batches, "
                                    f"{t['algorithmic_bytes_per_launch']} algorithmic bytes")
                else:
                    traffic = t
            except Exception:
                traffic = None
        line = {
            "metric": "rpc_calls_per_sec_batched_dispatch", "value": value, "unit": "calls/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "int64", "data": "synthetic",
            "config": {"workload": WORKLOAD_TEXT[args.workload], "batch_calls_per_gpu": int(per_launch_calls),
                       "resident_batches": nbuf,
                       "l2": f"{nbuf} distinct batches rotated: {nbuf * per_launch_calls * 64 / 1e6:.0f} MB "
                             f"working set > 126 MB L2",
                       "sharding": "mix64(request_id) % n_gpus, registry replica per GPU via ncclBroadcast"
                                   if multi else "single GPU",
                       "table": "shared memory" if not (args.flags & 4) else "global/L2"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "traffic_note": traffic_note,
                         "peak_kind": peak_kind,
                         "kernel": "dispatch_kernel (one launch per batch)" if (args.flags & 32) else
                                   "dispatch_multi_kernel (one persistent launch per 96 batches)", "algorithmic_bytes_per_call": BYTES_PER_CALL,
                         "frac_of_nominal_8TBs": achieved / 8000.0},
            "e2e": e2e_values if e2e_compact is not None else e2e_records,
            "e2e_reply_records": e2e_compact,
            "e2e_records": e2e_records,
            "gpu_launches": int(launches),
            "per_launch": per_launch,
            "host_enqueue_us_per_step": host_enqueue_s / args.steps * 1e6,
            "clocks": clocks,
        }
        if live:
            line["config"]["live_updates"] = {"count": len(update_ms), "every_steps": args.update_every,
                                              "host_ms_avg": float(np.mean(update_ms)) if update_ms else None,
                                              "transport": "on-device patch of the table (update_service)"
                                              if args.update_mode == "delta" else
                                              ("rebuilt image, ncclBroadcast" if multi else "rebuilt image, publish")}
        if cpu_baseline is not None:
            line["cpu_baseline"] = cpu_baseline
        print(json.dumps(line), flush=True)

    pin_req.free()
    pin_rep.free()
    for b in d_reqs + d_reps:
        b.free()
    ctx.close()
    if multi:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
