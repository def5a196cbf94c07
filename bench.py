#!/usr/bin/env python
"""bench.py — batched RPC dispatch throughput (BASELINE.json metric) on N B200s.

A step = one pass of the fused dispatch kernel over one batch of the named workload.
Default workload (N=1 and up): BASELINE config 2 — 2^20 mixed Add/Multiply calls,
1 service x 16 registered nodes, MaxConnections 3.  For N>1 the stream is sharded by
request-id hash (mix64(id) % N), a full registry replica per GPU installed by one
ncclBroadcast (inside the library); the data path has no collective ("scaling": "weak").
Barriers and the max-over-ranks reduction of the device times use a gloo group.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload c2|c2mesh|c3|c4|c5] [--no-extras]

`value`     calls/s with batches resident in HBM (CUDA events on the library's stream).
`e2e`       calls/s through the C ABI with pinned HOST buffers, H2D/D2H inside the timing.
`roofline`  64 algorithmic bytes per call / average launch duration vs measured HBM peak.
`parity`    every rank's reply buffers, checksummed on the device (order-independent 64-bit
            checksum), compared with the oracle's replies for the same shard; a mismatch on
            any rank withholds the number (exit status 1).
`extras`    short runs of the other single-GPU configs (C3, C4, C5) under the same rules, so
            that the default line carries them; at N>1 a C5 segment (live snapshot updates).
`cpu_baseline` the CPU oracle (a C port of the reference's path; Go itself cannot run on
            this image) on the box's host cores, bounded sample.
--impl reference times that same CPU port as the reference arm; that arm imports nothing of
the product (workloads.py and oracle/ only).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

BYTES_PER_CALL = 64  # 48 B request + 16 B reply
# calls per step (per GPU) and resident batches rotated, per workload: C3 is quoted on 16M calls
BATCH_OF = {"c2": 1 << 20, "c2mesh": 1 << 20, "c3": 1 << 24, "c4": 1 << 20, "c5": 1 << 20}
BUFFERS_OF = {"c2": 8, "c2mesh": 8, "c3": 2, "c4": 8, "c5": 8}
U64 = (1 << 64) - 1
# the kernel dispatch_batches launches per workload (launch_dispatch_multi in csrc/dispatch.cu picks by image)
TIMED_KERNEL = {"c2mesh": "dispatch_multi_kernel", "c4": "dispatch_bulk_multi_kernel"}


def measured_peak():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


def make_workload(name: str, n: int, seed: int, salt: int = 0):
    """(registry spec, call stream) of a BASELINE config.  Pure numpy: no product code."""
    import workloads as w
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


def common_config(workload: str, batch: int, nbuf: int) -> dict:
    """The `config` object: the same keys and values in both arms for the same arguments."""
    return {"workload": WORKLOAD_TEXT[workload], "batch_calls_per_gpu": int(batch), "resident_batches": int(nbuf),
            "l2": f"{nbuf} distinct batches rotated: {nbuf * batch * BYTES_PER_CALL / 1e6:.0f} MB working set "
                  f"> 126 MB L2"}


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


def in_tree_libs():
    """Shared objects of this repository mapped into the process (evidence of what an arm ran)."""
    try:
        maps = Path("/proc/self/maps").read_text()
    except OSError:
        return None
    return sorted({ln.split()[-1][len(str(ROOT)) + 1:] for ln in maps.splitlines()
                   if ln.rstrip().endswith(".so") and str(ROOT) in ln})


GO_NOTE = ("Go reference path not run - no Go toolchain on this image; C port of cluster/rpc.go + registry.go + "
           "example method bodies, in-process")


def run_reference(args, rank: int):
    """Reference arm: the reference's own path restated on the CPU (oracle port), all host
    threads, same config/metric.  Go itself cannot run here (no toolchain on this image).
    Nothing of the product is imported: inputs come from workloads.py, replies from oracle/."""
    if rank != 0:
        return
    from oracle import binding as ob
    batch = args.batch or BATCH_OF[args.workload]
    nbuf = args.buffers or BUFFERS_OF[args.workload]
    spec, _ = make_workload(args.workload, 32, 1)
    batches = [make_workload(args.workload, batch, 0x5EED0002 + b)[1] for b in range(nbuf)]
    kv, snap = ob.snapshot_from_spec(spec)
    threads = os.cpu_count() or 1
    rep = np.empty(max(len(b) for b in batches), dtype=ob.REPLY_DTYPE)
    for i in range(args.warmup):
        snap.dispatch_into(batches[i % nbuf], rep[:len(batches[i % nbuf])], 1 << 16, threads)
    t0 = time.perf_counter()
    calls = 0
    for i in range(args.steps):
        q = batches[i % nbuf]
        snap.dispatch_into(q, rep[:len(q)], 1 << 16, threads)
        calls += len(q)
    el = time.perf_counter() - t0
    value = calls / el
    reqs = batches[0]
    k = len(snap.selected("prime_worker" if args.workload == "c4" else "calculator") or [0]) \
        if args.workload not in ("c3", "c5") else 3
    lb1 = loopback_rate(reqs, 1, min(len(reqs), 30_000), k, 2 if args.workload == "c4" else 1)
    lbn = loopback_rate(reqs, threads, min(len(reqs) // threads, 20_000), k, 2 if args.workload == "c4" else 1)
    line = {
        "impl": "reference", "metric": "rpc_calls_per_sec_batched_dispatch", "value": value,
        "unit": "calls/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": el / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int64", "data": "synthetic",
        "config": common_config(args.workload, batch, nbuf),
        "note": GO_NOTE,
        "cpu_baseline": {"value": value, "unit": "calls/s", "cores": threads, "kind": "port",
                         "sample": f"{args.steps} passes, each over one of {nbuf} rotated {batch}-call batches",
                         "what": "in-process C port of the path's algorithm, no transport: the fastest CPU "
                                 "form, used for `value`"},
        "loopback_1_caller": lb1, "loopback_all_threads": lbn,
        "e2e": {"value": value, "unit": "calls/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "in_tree_libs_loaded": in_tree_libs(),
    }
    print(json.dumps(line), flush=True)


class Env:
    """Process-level plumbing of one rank: torch (device memory, events), gloo (barriers, reductions)."""

    def __init__(self, rank, local_rank, world, torch, dist, pkg, torch_nccl):
        self.rank, self.local_rank, self.world = rank, local_rank, world
        self.torch, self.dist, self.p = torch, dist, pkg
        self.multi = world > 1
        self.rdev = "cuda" if torch_nccl else "cpu"

    def barrier(self, ctx):
        if self.multi:
            self.dist.barrier()
        self.torch.cuda.synchronize()
        ctx.sync()

    def reduce(self, x: float, op: str) -> float:
        if not self.multi:
            return float(x)
        t = self.torch.tensor([x], device=self.rdev, dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX if op == "max" else self.dist.ReduceOp.SUM)
        return float(t.item())

    def gather(self, obj):
        """Every rank's object, in rank order, on every rank."""
        if not self.multi:
            return [obj]
        out = [None] * self.world
        self.dist.all_gather_object(out, obj)
        return out


def measure(ctx, env, workload: str, *, steps: int, warmup: int, nbuf: int, batch: int, flags: int = 0,
            update_every: int = 64, update_mode: str = "image", per_launch: bool = False,
            parity_epochs: int = 2) -> dict:
    """Install the workload's registry on every GPU, keep `nbuf` batches resident, warm every kernel
    that the timed region launches, time `steps` steps with CUDA events, then check every rank's
    replies against the oracle through the device-side checksum."""
    import workloads as wl
    from ptype_b200.sharding import shard_of
    p, torch, rank, world, multi = env.p, env.torch, env.rank, env.world, env.multi
    ob = None
    if rank == 0:
        from oracle import binding as ob   # the checker: rank 0 only, never on the timed path

    spec, _ = make_workload(workload, 32, 1)
    live = workload == "c5"
    builder = None
    if live and rank == 0:   # a builder kept alive: each update changes one key and re-emits
        builder = p.ImageBuilder(spec.local_addr, spec.max_connections)
        for svc, node, addr, port in spec.nodes:
            builder.register(svc, node, addr, port)
        for svc, recv in spec.expose:
            builder.expose(svc, recv)
        builder.set_device_only(True)
    epoch = [1]
    first_image = (builder.finish(1) if live else spec.image(epoch=1)) if rank == 0 or not multi else None
    if multi:
        image = ctx.broadcast_image(first_image, root=0)
    else:
        image = first_image
        ctx.publish(image)
    live_nodes = {(x[0], x[1]): x for x in spec.nodes} if live else {}
    live_rng = np.random.default_rng(0x5EED0005)
    update_ms = []

    def live_update():
        """One registry change on rank 0 (a random node leaves, or a departed one returns), a new
        device-only image, and its installation on every GPU: ncclBroadcast, or publish on 1 GPU."""
        t0 = time.perf_counter()
        epoch[0] += 1
        img = None
        # every rank draws the same event (same seed), like every host seeing the same etcd watch
        svc = f"svc_{int(live_rng.integers(0, 10_000)):05d}"
        node = f"node_{int(live_rng.integers(0, 64)):02d}"
        leaving = (svc, node) in live_nodes
        if leaving:
            del live_nodes[(svc, node)]
        else:
            live_nodes[(svc, node)] = (svc, node, "10.99.0.1", 9000)
        if update_mode == "delta":
            ctx.update_service(svc, -1 if leaving else +1)
            update_ms.append((time.perf_counter() - t0) * 1e3)
            return
        if rank == 0:
            if leaving:
                builder.unregister(svc, node)
            else:
                builder.register(svc, node, "10.99.0.1", 9000)
            img = builder.finish(epoch[0])
        if multi:
            ctx.broadcast_image(img, root=0)
        else:
            ctx.publish(img)
        update_ms.append((time.perf_counter() - t0) * 1e3)

    def oracle_snapshot():
        now = spec if not live else wl.RegistrySpec(sorted(live_nodes.values()), spec.expose, [],
                                                    spec.local_addr, spec.max_connections)
        return ob.snapshot_from_spec(now, image.salt)

    threads = os.cpu_count() or 1
    # ---- per-rank shard of each step's global stream, `nbuf` distinct steps resident in HBM
    d_reqs, d_reps, counts, hb = [], [], [], None
    expected, streams, owners_of, want0 = [], [], [], None
    snap = oracle_snapshot()[1] if rank == 0 and not live else None
    for b in range(nbuf):
        _, stream = make_workload(workload, batch * world, 0x5EED0002 + b, image.salt)
        owners = shard_of(np.arange(len(stream), dtype=np.uint64), world) if multi else None
        mine = stream[owners == rank] if multi else stream
        if b == 0:
            hb = mine
        counts.append(len(mine))
        d_reqs.append(ctx.device_buffer(len(mine) * 48).upload(mine))
        d_reps.append(ctx.device_buffer(len(mine) * 16))
        if rank == 0 and not live:
            want = snap.dispatch(stream, check_limit=1 << 16, threads=threads)
            expected.append([ob.reply_checksum(want[owners == r] if multi else want) for r in range(world)])
            if b == 0:
                want0 = want[owners == 0] if multi else want
        elif rank == 0:
            streams.append(stream)
            owners_of.append(owners)
    stream_ptr = ctx.stream()
    ext = torch.cuda.ExternalStream(stream_ptr, device=torch.device("cuda", env.local_rank))

    def step(i: int):
        b = i % nbuf
        ctx.dispatch_device(d_reqs[b], counts[b], d_reps[b], flags=flags, sync=False)

    def batches_of(lo: int, hi: int):
        return ctx.make_batches((d_reqs[i % nbuf], d_reps[i % nbuf], counts[i % nbuf]) for i in range(lo, hi))

    # the K timed steps are submitted by C calls (ptype_b200_dispatch_batches): no Python between launches
    every = max(update_every, 1)

    def plan_steps(k: int):
        """The ptype_batch arrays of k steps, built ahead so that no Python runs between launches."""
        if not live:
            return [batches_of(0, k)]
        return [batches_of(lo, min(lo + every, k)) for lo in range(0, k, every)]

    def run_steps(segments):
        for i, seg in enumerate(segments):
            ctx.dispatch_batches(seg, flags=flags, sync=False)
            if live and i + 1 < len(segments):
                live_update()

    # ---- warm-up: W single steps, then one pass of exactly what is timed (same kernel instances,
    # same launch shapes; for C5 the update path too), so no first-use cost lands between the events
    for i in range(max(warmup, 3)):
        step(i)
    if live:
        live_update()
    run_steps(plan_steps(steps if not live else min(steps, 2 * every + 1)))
    # ... and keep the GPU under this load for a while before timing: a B200 that was idle a moment
    # ago has not settled its clocks and power state after a few hundred microseconds of work, and
    # a K = 20 timed region is itself only ~0.2 ms long (PTYPE_BENCH_SETTLE_MS, default 50)
    settle_ms = float(os.environ.get("PTYPE_BENCH_SETTLE_MS", "50"))
    if settle_ms > 0 and not live:
        t_end = time.perf_counter() + settle_ms * 1e-3
        lap = batches_of(0, 4 * nbuf)
        while time.perf_counter() < t_end:
            ctx.dispatch_batches(lap, flags=flags, sync=True)
    timed = plan_steps(steps)
    preroll = batches_of(0, min(32, steps))
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    env.barrier(ctx)
    update_ms.clear()
    # A pre-roll (untimed, at most 32 batches) keeps the GPU busy while the host enqueues the start
    # event and the first timed launch: the timed region then starts when the pre-roll ends, not
    # after a host launch latency on an idle GPU.  (The event between the two launches keeps them
    # from overlapping.)
    ctx.dispatch_batches(preroll, flags=flags, sync=False)
    launches0 = ctx.launch_count()
    w0 = time.perf_counter()
    ev0.record(ext)
    h0 = time.perf_counter()
    run_steps(timed)
    host_enqueue_s = time.perf_counter() - h0   # the host's cost of submitting the K steps (and C5's updates)
    ev1.record(ext)
    env.barrier(ctx)
    w1 = time.perf_counter()
    launches = ctx.launch_count() - launches0
    ms = ev0.elapsed_time(ev1)
    calls_mine = sum(counts[i % nbuf] for i in range(steps))
    ms_max = env.reduce(ms, "max")
    calls_all = env.reduce(float(calls_mine), "sum")

    # ---- CUDA events around >= 50 individual launches, median and best.  Outside
    # the timed region; the events between launches keep consecutive kernels from overlapping, so
    # this is the duration of one launch on its own, not the pipelined step time above.
    lone = None
    if per_launch and not live:
        pairs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(64)]
        for i, (a, b) in enumerate(pairs):
            a.record(ext)
            step(i)
            b.record(ext)
        env.barrier(ctx)
        us = sorted(a.elapsed_time(b) * 1e3 for a, b in pairs)
        lone = {"launches": len(us), "median_us": us[len(us) // 2], "best_us": us[0],
                "what": "one dispatch launch per batch, CUDA events around each (rank 0's GPU)"}

    # ---- parity on EVERY rank: device-side checksum of each reply buffer vs the oracle's replies
    # for that rank's shard (rank 0 holds the oracle; the sums travel over gloo)
    checked = min(nbuf, steps)
    parity = {"ranks": world, "buffers_per_rank": checked, "checksum_ok": True, "mismatches": []}

    def compare(sums_by_rank, want_of):
        for r in range(world):
            for b in range(checked):
                if sums_by_rank[r][b] != want_of(b, r):
                    parity["checksum_ok"] = False
                    parity["mismatches"].append({"rank": r, "buffer": b})

    if not live:
        sums = env.gather([ctx.reply_checksum(d_reps[b], counts[b], device=True) for b in range(checked)])
        if rank == 0:
            compare(sums, lambda b, r: expected[b][r])
            got0 = d_reps[0].download(p.REPLY_DTYPE, counts[0])
            parity["rank0_buffer0_bytes_equal"] = bool(got0.tobytes() == want0.tobytes())
            parity["checksum_ok"] = parity["checksum_ok"] and parity["rank0_buffer0_bytes_equal"]
    else:
        # the registry keeps changing: replay `parity_epochs` more segments, one pass over the
        # resident batches each, and compare each against the oracle's view of the registry as it
        # stands in that epoch (untimed; same kernels, same update path as the timed region)
        parity["epochs_checked"] = parity_epochs
        for _ in range(parity_epochs):
            live_update()
            ctx.dispatch_batches(batches_of(0, checked), flags=flags, sync=True)
            sums = env.gather([ctx.reply_checksum(d_reps[b], counts[b], device=True) for b in range(checked)])
            if rank == 0:
                snap_now = oracle_snapshot()[1]
                wants = [snap_now.dispatch(streams[b], check_limit=1 << 16, threads=threads) for b in range(checked)]
                compare(sums, lambda b, r: ob.reply_checksum(wants[b][owners_of[b] == r] if multi else wants[b]))
    ok = env.gather(parity["checksum_ok"] if rank == 0 else None)[0]
    for buf in d_reqs + d_reps:
        buf.free()
    launch_s = (ms_max * 1e-3) / steps
    per_launch_calls = calls_all / world / steps
    peak, peak_kind = measured_peak()
    achieved = per_launch_calls * BYTES_PER_CALL / launch_s / 1e9
    out = {"workload": workload, "value": calls_all / (ms_max * 1e-3), "ms_max": ms_max, "steps": steps,
           "launches": int(launches), "host_enqueue_s": host_enqueue_s, "calls_all": calls_all,
           "per_launch_calls": per_launch_calls, "achieved_gbs": achieved, "peak": peak, "peak_kind": peak_kind,
           "parity": parity, "parity_ok": bool(ok), "window": (w0, w1), "lone": lone, "hb": hb, "image": image,
           "spec": spec if not live else wl.RegistrySpec(sorted(live_nodes.values()), spec.expose, [], spec.local_addr,
                                                        spec.max_connections),   # the registry as it stands now
           "nbuf": nbuf, "batch": batch}
    if live:
        out["live_updates"] = {"count": len(update_ms), "every_steps": update_every,
                               "host_ms_avg": float(np.mean(update_ms)) if update_ms else None,
                               "transport": "on-device patch of the table (update_service)"
                               if update_mode == "delta" else
                               ("rebuilt image, ncclBroadcast" if multi else "rebuilt image, publish")}
    return out


def measure_wire(ctx, env, n_streams: int = 4096, calls_per_stream: int = 256, repeat: int = 20) -> dict:
    """SURVEY.md 8f N1 on the device: the net/rpc request streams of `n_streams` connections decoded into
    records, dispatched, and the replies encoded as Response streams.  Inputs come from the loopback
    restatement's own encoders (oracle/loopback.c), and the encoded answer must equal its server bytes."""
    import workloads as wl
    from oracle import binding as ob
    p = env.p
    spec = wl.calculator_registry()
    ctx.publish(spec.image())
    c2s, s2c, c_off, s_off, call_off, per = [], [], [0], [0], [0], []
    for s in range(n_streams):
        reqs = wl.calculator_calls(calls_per_stream, seed=0x5EED0600 + s)
        a, _, b, _ = ob.wire_streams(reqs, receivers=1)
        c2s.append(a)
        s2c.append(b)
        c_off.append(c_off[-1] + len(a))
        s_off.append(s_off[-1] + len(b))
        call_off.append(call_off[-1] + len(reqs))
    c2s, s2c = np.concatenate(c2s), np.concatenate(s2c)
    streams = np.zeros(n_streams, dtype=ctx.WIRE_STREAM_DTYPE)
    streams["svc_key"], streams["seq0"], streams["seq_step"] = p.service_key("calculator"), 1, 1
    reqs, calls, got_off, dec_s = ctx.wire_decode(c2s, np.array(c_off, dtype=np.uint64), streams, repeat=repeat)
    replies = ctx.dispatch(reqs)
    out, out_off, _, enc_s = ctx.wire_encode(c2s, got_off, calls, replies, np.zeros(n_streams, dtype=np.uint8),
                                             repeat=repeat)
    ok = bool(np.array_equal(got_off, np.array(call_off, dtype=np.uint64)) and out.tobytes() == s2c.tobytes())
    n = len(reqs)
    peak, _ = measured_peak()
    dec_bytes = len(c2s) * 2 + n * (48 + 24)            # the stream is walked twice (count, decode)
    enc_bytes = 2 * n * (24 + 16) + len(out) + int(calls["sm_len"].sum())   # sizes pass + write pass
    return {"streams": n_streams, "calls": n, "request_bytes": int(len(c2s)), "response_bytes": int(len(out)),
            "decode_us": dec_s * 1e6, "encode_us": enc_s * 1e6,
            "decode_calls_per_s": n / dec_s, "encode_calls_per_s": n / enc_s,
            "decode_gbs": dec_bytes / dec_s / 1e9, "encode_gbs": enc_bytes / enc_s / 1e9,
            "decode_frac_of_hbm_peak": dec_bytes / dec_s / 1e9 / peak,
            "encode_frac_of_hbm_peak": enc_bytes / enc_s / 1e9 / peak,
            "timing": f"wall clock around the C-ABI call (it waits for the GPU), mean of {repeat}",
            "bytes_equal_loopback_streams": ok, "parity_ok": ok}


def measure_assign_seq(ctx, env, n: int = 1 << 24, reps: int = 10) -> dict:
    """SURVEY.md 8f N4 on the C3 registry: the balancers' counters drawn on the device for a batch that
    mixes 10 000 services - into the records (assign_seq), and into an array beside them followed by the
    dispatch that reads it (assign_seq_array + dispatch_seq).  CUDA events on the library's stream;
    the first draw is compared with the oracle's sequential rule."""
    import workloads as wl
    from oracle import binding as ob
    p, torch = env.p, env.torch
    spec = wl.stress_registry(10_000, 64)
    image = spec.image(epoch=1)
    ctx.publish(image)
    kv, snap = ob.snapshot_from_spec(spec, image.salt)
    reqs = wl.stress_calls(n, 10_000, seed=0x5EED0700, salt=image.salt)
    reqs["seq"] = 0
    want = reqs.copy()
    counters = np.zeros(snap.n_clients(), dtype=np.uint64)
    snap.assign_seq(want, counters)
    d_req = ctx.device_buffer(n * 48).upload(reqs)
    d_seq = ctx.device_buffer(n * 8)
    d_rep = ctx.device_buffer(n * 16)
    ext = torch.cuda.ExternalStream(ctx.stream(), device=torch.device("cuda", env.local_rank))
    ctx.assign_seq_array(d_req, n, d_seq, sync=False)
    ctx.dispatch_seq(d_req, d_seq, n, d_rep)
    ok = bool(np.array_equal(d_seq.download(np.uint64, n), want["seq"]) and
              d_rep.download(p.REPLY_DTYPE, n).tobytes() ==
              snap.dispatch(want, check_limit=1 << 16, threads=os.cpu_count() or 1).tobytes())

    def timed(fn):
        fn()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ctx.sync()
        a.record(ext)
        for _ in range(reps):
            fn()
        b.record(ext)
        ctx.sync()
        return a.elapsed_time(b) * 1e3 / reps

    in_records = timed(lambda: ctx.assign_seq(d_req, n, sync=False))
    to_array = timed(lambda: ctx.assign_seq_array(d_req, n, d_seq, sync=False))
    with_dispatch = timed(lambda: (ctx.assign_seq_array(d_req, n, d_seq, sync=False),
                                   ctx.dispatch_seq(d_req, d_seq, n, d_rep, sync=False)))
    for b in (d_req, d_seq, d_rep):
        b.free()
    peak, _ = measured_peak()
    return {"records": n, "services": 10_000,
            "assign_seq_us": in_records, "assign_seq_array_us": to_array,
            "assign_seq_array_plus_dispatch_seq_us": with_dispatch,
            "records_per_s": n / (in_records * 1e-6), "array_records_per_s": n / (to_array * 1e-6),
            "roofline": {"bound": "hbm", "algorithmic_bytes_per_record": 16,
                         "what": "8-byte key read + 8-byte sequence number written per record",
                         "achieved_gbs": n * 16 / (in_records * 1e-6) / 1e9, "peak": peak,
                         "frac": n * 16 / (in_records * 1e-6) / 1e9 / peak,
                         "frac_array_form": n * 16 / (to_array * 1e-6) / 1e9 / peak,
                         "note": "the records are 48-byte structures: reading an 8-byte key moves 48 bytes per "
                                 "record through DRAM, and storing 8 bytes into a record costs a 32-byte sector "
                                 "read and written (tools/partial_write_probe.cu); the array form avoids the latter"},
            "parity_ok": ok}


def extra_summary(m: dict) -> dict:
    """A sub-run as it appears under `extras` in the JSON line."""
    s = {"workload": WORKLOAD_TEXT[m["workload"]], "value": m["value"], "unit": "calls/s", "steps": m["steps"],
         "ms_per_step": m["ms_max"] / m["steps"], "batch_calls_per_gpu": m["batch"], "resident_batches": m["nbuf"],
         "roofline_frac": m["achieved_gbs"] / m["peak"], "achieved_gbs": m["achieved_gbs"],
         "gpu_launches": m["launches"], "parity": m["parity"]}
    if "live_updates" in m:
        s["live_updates"] = m["live_updates"]
    return s


def measure_e2e(ctx, env, args, m: dict) -> dict:
    """End to end through the C-ABI calls a host makes: pinned HOST buffers in, host buffers out,
    the copies inside the timed region.  Three forms: 48-byte records (ptype_b200_dispatch), one
    Client's compact batch with 16-byte replies (call_batch) and with 8-byte values + failed list
    (call_batch_values - what the CallBatch mirrors call).  Every result is compared with the oracle."""
    import workloads as wl
    p, rank, multi = env.p, env.rank, env.multi
    hb, image, spec, workload, batch = m["hb"], m["image"], m["spec"], m["workload"], m["batch"]
    flags = args.flags
    pin_req = p.PinnedArray(len(hb), p.REQ_DTYPE)
    pin_rep = p.PinnedArray(len(hb), p.REPLY_DTYPE)
    pin_req.array[:] = hb
    for _ in range(3):
        ctx.dispatch(pin_req.array, pin_rep.array, flags=flags)
    t0 = time.perf_counter()
    ctx.dispatch(pin_req.array, pin_rep.array, flags=flags)
    one = time.perf_counter() - t0
    e2e_steps = args.e2e_steps or max(5, min(2000, int(1.0 / max(one, 1e-6))))
    env.barrier(ctx)
    e0 = time.perf_counter()
    for _ in range(e2e_steps):
        ctx.dispatch(pin_req.array, pin_rep.array, flags=flags)
    env.barrier(ctx)
    e1 = time.perf_counter()
    e2e_s = env.reduce(e1 - e0, "max")
    e2e_calls = env.reduce(float(len(hb) * e2e_steps), "sum")
    records_bytes = pin_rep.array.tobytes()
    out = {"window": (e0, e1), "steps": e2e_steps,
           "records": {"value": e2e_calls / e2e_s, "unit": "calls/s", "h2d_bytes_per_step": int(len(hb) * 48),
                       "d2h_bytes_per_step": int(len(hb) * 16), "steps": e2e_steps,
                       "api": "ptype_b200_dispatch(host pinned 48-byte records -> host pinned replies)"},
           "compact": None, "values": None}

    # ---- ONE Client's batch in compact form: what (*Client).CallBatch submits - arguments plus a
    # method byte per call instead of 48-byte records.  Only for the single-service workloads.
    # At N > 1 every rank is a host process with its own Client (its own counter, rpc.go:129),
    # submitting its own batch to its own GPU over its own PCIe link: weak scaling, like `value`.
    cb = hb
    compact_bytes = values_got = vcount = None
    if workload not in ("c3", "c5"):
        if multi:
            cb = make_workload(workload, batch, 0x5EED0100 + rank, image.salt)[1]
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
        h2d = int(cargs.nbytes + (len(sel) if sel is not None else 0))
        for _ in range(3):
            ctx.call_batch(key, base, methods, pa, ps, pin_rep.array, flags=flags & 4)
        compact_bytes = pin_rep.array.tobytes()
        env.barrier(ctx)
        c0 = time.perf_counter()
        for _ in range(e2e_steps):
            ctx.call_batch(key, base, methods, pa, ps, pin_rep.array, flags=flags & 4)
        env.barrier(ctx)
        cs = env.reduce(time.perf_counter() - c0, "max")
        cb_calls = env.reduce(float(len(cb) * e2e_steps), "sum")
        # ... and with the replies CallBatch's caller reads: 8-byte values + the failed list
        pin_val = p.PinnedArray(len(cb), np.int64)
        for _ in range(3):
            _, _, vcount = ctx.call_batch_values(key, base, methods, pa, ps, pin_val.array,
                                                 failed_cap=1024, flags=flags & 4)
        values_got = pin_val.array.copy()
        env.barrier(ctx)
        v0 = time.perf_counter()
        for _ in range(e2e_steps):
            ctx.call_batch_values(key, base, methods, pa, ps, pin_val.array, failed_cap=1024, flags=flags & 4)
        env.barrier(ctx)
        vs = env.reduce(time.perf_counter() - v0, "max")
        out["values"] = {"value": cb_calls / vs, "unit": "calls/s", "h2d_bytes_per_step": h2d,
                         "d2h_bytes_per_step": int(len(cb) * 8), "steps": e2e_steps, "failed_calls": int(vcount),
                         "api": "ptype_b200_call_batch_values(args + method byte in, 8-byte values + failed list "
                                "out, pinned host buffers)"}
        out["compact"] = {"value": cb_calls / cs, "unit": "calls/s", "h2d_bytes_per_step": h2d,
                          "d2h_bytes_per_step": int(len(cb) * 16), "steps": e2e_steps,
                          "api": "ptype_b200_call_batch(one Client's batch: args + method byte per call, "
                                 "pinned host buffers)"}
        pin_args.free()
        pin_val.free()
        if pin_sel is not None:
            pin_sel.free()
    pin_req.free()
    pin_rep.free()

    # ---- every rank checks what its host-buffer calls returned (each rank's Client has its own batch)
    from oracle import binding as ob
    kv, snap = ob.snapshot_from_spec(spec, image.salt)
    threads = max(1, (os.cpu_count() or 1) // env.world)
    want = snap.dispatch(hb, check_limit=1 << 16, threads=threads)
    good = records_bytes == want.tobytes()
    if compact_bytes is not None:
        want_cb = want if cb is hb else snap.dispatch(cb, check_limit=1 << 16, threads=threads)
        good = good and compact_bytes == want_cb.tobytes() and np.array_equal(values_got, want_cb["value"]) \
            and vcount == int((want_cb["status"] != 0).sum())
    out["parity_ok"] = all(env.gather(bool(good)))
    return out


def mirror_e2e(env, args, workload: str, batch: int, steps: int):
    """End to end through the C++ mirror of the reference's Client (ptype_b200/host/cluster.hpp):
    tools/ptype_b200_bench drives cluster::Client::CallBatch - argument marshalling, sequence
    reservation, the C-ABI call and reply unpacking inside the timing - one process per GPU."""
    tool = ROOT / "tools" / "ptype_b200_bench"
    if workload not in ("c2", "c2mesh") or not tool.exists():
        return None
    cmd = [str(tool), "mirror", str(batch), str(steps), str(env.local_rank), "1" if workload == "c2mesh" else "0"]
    if env.multi:
        env.dist.barrier()
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
        kv = dict(zip(r.stdout.split()[0::2], r.stdout.split()[1::2])) if r.returncode == 0 else {}
    except Exception:
        kv = {}
    secs = float(kv.get("seconds", 0.0))
    calls = float(kv.get("calls", 0.0))
    all_ok = all(env.gather(bool(kv) and kv.get("parity") == "ok"))
    secs_max, calls_all = env.reduce(secs, "max"), env.reduce(calls, "sum")
    arr_secs = env.reduce(float(kv.get("array_seconds", 0.0)), "max")
    arr_calls = env.reduce(float(kv.get("array_calls", 0.0)), "sum")
    if not all_ok or secs_max <= 0:
        return {"value": None, "error": "tools/ptype_b200_bench mirror failed or its replies were wrong"}
    return {"value": calls_all / secs_max, "unit": "calls/s", "steps": steps,
            "array_form": {"value": arr_calls / arr_secs if arr_secs > 0 else None, "unit": "calls/s",
                           "api": "cluster::Client::CallBatch(method, int64[n][3], n, BatchReply*): arguments "
                                  "copied into the pooled batch, 16-byte reply records unpacked, one host thread"},
            "h2d_bytes_per_step": int(float(kv.get("h2d_bytes_per_step", 0))),
            "d2h_bytes_per_step": int(float(kv.get("d2h_bytes_per_step", 0))),
            "api": "cluster::Client::CallBatch (C++ mirror of cluster/rpc.go:59-67) -> ptype_b200_call_batch_values"}


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
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the short C3/C4/C5 sub-runs the default (c2) line carries")
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

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the dispatch path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    bind_to_gpu_numa_node(local_rank)
    multi = world > 1
    torch_nccl = os.environ.get("PTYPE_BENCH_TORCH_NCCL") == "1"
    if multi:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # Host-side plumbing (barriers, the max over ranks of the device times, the NCCL id) goes
        # over gloo.  The path's one collective, the registry-image broadcast, is NCCL inside the
        # library.  A torch NCCL process group is deliberately NOT created: with it (peer access
        # enabled by torch's NCCL init) the same HBM-bound kernel ran 12.1 us per step instead of
        # 9.5 us on every rank (measured at N=2; NCCL_P2P_DISABLE=1 restored 9.7 us).
        # PTYPE_BENCH_TORCH_NCCL=1 reproduces that.
        if torch_nccl:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        else:
            dist.init_process_group("gloo")
    env = Env(rank, local_rank, world, torch, dist, p, torch_nccl)

    sampler = ClockSampler(local_rank)
    sampler.start()

    ctx = p.Context([local_rank])
    if multi:
        ids = [p.Context.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.comm_init(ids[0], rank, world)

    nbuf = args.buffers or BUFFERS_OF[args.workload]
    batch = args.batch or BATCH_OF[args.workload]
    m = measure(ctx, env, args.workload, steps=args.steps, warmup=args.warmup, nbuf=nbuf, batch=batch,
                flags=args.flags, update_every=args.update_every, update_mode=args.update_mode, per_launch=True)
    e2e = measure_e2e(ctx, env, args, m)
    mirror = mirror_e2e(env, args, args.workload, batch, max(5, min(200, e2e["steps"])))
    windows = [m["window"], e2e["window"]]

    # ---- the other configs, short, under the same rules (the default line carries them)
    extras = {}
    if args.workload == "c2" and not args.no_extras and not args.flags:
        plan = [("c5", dict(steps=192, nbuf=4))] if multi else \
               [("c3", dict(steps=12, nbuf=2)), ("c4", dict(steps=200, nbuf=8)), ("c5", dict(steps=256, nbuf=8))]
        for name, kw in plan:
            x = measure(ctx, env, name, warmup=3, batch=BATCH_OF[name], update_every=64, **kw)
            windows.append(x["window"])
            extras[name] = x

    wire = ranks = None
    if args.workload == "c2" and not args.no_extras and not args.flags and not multi:
        wire = measure_wire(ctx, env)
        ranks = measure_assign_seq(ctx, env)

    sampler.stop()
    sampler.join(timeout=1.0)
    clocks = sampler.summary(windows)

    parity_ok = m["parity_ok"] and e2e["parity_ok"] and all(x["parity_ok"] for x in extras.values()) and \
        (wire is None or wire["parity_ok"]) and (ranks is None or ranks["parity_ok"])
    if not parity_ok:
        if rank == 0:
            bad = {"primary": m["parity"], "e2e_ok": e2e["parity_ok"],
                   "extras": {k: x["parity"] for k, x in extras.items()}, "wire": wire and wire["parity_ok"],
                   "assign_seq": ranks and ranks["parity_ok"]}
            print(json.dumps({"error": "CUDA replies differ from the oracle - number withheld", "parity": bad}),
                  file=sys.stderr, flush=True)
        ctx.close()
        if multi:
            dist.destroy_process_group()
        raise SystemExit(1)

    # ---- the CPU baseline (rank 0, N=1 only)
    cpu_baseline = None
    if rank == 0 and not multi:
        restore_affinity()
        spec, hb = m["spec"], m["hb"]
        if args.workload == "c5":   # the registry as first installed: the CPU port has no live updates to pay for
            spec = make_workload("c3", 32, 1)[0]
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
            "metric": "rpc_calls_per_sec_batched_dispatch", "value": m["value"], "unit": "calls/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": m["ms_max"] / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "int64", "data": "synthetic",
            "config": common_config(args.workload, batch, nbuf),
            "placement": {"sharding": "mix64(request_id) % n_gpus, registry replica per GPU via ncclBroadcast"
                                      if multi else "single GPU",
                          "table": "shared memory" if not (args.flags & 4) else "global/L2",
                          "calls_per_gpu_per_step": m["per_launch_calls"],
                          "warm": "W single-launch steps, one untimed pass of the timed launch list, a pre-roll of "
                                  "min(32, K) batches in flight when the start event is recorded"},
            "roofline": {"bound": "hbm", "achieved": m["achieved_gbs"], "peak": m["peak"], "unit": "GB/s",
                         "frac": m["achieved_gbs"] / m["peak"], "traffic": traffic, "traffic_note": traffic_note,
                         "peak_kind": m["peak_kind"],
                         "kernel": "dispatch_kernel (one launch per batch)" if (args.flags & 32) else
                                   TIMED_KERNEL.get(args.workload, "dispatch_multi_prefetch_kernel") +
                                   " (one persistent launch per <= 96 batches)",
                         "algorithmic_bytes_per_call": BYTES_PER_CALL,
                         "frac_of_nominal_8TBs": m["achieved_gbs"] / 8000.0},
            "e2e": e2e["values"] if e2e["values"] is not None else e2e["records"],
            "e2e_mirror": mirror,
            "e2e_reply_records": e2e["compact"],
            "e2e_records": e2e["records"],
            "parity": m["parity"],
            "gpu_launches": m["launches"],
            "per_launch": m["lone"],
            "host_enqueue_us_per_step": m["host_enqueue_s"] / args.steps * 1e6,
            "extras": {k: extra_summary(x) for k, x in extras.items()},
            "wire_codec": wire,
            "assign_seq": ranks,
            "clocks": clocks,
        }
        if "live_updates" in m:
            line["placement"]["live_updates"] = m["live_updates"]
        if cpu_baseline is not None:
            line["cpu_baseline"] = cpu_baseline
        print(json.dumps(line), flush=True)

    ctx.close()
    if multi:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
