"""Parity of the CUDA dispatch path with the CPU oracle, through the C ABI.

Bit-exact bar: every reply record (value, node, status) must equal the oracle's on
the same seeded inputs.  Sizes the oracle finishes in seconds are compared in full;
BASELINE.json's full sizes are also checked through size-independent properties
(idempotence, shard-sum of checksums, status accounting)."""
import threading
from pathlib import Path

import numpy as np
import pytest

import ptype_b200 as p
import workloads as w
from ptype_b200.sharding import shard_indices

pytestmark = pytest.mark.gpu
GOLD = Path(__file__).parent / "golden"
LIMIT = 1 << 16  # the context's default Prime.Check iteration limit


def assert_same(got: np.ndarray, want: np.ndarray, reqs=None):
    if got.tobytes() == want.tobytes():
        return
    bad = np.flatnonzero((got["value"] != want["value"]) | (got["node"] != want["node"]) |
                         (got["status"] != want["status"]))
    j = int(bad[0])
    raise AssertionError(f"{len(bad)} of {len(got)} replies differ; first at {j}: got {got[j]} want {want[j]}"
                         + (f" req {reqs[j]}" if reqs is not None else ""))


def device_roundtrip(ctx, reqs, flags=0):
    """Upload, dispatch with device pointers (one launch), download."""
    n = len(reqs)
    d_req = ctx.device_buffer(max(n, 1) * 48).upload(reqs)
    d_rep = ctx.device_buffer(max(n, 1) * 16)
    try:
        ctx.dispatch_device(d_req, n, d_rep, flags=flags)
        return d_rep.download(p.REPLY_DTYPE, n)
    finally:
        d_req.free()
        d_rep.free()


# kernel variants selectable by flag: table in shared memory or L2, requests staged by
# bulk-async copies or by LDG + shuffle transpose
VARIANTS = [0, p.F_GLOBAL_TABLE, p.F_LDG_LOADS, p.F_GLOBAL_TABLE | p.F_LDG_LOADS, p.F_BULK_LOADS,
            p.F_GLOBAL_TABLE | p.F_BULK_LOADS]

SPECS = {
    "calc3": lambda: w.calculator_registry(),
    "calcmesh": lambda: w.calculator_registry(mesh=True),
    "optimus": lambda: w.optimus_registry(),
}


@pytest.mark.parametrize("name", list(SPECS))
def test_golden_fixtures(ctx, name):
    """CUDA replies equal the committed fixture replies (tests/golden/make_fixtures.py)."""
    fx = np.load(GOLD / "dispatch_fixtures.npz")
    ctx.set_check_limit(LIMIT)
    ctx.publish(SPECS[name]().image())
    req = fx[f"{name}_req"].view(p.REQ_DTYPE).reshape(-1)
    want = fx[f"{name}_rep"].view(p.REPLY_DTYPE).reshape(-1)
    assert_same(ctx.dispatch(req), want, req)
    assert_same(device_roundtrip(ctx, req), want, req)
    for flags in VARIANTS:
        assert_same(device_roundtrip(ctx, req, flags=flags), want, req)
        assert_same(ctx.dispatch(req, flags=flags), want, req)


@pytest.mark.parametrize("name", list(SPECS))
@pytest.mark.parametrize("limit", [LIMIT, 1 << 20, 0])
def test_edge_cases_every_status(ctx, oracle, name, limit):
    spec = SPECS[name]()
    spec.declare = ["declared_but_empty"]
    services = sorted({s for s, *_ in spec.nodes}) + ["declared_but_empty"]
    ctx.set_check_limit(limit)
    ctx.publish(spec.image())
    kv, snap = oracle.snapshot_from_spec(spec)
    req = w.edge_case_calls(services)
    want = snap.dispatch(req, check_limit=limit)
    seen = set(np.unique(want["status"]).tolist())
    assert {0, 1, 2, 3, 4, 5, 6} <= seen
    assert_same(ctx.dispatch(req), want, req)
    for flags in VARIANTS:
        assert_same(device_roundtrip(ctx, req, flags=flags), want, req)
    ctx.set_check_limit(LIMIT)


@pytest.mark.parametrize("n", [0, 1, 2, 31, 32, 33, 63, 64, 65, 95, 96, 97, 1000, 4097,
                               128 * 1024 - 1, 128 * 1024, 128 * 1024 + 1, 3 * 128 * 1024 + 77])
def test_ragged_sizes(ctx, oracle, n):
    """Empty, sub-warp, warp-boundary and pipeline-chunk-boundary batch sizes."""
    spec = w.calculator_registry()
    ctx.publish(spec.image())
    kv, snap = oracle.snapshot_from_spec(spec)
    req = w.calculator_calls(n, seed=0x5EED0002 + n)
    want = snap.dispatch(req)
    assert_same(ctx.dispatch(req), want, req)
    for flags in VARIANTS:
        assert_same(device_roundtrip(ctx, req, flags=flags), want, req)


@pytest.mark.parametrize("mesh", [False, True])
def test_c2_calculator_1m(ctx, oracle, mesh):
    """BASELINE config 2 at full size: 2^20 mixed Add/Multiply calls, 1 service x 16 nodes."""
    spec = w.calculator_registry(mesh=mesh)
    ctx.publish(spec.image())
    kv, snap = oracle.snapshot_from_spec(spec)
    req = w.calculator_calls(1 << 20)
    want = snap.dispatch(req, threads=8)
    got = device_roundtrip(ctx, req)
    assert_same(got, want, req)
    sel = np.array(snap.selected("calculator"))
    assert np.array_equal(np.unique(got["node"]), np.unique(sel))
    assert np.array_equal(got["node"], sel[(req["seq"] % np.uint64(len(sel))).astype(np.int64)])
    assert_same(ctx.dispatch(req), want, req)           # host-buffer pipeline, 8 chunks
    assert_same(device_roundtrip(ctx, req), got)         # idempotent


@pytest.fixture(scope="module")
def stress(oracle):
    spec = w.stress_registry(10_000, 64)
    kv, snap = oracle.snapshot_from_spec(spec)
    return spec, spec.image(), snap, kv


@pytest.mark.parametrize("zipf", [False, True])
def test_c3_registry_stress(ctx, stress, oracle, zipf):
    """BASELINE config 3: 10k services x 64 nodes (12 288-slot table staged in shared
    memory), 2^22 routed calls compared in full; 1 % name an unregistered service."""
    spec, img, snap, _ = stress
    assert img.header.n_slots == 12_288
    ctx.publish(img)
    n = 1 << 22
    req = w.stress_calls(n, 10_000, zipf=zipf)
    want = snap.dispatch(req, threads=8)
    got = device_roundtrip(ctx, req)
    assert_same(got, want, req)
    for flags in VARIANTS[1:] * 3:   # repeated: the staged-request ring once raced here
        assert_same(device_roundtrip(ctx, req, flags=flags), want, req)
    miss = got["status"] == p.ST_NO_SERVICE
    assert 0.005 < miss.mean() < 0.02 and np.all(got["node"][miss] == p.NO_NODE)
    ok = got["status"] == p.ST_OK
    assert ok.sum() + miss.sum() == n
    # every routed node belongs to the service the call named
    assert np.all((got["node"][ok] // 64) < 10_000)


def test_c3_full_size_properties(ctx, stress, oracle):
    """2^24 calls (config 3's size): full compare against the 8-thread oracle, plus
    checksum-of-shards == checksum-of-whole computed on the device."""
    spec, img, snap, _ = stress
    ctx.publish(img)
    n = 1 << 24
    req = w.stress_calls(n, 10_000)
    d_req = ctx.device_buffer(n * 48).upload(req)
    d_rep = ctx.device_buffer(n * 16)
    ctx.dispatch_device(d_req, n, d_rep)
    whole = ctx.reply_checksum(d_rep, n, device=True)
    got = d_rep.download(p.REPLY_DTYPE, n)
    want = snap.dispatch(req, threads=8)
    assert_same(got, want)
    assert whole == oracle.reply_checksum(want)
    # shard the stream 4 ways by request-id hash; the order-independent checksum adds up
    total = 0
    for rank in range(4):
        idx = shard_indices(n, 4, rank)
        part = ctx.dispatch(req[idx])
        assert_same(part, want[idx])
        # shard terms are indexed by their position in the whole stream
        total = (total + sum_terms(oracle, part, idx)) & ((1 << 64) - 1)
    assert total == whole
    d_req.free()
    d_rep.free()


def sum_terms(oracle, part, idx):
    """Checksum of shard replies under their global stream positions (numpy restatement of
    the definition in include/ptype_b200.h; checked against the oracle on a prefix)."""
    from ptype_b200.sharding import mix64
    with np.errstate(over="ignore"):
        pos = idx.astype(np.uint64)
        w0 = part["value"].view(np.uint64)
        w1 = part["node"].astype(np.uint64) | (part["status"].astype(np.uint64) << np.uint64(32))
        t = mix64(w0 + np.uint64(0x9E3779B97F4A7C15) * (pos + np.uint64(1))) ^ \
            mix64(w1 ^ (pos * np.uint64(0xC2B2AE3D27D4EB4F) + np.uint64(0x165667B19E3779F9)))
        total = int(t.sum(dtype=np.uint64))
    assert int(t[:1].sum(dtype=np.uint64)) == oracle.reply_checksum(part[:1], int(idx[0]))
    return total


def test_c3_mesh_variant(ctx, oracle):
    spec = w.stress_registry(2_000, 64, mesh=True)
    ctx.publish(spec.image())
    kv, snap = oracle.snapshot_from_spec(spec)
    req = w.stress_calls(1 << 20, 2_000)
    assert_same(device_roundtrip(ctx, req), snap.dispatch(req, threads=8), req)


def test_c4_optimus_replay(ctx, oracle):
    """BASELINE config 4: the coordinator's splitWork stream (>= 2^20 Prime.Check calls)."""
    spec = w.optimus_registry()
    ctx.publish(spec.image())
    kv, snap = oracle.snapshot_from_spec(spec)
    req = w.optimus_calls(1 << 20)
    want = snap.dispatch(req, threads=8)
    got = device_roundtrip(ctx, req)
    assert_same(got, want, req)
    # watchReplies (coordinator.go:91-99): a target is composite iff some chunk reply != target
    composite = got["value"] != req["a2"]
    assert composite.any() and (~composite).any()
    assert np.all(req["a2"][composite] % got["value"][composite] == 0)


def test_stats_and_failed_list(ctx, oracle):
    spec = w.calculator_registry()
    spec.declare = ["declared_but_empty"]
    ctx.publish(spec.image())
    kv, snap = oracle.snapshot_from_spec(spec)
    req = np.concatenate([w.edge_case_calls(["calculator", "declared_but_empty"]),
                          w.calculator_calls(300_000)])
    want = snap.dispatch(req, check_limit=LIMIT)
    n_fail = int((want["status"] != 0).sum())
    rep, failed, stats = ctx.dispatch_ex(req, failed_cap=n_fail + 10)
    assert_same(rep, want, req)
    assert stats.calls == len(req) and stats.failed_count == n_fail
    assert list(stats.by_status) == np.bincount(want["status"], minlength=8).tolist()
    assert np.array_equal(np.sort(failed), np.flatnonzero(want["status"] != 0))
    # a list shorter than the failures: count still exact, entries valid and distinct
    rep, failed, stats = ctx.dispatch_ex(req, failed_cap=7)
    assert stats.failed_count == n_fail and len(failed) == 7 and len(set(failed.tolist())) == 7
    assert np.all(want["status"][failed] != 0)
    # device-pointer flavour
    n = len(req)
    d_req, d_rep = ctx.device_buffer(n * 48).upload(req), ctx.device_buffer(n * 16)
    d_failed, d_stats = ctx.device_buffer((n_fail + 10) * 4), ctx.device_buffer(80)
    ctx.dispatch_device(d_req, n, d_rep, failed=d_failed, failed_cap=n_fail + 10, stats=d_stats)
    words = d_stats.download(np.uint64, 10)
    assert words[0] == n and words[9] == n_fail
    assert np.array_equal(np.sort(d_failed.download(np.uint32, n_fail)), np.flatnonzero(want["status"] != 0))
    assert_same(d_rep.download(p.REPLY_DTYPE, n), want)
    for b in (d_req, d_rep, d_failed, d_stats):
        b.free()


@pytest.mark.parametrize("n", [0, 1, 33, 100_000, (1 << 20) + 3])
def test_reply_checksum(ctx, oracle, n):
    rng = np.random.default_rng(n)
    rep = np.zeros(n, dtype=p.REPLY_DTYPE)
    rep["value"] = rng.integers(-(1 << 63), (1 << 63) - 1, n, dtype=np.int64)
    rep["node"] = rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32)
    rep["status"] = rng.integers(0, 8, n, dtype=np.uint32)
    for base in (0, 12345, (1 << 63) + 5):
        assert ctx.reply_checksum(rep, index_base=base) == oracle.reply_checksum(rep, base)
    if n > 1:
        swapped = rep.copy()
        swapped[[0, n - 1]] = swapped[[n - 1, 0]]
        if swapped[0] != rep[0]:
            assert ctx.reply_checksum(swapped) != ctx.reply_checksum(rep)  # position matters


def test_snapshot_flip_between_batches(ctx, oracle):
    """publish -> epoch flip mirrors handleNewNodes' swap (cluster/rpc.go:226-236): batches
    before the flip see the old node list, batches after see the new one."""
    req = w.calculator_calls(50_000)
    spec_a = w.calculator_registry(16)
    spec_b = w.calculator_registry(5)           # nodes 06..16 deregistered
    for epoch, spec in ((11, spec_a), (12, spec_b), (13, spec_a), (14, spec_b)):
        ctx.publish(spec.image(epoch=epoch))
        assert ctx.epoch() == epoch
        kv, snap = oracle.snapshot_from_spec(spec)
        assert ctx.image.selected("calculator") == snap.selected("calculator")
        assert_same(ctx.dispatch(req), snap.dispatch(req), req)
    # an image with no services at all: every call gets NO_SERVICE
    ctx.publish(w.RegistrySpec([], [], [], "", 3).image(epoch=15))
    rep = ctx.dispatch(req[:1000])
    assert np.all(rep["status"] == p.ST_NO_SERVICE) and np.all(rep["node"] == p.NO_NODE)
    assert np.all(rep["value"] == 0)


def test_live_updates_without_stopping_the_stream(ctx, oracle):
    """BASELINE config 5 in miniature: snapshots are published while batches are in flight
    (no host synchronisation in between); every batch is answered from the snapshot that was
    current when it was submitted - the epoch flips at batch boundaries (rpc.go:233-236)."""
    base = w.stress_registry(400, 6)
    b = p.ImageBuilder(base.local_addr, base.max_connections)
    for svc, node, addr, port in base.nodes:
        b.register(svc, node, addr, port)
    for svc, recv in base.expose:
        b.expose(svc, recv)
    b.set_device_only(True)
    nodes = list(base.nodes)
    n, per_epoch = 200_000, 4
    images, d_reqs, d_reps, wants = [], [], [], []
    rng = np.random.default_rng(5)
    for epoch in range(1, 7):
        if epoch > 1:     # one node of a random service leaves, or comes back
            s = int(rng.integers(0, 400))
            name = (f"svc_{s:05d}", f"node_{int(rng.integers(0, 6)):02d}")
            hit = [x for x in nodes if (x[0], x[1]) == name]
            if hit:
                nodes.remove(hit[0])
                b.unregister(*name)
            else:
                nodes.append((name[0], name[1], "10.9.9.9", 9999))
                b.register(name[0], name[1], "10.9.9.9", 9999)
        images.append(b.finish(epoch))
        spec = w.RegistrySpec(sorted(nodes), base.expose, [], base.local_addr, base.max_connections)
        kv, snap = oracle.snapshot_from_spec(spec)
        for k in range(per_epoch):
            req = w.stress_calls(n, 400, seed=1000 * epoch + k)
            d_reqs.append(ctx.device_buffer(n * 48).upload(req))
            d_reps.append(ctx.device_buffer(n * 16))
            wants.append(snap.dispatch(req, threads=8))
    assert len({img.blob[192:img.header.device_bytes] for img in images}) == len(images)
    # everything is resident: now publish and submit back to back, no synchronisation
    for e, img in enumerate(images):
        ctx.publish(img)
        lo = e * per_epoch
        ctx.dispatch_batches(ctx.make_batches(zip(d_reqs[lo:lo + per_epoch], d_reps[lo:lo + per_epoch],
                                                  [n] * per_epoch)), sync=False)
    assert ctx.epoch() == len(images)
    ctx.sync()
    for d, want in zip(d_reps, wants):
        assert_same(d.download(p.REPLY_DTYPE, n), want)
    for d in d_reqs + d_reps:
        d.free()


def test_update_service_patches_the_table_on_the_device(ctx, oracle):
    """ptype_b200_update_service: one node joins or leaves; the table is patched on the device
    (node count of the service, node ids of every service after it) and batches submitted after
    the update - with none of them synchronised - are answered as from a freshly built image."""
    base = w.stress_registry(300, 5)
    base.declare = ["svc_00100x_empty"]
    ctx.publish(base.image(epoch=1))
    nodes = {(s, nd): (s, nd, a, pt) for s, nd, a, pt in base.nodes}
    rng = np.random.default_rng(11)
    n, rounds = 100_000, 12
    d_reqs, d_reps, wants = [], [], []
    plan = []
    for k in range(rounds):
        svc = f"svc_{int(rng.integers(0, 300)):05d}" if k != 3 else "svc_00100x_empty"
        present = sorted(nd for (s, nd) in nodes if s == svc)
        if present and (rng.integers(0, 2) == 0 or len(present) >= 5):
            del nodes[(svc, present[int(rng.integers(0, len(present)))])]
            plan.append((svc, -1))
        else:
            name = f"node_{int(rng.integers(5, 99)):02d}"
            nodes[(svc, name)] = (svc, name, "10.7.7.7", 7777)
            plan.append((svc, +1))
        spec = w.RegistrySpec(sorted(nodes.values()), base.expose + [("svc_00100x_empty", "Calculator")],
                              ["svc_00100x_empty"], base.local_addr, base.max_connections)
        kv, snap = oracle.snapshot_from_spec(spec)
        req = w.stress_calls(n, 300, seed=50 + k)
        req["svc_key"][::97] = p.service_key("svc_00100x_empty")
        d_reqs.append(ctx.device_buffer(n * 48).upload(req))
        d_reps.append(ctx.device_buffer(n * 16))
        wants.append(snap.dispatch(req, threads=8))
    # the image does not expose Calculator on the empty service: rebuild the first image with it
    first = w.RegistrySpec(base.nodes, base.expose + [("svc_00100x_empty", "Calculator")],
                           ["svc_00100x_empty"], base.local_addr, base.max_connections)
    ctx.publish(first.image(epoch=1))
    for k, (svc, delta) in enumerate(plan):
        ctx.update_service(svc, delta)
        assert ctx.epoch() == 2 + k
        ctx.dispatch_batches(ctx.make_batches([(d_reqs[k], d_reps[k], n)]), sync=False)
    ctx.sync()
    for k in range(rounds):
        assert_same(d_reps[k].download(p.REPLY_DTYPE, n), wants[k])
    img = ctx.current_image()            # patched host copy: valid, device-only
    assert img.epoch == 1 + rounds and img.header.host_sections == 0 and img.header.n_nodes == len(nodes)
    with pytest.raises(p.PtypeError):
        ctx.update_service("never_registered", +1)
    with pytest.raises(p.PtypeError):
        ctx.update_service("svc_00000", -1000)
    for d in d_reqs + d_reps:
        d.free()


def test_max_connections_larger_than_shared_cache(ctx, oracle):
    """MaxConnections above the 256 hashes kept in shared memory: hashes read from global."""
    nodes = [("wide", f"n{i:04d}", f"10.1.{i >> 8}.{i & 255}", 7000) for i in range(700)]
    spec = w.RegistrySpec(nodes, [("wide", "Calculator")], [], "10.9.9.9", 300)
    ctx.publish(spec.image())
    kv, snap = oracle.snapshot_from_spec(spec)
    req = w.calculator_calls(20_000)
    req["svc_key"] = p.service_key("wide")
    assert_same(ctx.dispatch(req), snap.dispatch(req), req)


def test_colliding_table_slots(ctx, oracle):
    """A table forced to ~89 % load: keys displaced to their second and third candidate
    slots, and misses that probe all three."""
    spec = w.stress_registry(500, 2)
    img = spec.image(n_slots=560)
    ctx.publish(img)
    kv, snap = oracle.snapshot_from_spec(spec)
    req = w.stress_calls(100_000, 500, miss_rate=0.2)
    for flags in VARIANTS:
        assert_same(device_roundtrip(ctx, req, flags=flags), snap.dispatch(req), req)


def test_calls_from_many_threads(ctx, oracle):
    """cgo calls arrive on arbitrary OS threads: no thread-local CUDA state in the library."""
    spec = w.calculator_registry()
    ctx.publish(spec.image())
    kv, snap = oracle.snapshot_from_spec(spec)
    errors = []

    def worker(seed):
        try:
            req = w.calculator_calls(70_000, seed=seed)
            assert_same(ctx.dispatch(req), snap.dispatch(req), req)
        except Exception as e:  # noqa: BLE001
            errors.append(e)

    ts = [threading.Thread(target=worker, args=(100 + i,)) for i in range(6)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errors, errors


def test_dispatch_batches_equals_one_by_one(ctx, oracle):
    """Many device-resident batches enqueued by one C call (kernels overlap through
    programmatic dependent launch) give the replies of separate dispatches."""
    spec = w.calculator_registry()
    ctx.publish(spec.image())
    kv, snap = oracle.snapshot_from_spec(spec)
    sizes = [1 << 18, 5, 0, 77_777, 1 << 20, 33, 1 << 16] * 3
    reqs = [w.calculator_calls(n, seed=900 + i, seq_base=i * 1000) for i, n in enumerate(sizes)]
    d_reqs = [ctx.device_buffer(max(n, 1) * 48).upload(r) for n, r in zip(sizes, reqs)]
    d_reps = [ctx.device_buffer(max(n, 1) * 16) for n in sizes]
    batches = ctx.make_batches(zip(d_reqs, d_reps, sizes))
    for flags in (0, p.F_ONE_LAUNCH_PER_BATCH, p.F_BULK_LOADS, p.F_GLOBAL_TABLE):
        for d in d_reps:                       # poison, so stale replies cannot pass
            d.upload(np.full(4, 0xAB, dtype=np.uint8))
        ctx.dispatch_batches(batches, flags=flags, sync=False)
        ctx.sync()
        for n, r, d in zip(sizes, reqs, d_reps):
            assert_same(d.download(p.REPLY_DTYPE, n), snap.dispatch(r), r)
    for b in d_reqs + d_reps:
        b.free()


@pytest.mark.parametrize("shape", ["pipelined_256", "unrolled_mesh", "pipelined_1024"])
def test_ragged_batch_lists_in_every_multi_batch_loop(ctx, oracle, shape):
    """130 batches (more than one launch holds) of 0..40, 32k+1 and odd sizes, all carved out of ONE
    request buffer and ONE reply buffer with 16 guard bytes between neighbours: every loop of the
    multi-batch kernel has to find each batch's tiles, stop at its last call and leave the bytes
    between the reply regions alone."""
    if shape == "pipelined_256":
        spec, n_services = w.calculator_registry(), None
    elif shape == "unrolled_mesh":
        spec, n_services = w.calculator_registry(mesh=True), None
    else:
        spec, n_services = w.stress_registry(2000, 5), 2000      # 62 KiB table: one 1024-thread CTA per SM
    ctx.publish(spec.image())
    kv, snap = oracle.snapshot_from_spec(spec)
    rng = np.random.default_rng(7)
    sizes = [int(x) for x in rng.integers(0, 41, size=100)] + [32, 64, 31, 33, 1, 0, 4096 + 1, 32 * 1024 + 1,
                                                              100_003, 65_536] + [int(x) for x in rng.integers(1, 3000, size=20)]
    assert len(sizes) == 130
    reqs = [w.stress_calls(n, n_services, seed=40 + i) if n_services else w.calculator_calls(n, seed=40 + i, seq_base=i)
            for i, n in enumerate(sizes)]
    total = sum(sizes)
    d_req = ctx.device_buffer(total * 48 + 48).upload(np.concatenate(reqs))
    guard = 16
    rep_off = np.cumsum([0] + [n * 16 + guard for n in sizes])
    d_rep = ctx.device_buffer(int(rep_off[-1]) + 16)
    poison = np.full(int(rep_off[-1]) + 16, 0xAB, dtype=np.uint8)
    req_off = np.cumsum([0] + [n * 48 for n in sizes])
    triples = [(d_req.ptr + int(req_off[i]), d_rep.ptr + int(rep_off[i]), n) for i, n in enumerate(sizes)]
    want = [snap.dispatch(r) for r in reqs]
    for flags in (0, p.F_ONE_LAUNCH_PER_BATCH):
        d_rep.upload(poison)
        ctx.dispatch_batches(ctx.make_batches(triples), flags=flags)
        got = d_rep.download(np.uint8, len(poison))
        for i, n in enumerate(sizes):
            lo = int(rep_off[i])
            assert got[lo:lo + n * 16].tobytes() == want[i].tobytes(), (shape, flags, i, n)
            assert (got[lo + n * 16:lo + n * 16 + guard] == 0xAB).all(), (shape, flags, i, n)
    d_req.free()
    d_rep.free()


@pytest.mark.parametrize("sizes", [(700_000, 300_017, 50_000), (1 << 20, 1 << 20), (20_000, 1_000_000, 31, 1),
                                   (524_288, 1) + (40_000,) * 20])
def test_lists_whose_end_runs_as_a_second_launch(ctx, oracle, sizes):
    """A list of 2..48 batches is cut in two launches (csrc/dispatch.cu, launch_dispatch_multi): the
    last 8 % of its tiles go to short CTAs behind the persistent grid.  The cut may fall inside a
    batch (on a tile boundary), on a batch boundary, or leave tiny batches whole on either side;
    every reply must be there exactly once and the guard bytes between the regions untouched."""
    spec = w.calculator_registry()
    ctx.publish(spec.image())
    kv, snap = oracle.snapshot_from_spec(spec)
    reqs = [w.calculator_calls(n, seed=7000 + i, seq_base=i << 21) for i, n in enumerate(sizes)]
    guard = 32
    req_off = np.cumsum([0] + [n * 48 for n in sizes])
    rep_off = np.cumsum([0] + [n * 16 + guard for n in sizes])
    d_req = ctx.device_buffer(int(req_off[-1]) + 48).upload(np.concatenate(reqs))
    d_rep = ctx.device_buffer(int(rep_off[-1]))
    d_rep.upload(np.full(int(rep_off[-1]), 0xCD, dtype=np.uint8))
    ctx.dispatch_batches(ctx.make_batches(
        (d_req.ptr + int(req_off[i]), d_rep.ptr + int(rep_off[i]), n) for i, n in enumerate(sizes)), sync=False)
    ctx.sync()
    got = d_rep.download(np.uint8, int(rep_off[-1]))
    for i, n in enumerate(sizes):
        lo = int(rep_off[i])
        assert got[lo:lo + n * 16].tobytes() == snap.dispatch(reqs[i], threads=8).tobytes(), (i, n)
        assert (got[lo + n * 16:lo + n * 16 + guard] == 0xCD).all(), (i, n)
    d_req.free()
    d_rep.free()


@pytest.mark.parametrize("n", [1, 33, 4097, 128 * 1024 + 1, 1 << 20])
@pytest.mark.parametrize("seq_base", [0, (1 << 32) - 7, (1 << 63) - 5])
def test_call_batch_compact_form(ctx, oracle, n, seq_base):
    """ptype_b200_call_batch (one Client's batch: args + a method byte per call) gives the
    replies of dispatching the equivalent 48-byte records, for host and device buffers."""
    spec = w.calculator_registry()
    ctx.publish(spec.image())
    kv, snap = oracle.snapshot_from_spec(spec)
    reqs = w.calculator_calls(n, seed=31 + n, seq_base=seq_base)
    want = snap.dispatch(reqs)
    key, base, methods, sel, args = w.as_client_batch(reqs)
    assert base == seq_base and args.shape == (n, 2) and (sel is not None or n == 1)
    assert_same(ctx.call_batch(key, base, methods, args, sel), want, reqs)
    d_args = ctx.device_buffer(args.nbytes).upload(args)
    d_sel = ctx.device_buffer(max(n, 1)).upload(sel) if sel is not None else None
    d_rep = ctx.device_buffer(n * 16)
    ctx.call_batch(key, base, methods, d_args, d_sel, d_rep, n=n, args_per_call=2)
    assert_same(d_rep.download(p.REPLY_DTYPE, n), want, reqs)
    for b in (d_args, d_sel, d_rep):
        if b is not None:
            b.free()


def test_call_batch_prime_and_errors(ctx, oracle):
    spec = w.optimus_registry()
    ctx.publish(spec.image())
    kv, snap = oracle.snapshot_from_spec(spec)
    reqs = w.optimus_calls(300_000)
    key, base, methods, sel, args = w.as_client_batch(reqs)
    assert sel is None and args.shape[1] == 3
    assert_same(ctx.call_batch(key, base, methods, args), snap.dispatch(reqs, threads=8), reqs)
    # a method the nodes never registered, an ill-formed name, and a service nobody registered
    for sm, status in (("Prime.Sieve", p.ST_CANT_FIND_METHOD), ("Calculator.Add", p.ST_CANT_FIND_SERVICE),
                       ("nodot", p.ST_ILL_FORMED)):
        rep = ctx.call_batch(key, 0, [p.method_key(sm)], args[:1000])
        assert np.all(rep["status"] == status) and np.all(rep["value"] == 0)
        assert np.all(rep["node"] == np.array(snap.selected("prime_worker"))[np.arange(1, 1001) % 2])
    rep = ctx.call_batch(p.service_key("nobody"), 0, methods, args[:1000])
    assert np.all(rep["status"] == p.ST_NO_SERVICE) and np.all(rep["node"] == p.NO_NODE)


@pytest.mark.parametrize("kind", ["pageable", "pinned", "device"])
def test_call_batch_values_only(ctx, oracle, kind):
    """ptype_b200_call_batch_values: 8-byte values plus an (index, status) list of the failed
    calls - what CallBatch's caller reads - equal the full reply records' value/status."""
    spec = w.calculator_registry()
    ctx.publish(spec.image())
    kv, snap = oracle.snapshot_from_spec(spec)
    n = 300_017
    reqs = w.calculator_calls(n, seed=77)
    bad = np.arange(0, n, 9973)
    reqs["method"][bad] = 0xDEAD0000 | (bad & 0xFFFF).astype(np.uint32) % 2   # two unknown methods
    want = snap.dispatch(reqs)
    assert (want["status"] != 0).sum() == len(bad)
    key, base, methods, sel, args = w.as_client_batch(reqs)
    assert len(methods) == 4
    if kind == "pageable":
        values, failed, count = ctx.call_batch_values(key, base, methods, args, sel, failed_cap=64)
    elif kind == "pinned":
        pa, ps, pv = p.PinnedArray(args.size, np.int64), p.PinnedArray(n, np.uint8), p.PinnedArray(n, np.int64)
        pa.array[:] = args.reshape(-1)
        ps.array[:] = sel
        launches = ctx.launch_count()
        _, failed, count = ctx.call_batch_values(key, base, methods, pa.array.reshape(args.shape), ps.array,
                                                 pv.array, failed_cap=64)
        assert ctx.launch_count() == launches + 1
        values = pv.array.copy()
        for b in (pa, ps, pv):
            b.free()
    else:
        d_args = ctx.device_buffer(args.nbytes).upload(args)
        d_sel = ctx.device_buffer(n).upload(sel)
        d_val = ctx.device_buffer(n * 8)
        _, failed, count = ctx.call_batch_values(key, base, methods, d_args, d_sel, d_val, failed_cap=64,
                                                 n=n, args_per_call=2)
        values = d_val.download(np.int64, n)
        for b in (d_args, d_sel, d_val):
            b.free()
    assert np.array_equal(values, want["value"])
    assert count == len(bad) and len(failed) == len(bad)
    order = np.argsort(failed["index"])
    assert np.array_equal(failed["index"][order], bad.astype(np.uint32))
    assert np.array_equal(failed["status"][order], want["status"][bad])
    # a list shorter than the failures: exact count, valid distinct entries
    _, few, count = ctx.call_batch_values(key, base, methods, args, sel, failed_cap=5)
    assert count == len(bad) and len(few) == 5 and len(set(few["index"].tolist())) == 5
    assert np.all(np.isin(few["index"], bad))


def test_call_batch_against_the_big_table(ctx, stress, oracle):
    """The compact form with the 12 288-slot table staged in shared memory (1024-thread CTAs)."""
    spec, img, snap, _ = stress
    ctx.publish(img)
    n = 500_000
    reqs = w.calculator_calls(n, seed=5)
    reqs["svc_key"] = p.service_key("svc_01234")
    key, base, methods, sel, args = w.as_client_batch(reqs)
    assert_same(ctx.call_batch(key, base, methods, args, sel), snap.dispatch(reqs, threads=8), reqs)
    assert_same(ctx.call_batch(key, base, methods, args, sel, flags=p.F_GLOBAL_TABLE),
                snap.dispatch(reqs, threads=8), reqs)


def test_pinned_buffers_take_the_zero_copy_path(ctx, oracle):
    """Page-locked host buffers (ptype_b200_alloc_pinned) are read and written by the kernel
    across PCIe directly; results equal the staged-copy path's and the oracle's."""
    spec = w.calculator_registry()
    ctx.publish(spec.image())
    kv, snap = oracle.snapshot_from_spec(spec)
    n = 300_001
    reqs = np.concatenate([w.calculator_calls(n - 5000), w.edge_case_calls(["calculator"])])[:n]
    want = snap.dispatch(reqs, check_limit=LIMIT)
    pq, pr = p.PinnedArray(n, p.REQ_DTYPE), p.PinnedArray(n, p.REPLY_DTYPE)
    pq.array[:] = reqs
    launches = ctx.launch_count()
    ctx.dispatch(pq.array, pr.array)
    assert ctx.launch_count() == launches + 1          # one launch, no chunk pipeline
    assert_same(pr.array.copy(), want, reqs)
    assert_same(ctx.dispatch(reqs), want, reqs)         # pageable: staged copies, 3 chunks
    cal = w.calculator_calls(n, seed=9)
    key, base, methods, sel, args = w.as_client_batch(cal)
    pa, ps = p.PinnedArray(args.size, np.int64), p.PinnedArray(n, np.uint8)
    pa.array[:] = args.reshape(-1)
    ps.array[:] = sel
    pr.array[:] = np.zeros(1, dtype=p.REPLY_DTYPE)
    ctx.call_batch(key, base, methods, pa.array.reshape(args.shape), ps.array, pr.array)
    assert_same(pr.array.copy(), snap.dispatch(cal), cal)
    for b in (pq, pr, pa, ps):
        b.free()


def test_misaligned_buffers_are_rejected(ctx):
    """Record arrays must be 16-byte aligned (128-bit loads/stores): a misaligned pointer is an
    argument error, not a device fault."""
    from ptype_b200._abi import lib, E_INVAL, F_DEVICE_PTRS, Batch
    ctx.publish(w.calculator_registry().image())
    d = ctx.device_buffer(4096)
    assert lib.ptype_b200_dispatch(ctx._h, 0, d.ptr + 8, 10, d.ptr + 1024, F_DEVICE_PTRS) == E_INVAL
    assert lib.ptype_b200_dispatch(ctx._h, 0, d.ptr, 10, d.ptr + 1032, F_DEVICE_PTRS) == E_INVAL
    one = (Batch * 1)()
    one[0].reqs, one[0].replies, one[0].n = d.ptr + 4, d.ptr + 1024, 10
    assert lib.ptype_b200_dispatch_batches(ctx._h, 0, one, 1, F_DEVICE_PTRS) == E_INVAL
    d.free()
    ctx.dispatch(w.calculator_calls(10))        # the context is still usable


def test_dispatch_without_image_fails():
    c = p.Context([0])
    try:
        with pytest.raises(p.PtypeError) as e:
            c.dispatch(w.calculator_calls(10))
        assert e.value.code == 4  # PTYPE_E_NO_IMAGE
    finally:
        c.close()
