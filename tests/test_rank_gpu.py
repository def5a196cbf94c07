"""The balancer's counters on the device (ptype_b200_assign_seq) against the
oracle's restatement of N sequential Client.Call()s (rpc.go:129,172-183): seq fields and counters
bit-exact, through batches, registry flips and table re-placements."""
import numpy as np
import pytest

import ptype_b200 as p
import workloads as w
from test_fuzz_gpu import random_spec

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def rctx():
    """A context of its own: once counters are in use every flip also carries them over."""
    c = p.Context([0])
    yield c
    c.close()


@pytest.fixture(scope="module")
def stress(oracle):
    spec = w.stress_registry()
    img = spec.image()
    kv, snap = oracle.snapshot_from_spec(spec, img.salt)
    return spec, img, kv, snap


def device_assign(ctx, reqs):
    n = len(reqs)
    d = ctx.device_buffer(max(n, 1) * 48).upload(reqs)
    ctx.assign_seq(d, n)
    out = d.download(p.REQ_DTYPE, n)
    d.free()
    return out


def check_counters(ctx, snap, counters, services):
    for svc in services:
        i = snap.client_index(svc)
        assert ctx.seq_get(svc) == (int(counters[i]) if i >= 0 else 0), svc


def test_single_service_batch_counts_from_one(rctx, oracle):
    """C2: one Client, 2^20 calls - every lane of every warp lands on the same counter."""
    spec = w.calculator_registry()
    img = spec.image()
    rctx.publish(img)
    reqs = w.calculator_calls(1 << 20, salt=img.salt)
    want = reqs["seq"].copy()           # the generator's 1..N
    reqs["seq"] = 0xDEADBEEF
    got = device_assign(rctx, reqs)
    assert np.array_equal(got["seq"], want)
    assert rctx.seq_get("calculator") == 1 << 20
    other = got.copy()
    other["seq"] = reqs["seq"]
    assert other.tobytes() == reqs.tobytes()          # nothing but seq was touched
    again = device_assign(rctx, reqs[:1000])          # the counter persists: next batch continues
    assert np.array_equal(again["seq"], want[:1000] + np.uint64(1 << 20))


@pytest.mark.parametrize("n", [0, 1, 31, 32, 33, 255, 256, 257, 5000, 1 << 18])
def test_stress_registry_ragged_sizes(stress, n):
    """C3 registry (10 000 services, 12 288 slots): uniform services, 1 % unregistered."""
    spec, img, kv, snap = stress
    with p.Context([0]) as ctx:          # counters start at zero, like the oracle's
        ctx.publish(img)
        counters = np.zeros(snap.n_clients(), dtype=np.uint64)
        rng = np.random.default_rng(n)
        for svc in ("svc_00000", "svc_09999", "svc_01234"):
            v = int(rng.integers(0, 1 << 62))
            ctx.seq_set(svc, v)
            counters[snap.client_index(svc)] = v
        for rep in range(2):
            reqs = w.stress_calls(n, seed=77 + rep, salt=img.salt)
            reqs["seq"] = 5
            want = reqs.copy()
            snap.assign_seq(want, counters)
            got = device_assign(ctx, reqs)
            assert got.tobytes() == want.tobytes(), (n, rep)
        check_counters(ctx, snap, counters, ["svc_00000", "svc_09999", "svc_01234", "svc_00500"])


def test_stress_2_24_and_dispatch(stress):
    """BASELINE config 3 at full size: ranks on the device, then the dispatch that uses them."""
    spec, img, kv, snap = stress
    n = 1 << 24
    reqs = w.stress_calls(n, salt=img.salt)
    reqs["seq"] = 0
    want = reqs.copy()
    counters = np.zeros(snap.n_clients(), dtype=np.uint64)
    snap.assign_seq(want, counters)
    assert counters.sum() == (want["seq"] != 0).sum()
    with p.Context([0]) as ctx:
        ctx.publish(img)
        d_req = ctx.device_buffer(n * 48).upload(reqs)
        d_rep = ctx.device_buffer(n * 16)
        ctx.assign_seq(d_req, n, sync=False)
        ctx.dispatch_device(d_req, n, d_rep)
        got = d_req.download(p.REQ_DTYPE, n)
        assert np.array_equal(got["seq"], want["seq"])
        rep = d_rep.download(p.REPLY_DTYPE, n)
        assert rep.tobytes() == snap.dispatch(want, threads=8).tobytes()
        d_req.free()
        d_rep.free()
        for svc in ("svc_00000", "svc_04321", "svc_09999"):
            assert ctx.seq_get(svc) == int(counters[snap.client_index(svc)])


@pytest.mark.parametrize("seed", range(6))
def test_random_registries(oracle, seed):
    """Awkward names, services without nodes (they consume nothing), unknown services."""
    rng = np.random.default_rng(4200 + seed)
    spec, names = random_spec(rng)
    img = spec.image()
    with p.Context([0]) as ctx:
        ctx.publish(img)
        kv, snap = oracle.snapshot_from_spec(spec, img.salt)
        counters = np.zeros(snap.n_clients(), dtype=np.uint64)
        keys = np.array([p.service_key(s, img.salt) for s in names + ["no such service"]], dtype=np.uint64)
        for n in (1, 777, 40_000):
            reqs = np.zeros(n, dtype=p.REQ_DTYPE)
            reqs["svc_key"] = keys[rng.integers(0, len(keys), n)]
            reqs["a0"] = rng.integers(-5, 5, n)
            want = reqs.copy()
            snap.assign_seq(want, counters)
            got = device_assign(ctx, reqs)
            assert got.tobytes() == want.tobytes(), (seed, n)
        check_counters(ctx, snap, counters, names)
        with pytest.raises(p.PtypeError):
            ctx.seq_get("no such service")


def test_counters_follow_their_services_through_flips(oracle):
    """update_service keeps the placement, a rebuilt image with more services changes it, and a
    broadcast-free publish of an image without a service drops that service's counter."""
    spec = w.RegistrySpec([(f"s{i:03d}", f"n{j}", "h", 1000 + j) for i in range(200) for j in range(3)],
                          [], [], "10.0.0.5", 3)
    img = spec.image()
    with p.Context([0]) as ctx:
        ctx.publish(img)
        kv, snap = oracle.snapshot_from_spec(spec, img.salt)
        counters = np.zeros(snap.n_clients(), dtype=np.uint64)
        rng = np.random.default_rng(5)
        keys = np.array([p.service_key(f"s{i:03d}", img.salt) for i in range(200)], dtype=np.uint64)

        def batch(n):
            reqs = np.zeros(n, dtype=p.REQ_DTYPE)
            reqs["svc_key"] = keys[rng.integers(0, 200, n)]
            return reqs

        def step(snap_now, counters_now, n=20_000):
            reqs = batch(n)
            want = reqs.copy()
            snap_now.assign_seq(want, counters_now)
            assert device_assign(ctx, reqs).tobytes() == want.tobytes()

        step(snap, counters)
        # a node leaves s007 until none is left: its calls stop consuming the counter
        for _ in range(3):
            ctx.update_service("s007", -1)
        for j in range(3):
            assert kv.delete(f"services/s007/n{j}/") == 1
        snap2 = oracle.Snapshot(kv, spec.local_addr, spec.max_connections, ["s007"], img.salt)
        c2 = np.zeros(snap2.n_clients(), dtype=np.uint64)
        for i in range(200):
            c2[snap2.client_index(f"s{i:03d}")] = counters[snap.client_index(f"s{i:03d}")]
        frozen = int(c2[snap2.client_index("s007")])
        step(snap2, c2)
        assert ctx.seq_get("s007") == frozen > 0
        # 300 more services: a bigger table, every key placed anew
        nodes3 = [x for x in spec.nodes if x[0] != "s007"] + [(f"t{i:03d}", "n0", "h", 1) for i in range(300)]
        spec3 = w.RegistrySpec(nodes3, [], ["s007"], "10.0.0.5", 3)
        img3 = spec3.image(epoch=9)
        assert img3.salt == img.salt and img3.header.n_slots != img.header.n_slots
        ctx.publish(img3)
        kv3, snap3 = oracle.snapshot_from_spec(spec3, img3.salt)
        c3 = np.zeros(snap3.n_clients(), dtype=np.uint64)
        for i in range(200):
            c3[snap3.client_index(f"s{i:03d}")] = c2[snap2.client_index(f"s{i:03d}")]
        step(snap3, c3)
        assert ctx.seq_get("s007") == frozen and ctx.seq_get("t000") == 0
        for i in (0, 7, 42, 199):
            assert ctx.seq_get(f"s{i:03d}") == int(c3[snap3.client_index(f"s{i:03d}")])
        # an image that no longer knows s042: its counter is gone when the service comes back
        spec4 = w.RegistrySpec([x for x in nodes3 if x[0] != "s042"], [], ["s007"], "10.0.0.5", 3)
        ctx.publish(spec4.image(epoch=10))
        with pytest.raises(p.PtypeError):
            ctx.seq_get("s042")
        ctx.publish(img3)
        assert ctx.seq_get("s042") == 0 and ctx.seq_get("s041") == int(c3[snap3.client_index("s041")])


def test_tables_beyond_the_histogram_are_refused(oracle):
    spec = w.RegistrySpec([(f"s{i}", "n", "h", 1) for i in range(8)], [], [], "", 3)
    img = p.build_image(spec.nodes, n_slots=60_000)
    with p.Context([0]) as ctx:
        ctx.publish(img)
        d = ctx.device_buffer(48 * 4)
        with pytest.raises(p.PtypeError):
            ctx.assign_seq(d, 4)
        d.free()


@pytest.mark.parametrize("n_slots", [18_000, 19_000, 30_000, 51_200])
def test_large_tables_either_side_of_the_staging_limit(oracle, n_slots):
    """Up to ~18 500 slots the keys sit in shared memory beside the histogram; beyond that (to
    51 200) the table is probed in L2.  Same answers either way."""
    nodes = [(f"svc{i:04d}", f"n{j}", "h", 1) for i in range(1500) for j in range(1 + i % 2)]
    spec = w.RegistrySpec(nodes, [], ["idle"], "10.0.0.5", 3)
    img = p.build_image(spec.nodes, local_addr=spec.local_addr, max_connections=3, declare=spec.declare,
                        n_slots=n_slots)
    assert img.header.n_slots == n_slots
    kv, snap = oracle.snapshot_from_spec(spec, img.salt)
    rng = np.random.default_rng(n_slots)
    keys = np.array([p.service_key(f"svc{i:04d}", img.salt) for i in range(1500)] +
                    [p.service_key("idle", img.salt), p.service_key("nobody", img.salt)], dtype=np.uint64)
    with p.Context([0]) as ctx:
        ctx.publish(img)
        counters = np.zeros(snap.n_clients(), dtype=np.uint64)
        for n in (100_000, 3):
            reqs = np.zeros(n, dtype=p.REQ_DTYPE)
            reqs["svc_key"] = keys[rng.integers(0, len(keys), n)]
            want = reqs.copy()
            snap.assign_seq(want, counters)
            assert device_assign(ctx, reqs).tobytes() == want.tobytes(), (n_slots, n)
        assert ctx.seq_get("svc0007") == int(counters[snap.client_index("svc0007")]) > 0
        assert ctx.seq_get("idle") == 0


@pytest.mark.parametrize("n", [0, 1, 33, 5000, (1 << 20) + 7])
@pytest.mark.parametrize("mesh", [False, True])
def test_seq_array_form_equals_the_record_form(stress, oracle, n, mesh):
    """ptype_b200_assign_seq_array + ptype_b200_dispatch_seq: the numbers go to a plain array and
    the dispatch takes them from there.  Same numbers, same replies as writing them into the
    records and dispatching those - and the records are left as they were."""
    if mesh:   # the 256-thread shape with identity selection (the other loop of the multi-batch kernel)
        spec = w.calculator_registry(mesh=True)
        img = spec.image()
        kv, snap = oracle.snapshot_from_spec(spec, img.salt)
        reqs = w.calculator_calls(n, seed=3, salt=img.salt)
    else:
        spec, img, kv, snap = stress
        reqs = w.stress_calls(n, seed=31, salt=img.salt)
    reqs["seq"] = 0xABCD
    want = reqs.copy()
    counters = np.zeros(snap.n_clients(), dtype=np.uint64)
    snap.assign_seq(want, counters)
    with p.Context([0]) as ctx:
        ctx.publish(img)
        d_req = ctx.device_buffer(max(n, 1) * 48).upload(reqs)
        d_seq = ctx.device_buffer(max(n, 1) * 8)
        d_rep = ctx.device_buffer(max(n, 1) * 16)
        for rep in range(2):   # the counters carry on from the first batch to the second
            if rep:
                snap.assign_seq(want, counters)
            ctx.assign_seq_array(d_req, n, d_seq, sync=False)
            ctx.dispatch_seq(d_req, d_seq, n, d_rep)
            assert np.array_equal(d_seq.download(np.uint64, n), want["seq"])
            assert d_req.download(p.REQ_DTYPE, n).tobytes() == reqs.tobytes()
            assert d_rep.download(p.REPLY_DTYPE, n).tobytes() == snap.dispatch(want, threads=8).tobytes()
        for b in (d_req, d_seq, d_rep):
            b.free()
