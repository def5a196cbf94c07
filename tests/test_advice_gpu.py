"""GPU parity tests for corner cases a review of the first round found untested: every launch
shape around the staged-table sizes where the CTA shape changes, node-less neighbours under update_service, service
names whose keys fold alike, and dispatches that follow a producer on the same stream."""
import numpy as np
import pytest

import ptype_b200 as p
import workloads as w

pytestmark = pytest.mark.gpu


def assert_same(got: np.ndarray, want: np.ndarray):
    if got.tobytes() != want.tobytes():
        bad = np.flatnonzero((got["value"] != want["value"]) | (got["node"] != want["node"]) |
                             (got["status"] != want["status"]))
        j = int(bad[0])
        raise AssertionError(f"{len(bad)} of {len(got)} replies differ; first at {j}: got {got[j]} want {want[j]}")


@pytest.mark.parametrize("n_services", [1400, 1500, 2000, 3100])
def test_every_entry_point_on_mid_sized_tables(ctx, oracle, n_services):
    """Slot tables of 44-100 KiB: the largest that four 256-thread CTAs stage side by side (1400
    services), the smallest that goes to one 1024-thread CTA (1500), and the band above 48 KiB
    where a kernel needs the opt-in for large dynamic shared memory.  dispatch, dispatch_batches
    (both launch modes), call_batch and call_batch_values must all run there, not just dispatch."""
    spec = w.stress_registry(n_services, 4)
    image = spec.image()
    table_bytes = image.header.n_slots * 16
    assert 43 * 1024 < table_bytes <= 100 * 1024, table_bytes
    ctx.publish(image)
    kv, snap = oracle.snapshot_from_spec(spec)
    n = 300_001
    reqs = w.stress_calls(n, n_services, seed=n_services)
    want = snap.dispatch(reqs, threads=8)
    d_req = ctx.device_buffer(n * 48).upload(reqs)
    d_rep = ctx.device_buffer(n * 16)
    for flags in (0, p.F_ONE_LAUNCH_PER_BATCH, p.F_LDG_LOADS):
        d_rep.upload(np.full(64, 0xAB, dtype=np.uint8))
        ctx.dispatch_batches(ctx.make_batches([(d_req, d_rep, n), (d_req, d_rep, 0)]), flags=flags)
        assert_same(d_rep.download(p.REPLY_DTYPE, n), want)
    assert_same(ctx.dispatch(reqs), want)
    replies, failed, stats = ctx.dispatch_ex(reqs, failed_cap=n)
    assert_same(replies, want)
    assert stats.failed_count == int((want["status"] != 0).sum())
    # one Client of one of the services, compact form
    svc = f"svc_{n_services // 2:05d}"
    one = w.calculator_calls(200_003, seed=9)
    one["svc_key"] = p.service_key(svc)
    want_one = snap.dispatch(one, threads=8)
    key, base, methods, sel, args = w.as_client_batch(one)
    assert_same(ctx.call_batch(key, base, methods, args, sel), want_one)
    values, vfailed, vcount = ctx.call_batch_values(key, base, methods, args, sel)
    assert np.array_equal(values, want_one["value"]) and vcount == int((want_one["status"] != 0).sum())
    d_args = ctx.device_buffer(args.nbytes).upload(args)
    d_sel = ctx.device_buffer(len(sel)).upload(sel)
    d_out = ctx.device_buffer(len(one) * 16)
    ctx.call_batch(key, base, methods, d_args, d_sel, d_out, n=len(one), args_per_call=2)
    assert_same(d_out.download(p.REPLY_DTYPE, len(one)), want_one)
    for b in (d_req, d_rep, d_args, d_sel, d_out):
        b.free()


def test_update_service_with_node_less_neighbours(ctx, oracle):
    """Two adjacent services lose their only node, then get one back in reverse order.  While
    both are empty they share a node base, so only their key order says whose ids move: every
    intermediate table must answer like an image rebuilt from the changed registry."""
    names = ["aaa", "bbb", "ccc", "ddd", "eee"]
    reg = {(svc, node): (svc, node, "10.1.1.1", 4000)
           for svc, node in [("aaa", "n1"), ("aaa", "n2"), ("bbb", "n1"), ("ccc", "n1"), ("ddd", "n1"),
                             ("ddd", "n2"), ("ddd", "n3"), ("eee", "n1")]}
    expose = [(s, "Calculator") for s in names]

    def spec_now():
        return w.RegistrySpec(sorted(reg.values()), expose, list(names), "10.0.0.5", 3)

    ctx.publish(spec_now().image(epoch=1))
    reqs = w.calculator_calls(40_000, seed=77)
    keys = np.array([p.service_key(s) for s in names], dtype=np.uint64)
    reqs["svc_key"] = keys[np.arange(len(reqs)) % len(names)]
    events = [("bbb", "n1", -1), ("ccc", "n1", -1),      # both empty now, same base
              ("ccc", "n9", +1), ("bbb", "n9", +1),      # back, the later one first
              ("bbb", "n9", -1), ("ccc", "n9", -1), ("bbb", "n5", +1), ("ccc", "n5", +1),
              ("aaa", "n1", -1), ("aaa", "n2", -1), ("bbb", "n5", -1), ("aaa", "n7", +1), ("bbb", "n7", +1)]
    for svc, node, delta in events:
        if delta < 0:
            del reg[(svc, node)]
        else:
            reg[(svc, node)] = (svc, node, "10.2.2.2", 5000)
        ctx.update_service(svc, delta)
        kv, snap = oracle.snapshot_from_spec(spec_now())
        assert_same(ctx.dispatch(reqs), snap.dispatch(reqs))
        # the context's host copy of the table went through the same patch
        rebuilt = spec_now().image(epoch=1)
        patched = ctx.current_image()
        h0, h1 = rebuilt.header, patched.header
        slots0 = rebuilt.blob[h0.off_slots:h0.off_slots + 16 * h0.n_slots]
        slots1 = patched.blob[h1.off_slots:h1.off_slots + 16 * h1.n_slots]
        assert slots0 == slots1


def test_names_whose_keys_fold_alike(ctx, oracle):
    """Four service names whose FNV-1a-64 keys xor-fold to the same 32 bits share all three
    candidate slots under every table seed; the builder has to move on to another key salt, and
    requests hashed with that salt route as the oracle says."""
    names = ["adv13608795", "adv31967732", "adv51281479", "adv54581514"]
    folds = {(w.service_key(s) & 0xFFFFFFFF) ^ (w.service_key(s) >> 32) for s in names}
    assert len(folds) == 1
    spec = w.RegistrySpec([(s, f"node_{k}", f"10.0.{i}.{k}", 7000 + k) for i, s in enumerate(names) for k in range(5)],
                          [(s, "Calculator") for s in names], [], "10.0.0.5", 3)
    image = spec.image()
    assert image.salt != 0
    ctx.publish(image)
    kv, snap = oracle.snapshot_from_spec(spec, image.salt)
    reqs = w.calculator_calls(50_000, seed=5)
    keys = np.array([p.service_key(s, image.salt) for s in names] + [p.service_key("adv00000000", image.salt)],
                    dtype=np.uint64)
    reqs["svc_key"] = keys[np.arange(len(reqs)) % len(keys)]
    want = snap.dispatch(reqs)
    assert_same(ctx.dispatch(reqs), want)
    assert (want["status"] == p.ST_NO_SERVICE).sum() == len(reqs) // len(keys)


def test_dispatch_right_behind_a_producer_on_the_stream(ctx, oracle):
    """assign_seq (asynchronous) rewrites the seq field of the very records the next dispatch
    reads, on the same stream and with nothing in between.  The dispatch must not overlap it -
    repeated, because an overlap would only sometimes be caught."""
    spec = w.stress_registry(300, 7)
    ctx.publish(spec.image())
    kv, snap = oracle.snapshot_from_spec(spec)
    n = 1 << 20
    reqs = w.stress_calls(n, 300, seed=123)
    d_req = ctx.device_buffer(n * 48)
    d_rep = ctx.device_buffer(n * 16)
    for rep in range(6):
        scrambled = reqs.copy()
        scrambled["seq"] = 0xDEAD0000 + rep
        d_req.upload(scrambled)
        ctx.dispatch_batches(ctx.make_batches([(d_req, d_rep, n)]), sync=False)   # a dispatch in flight first
        ctx.assign_seq(d_req, n, sync=False)
        ctx.dispatch_device(d_req, n, d_rep, sync=False, flags=p.F_AFTER_PRODUCER if rep % 2 else 0)
        ctx.sync()
        got_req = d_req.download(p.REQ_DTYPE, n)
        assert_same(d_rep.download(p.REPLY_DTYPE, n), snap.dispatch(got_req, threads=8))
        assert got_req["seq"].max() < 0xDEAD0000   # the counters' values, not the scrambled ones


@pytest.mark.parametrize("n_services", [1023, 1025, 1449, 1451, 1535, 1537, 3199, 3201, 6143, 6145, 10444, 10445])
def test_table_sizes_either_side_of_every_threshold(ctx, oracle, n_services):
    """Registries one step either side of each size at which something changes: 32 KiB of table
    (the shared-memory opt-in), 45.3 KiB (four 256-thread CTAs -> one 1024-thread CTA), 48 KiB (the
    default shared-memory limit), 100 KiB (the old shape switch), 192 KiB (half-full tables end,
    then the staging budget ends and the table is probed in L2).  Dispatch in one launch and as a
    batch list, the compact client form, and the balancer's counters drawn on the device."""
    spec = w.stress_registry(n_services, 3)
    image = spec.image()
    ctx.publish(image)
    kv, snap = oracle.snapshot_from_spec(spec)
    n = 70_001
    reqs = w.stress_calls(n, n_services, seed=n_services)
    want = snap.dispatch(reqs, threads=8)
    assert_same(ctx.dispatch(reqs), want)
    d_req = ctx.device_buffer(n * 48).upload(reqs)
    d_rep = ctx.device_buffer(n * 16)
    ctx.dispatch_batches(ctx.make_batches([(d_req, d_rep, n), (d_req, d_rep, 0)]))
    assert_same(d_rep.download(p.REPLY_DTYPE, n), want)
    one = w.calculator_calls(30_011, seed=3)
    one["svc_key"] = p.service_key(f"svc_{n_services - 1:05d}")
    key, base, methods, sel, args = w.as_client_batch(one)
    assert_same(ctx.call_batch(key, base, methods, args, sel), snap.dispatch(one, threads=8))
    d_req.free()
    d_rep.free()
    # the balancer's counters drawn on the device (a context of its own: counters start at zero),
    # then the dispatch that uses them
    assert image.salt == 0
    blank = reqs.copy()
    blank["seq"] = 0
    want_seq = blank.copy()
    snap.assign_seq(want_seq, np.zeros(snap.n_clients(), dtype=np.uint64))
    with p.Context([0]) as fresh:
        fresh.publish(image)
        d_req = fresh.device_buffer(n * 48).upload(blank)
        d_rep = fresh.device_buffer(n * 16)
        fresh.assign_seq(d_req, n)
        drawn = d_req.download(p.REQ_DTYPE, n)
        assert np.array_equal(drawn["seq"], want_seq["seq"])
        fresh.dispatch_device(d_req, n, d_rep)
        assert_same(d_rep.download(p.REPLY_DTYPE, n), snap.dispatch(want_seq, threads=8))
        d_req.free()
        d_rep.free()
