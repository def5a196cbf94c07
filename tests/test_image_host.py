"""Host-side logic: the registry image builder against the oracle's snapshot, the C ABI's
exported symbols against include/ptype_b200.h, workload generators.  No GPU."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

import ptype_b200 as p
from ptype_b200 import _abi
import workloads as w
from ptype_b200.sharding import shard_of, shard_indices

ROOT = Path(__file__).resolve().parent.parent


def test_header_symbols_exported():
    """Every function include/ptype_b200.h declares is exported by the library and bound."""
    text = (ROOT / "include" / "ptype_b200.h").read_text()
    declared = set(re.findall(r"\b(ptype_b200_\w+)\s*\(", text))
    bound = {name for name, _, _ in _abi.SIGNATURES}
    assert declared == bound, (declared - bound, bound - declared)
    lib = C.CDLL(str(p.LIB_PATH))
    for name in declared:
        assert hasattr(lib, name), name
    assert _abi.lib.ptype_b200_abi_version() == 1


def test_record_layouts():
    assert p.REQ_DTYPE.itemsize == 48 and p.REPLY_DTYPE.itemsize == 16
    assert [p.REQ_DTYPE.fields[f][1] for f in ("svc_key", "seq", "method", "recv", "a0", "a1", "a2")] == \
        [0, 8, 16, 20, 24, 32, 40]
    assert [p.REPLY_DTYPE.fields[f][1] for f in ("value", "node", "status")] == [0, 8, 12]


def test_workload_generators_hash_like_the_library():
    """workloads.py hashes names in pure Python (so the CPU arm of bench.py needs no product code);
    the keys it writes into request records must be the library's."""
    assert w.REQ_DTYPE == p.REQ_DTYPE and w.RECV_ILL_FORMED == p.RECV_ILL_FORMED
    for name in ["", "calculator", "prime_worker", "svc_00042", "ghost_00001", "a/b", "é"]:
        for salt in (0, 1, 0x1234_5678_9ABC_DEF0):
            assert w.service_key(name, salt) == p.service_key(name, salt)
    for sm in ["Calculator.Multiply", "Calculator.Add", "Prime.Check", "nodot", "", ".", "a.b.c", ".x", "x."]:
        assert w.method_key(sm) == p.method_key(sm), sm


def test_hash_helpers_match_oracle(oracle):
    for s in ["", "0", "calculator", "prime_worker", "svc_00042", "Calculator.Multiply", "x" * 300]:
        b = s.encode()
        assert _abi.lib.ptype_b200_fnv1a32(b, len(b)) == oracle.lib().po_fnv1a32(b, len(b))
        for salt in (0, 1, 0xDEADBEEF):
            assert p.service_key(s, salt) == oracle.lib().po_fnv1a64_salted(b, len(b), salt)
    for sm in ["Calculator.Multiply", "Prime.Check", "nodot", "", ".", "a.b.c", "Calculator."]:
        m, r = C.c_uint32(), C.c_uint32()
        oracle.lib().po_method_key(sm.encode(), C.byref(m), C.byref(r))
        assert p.method_key(sm) == (m.value, r.value)
    assert p.method_key("nodot")[1] == p.RECV_ILL_FORMED
    for addr in ["", "10.0.0.5", "127.0.0.1"]:
        for i in (0, 1, 2, 10, 123):
            for n in (1, 4, 16, 64):
                assert p.conn_hash(addr, i) % n == oracle.lib().po_hash_connection_index(addr.encode(), i, n)


SPECS = {
    "calc3": lambda: w.calculator_registry(),
    "calcmesh": lambda: w.calculator_registry(mesh=True),
    "optimus": lambda: w.optimus_registry(),
    "stress_small": lambda: w.stress_registry(200, 7),
    "stress_mesh_small": lambda: w.stress_registry(50, 9, mesh=True),
}


@pytest.mark.parametrize("name", list(SPECS))
def test_image_selection_matches_oracle(oracle, name):
    """getSelectedNodes() per service, and the global node numbering, agree with the oracle."""
    spec = SPECS[name]()
    spec.declare = ["declared_but_empty"]
    img = spec.image(epoch=7)
    kv, snap = oracle.snapshot_from_spec(spec)
    assert img.epoch == 7 and img.header.n_nodes == len(spec.nodes)
    services = sorted({s for s, *_ in spec.nodes})
    for svc in services[:64] + services[-8:]:
        assert img.selected(svc) == snap.selected(svc), svc
    assert img.selected("declared_but_empty") == [] == snap.selected("declared_but_empty")
    assert img.selected("never_heard_of_it") is None and snap.selected("never_heard_of_it") is None
    for nid in list(range(min(40, len(spec.nodes)))) + [len(spec.nodes) - 1]:
        a, port, svc, _ = img.node(nid)
        assert (a, port, svc) == snap.node(nid)


def test_node_order_is_etcd_key_order(oracle):
    """node10 sorts before node2, and 'foo-x' before 'foo' (because '-' < '/')."""
    nodes = [("foo", "node2", "h2", 2), ("foo", "node10", "h10", 10), ("foo", "node1", "h1", 1),
             ("foo-x", "a", "hx", 5), ("foo_y", "a", "hy", 6), ("bar", "z", "hb", 7)]
    spec = w.RegistrySpec(nodes, [], [], "", 0)
    img = spec.image()
    kv, snap = oracle.snapshot_from_spec(spec)
    got = [img.node(i)[:3] for i in range(len(nodes))]
    assert got == [snap.node(i) for i in range(len(nodes))]
    assert [g[0] for g in got] == ["hb", "hx", "h1", "h10", "h2", "hy"]
    assert img.selected("foo") == [2, 3, 4]


def test_reregister_overwrites_and_validation():
    b = p.ImageBuilder("", 3)
    b.register("s", "n", "old", 1).register("s", "n", "new", 2)
    img = b.finish(3)
    assert img.node(0)[:2] == ("new", 2) and img.header.n_nodes == 1
    for bad in ("", "a/b", ".", ".."):
        with pytest.raises(p.PtypeError):
            p.ImageBuilder().register(bad, "n", "h", 1)
    with pytest.raises(p.PtypeError):
        p.ImageBuilder("", -1)          # Go would panic in make([]Node, 0, -1) (rpc.go:257)
    with pytest.raises(p.PtypeError):
        p.ImageBuilder().expose("s", "NoSuchReceiver")
    blob = bytearray(img.blob)
    blob[200] ^= 1
    with pytest.raises(p.PtypeError):
        p.Image(bytes(blob))
    with pytest.raises(p.PtypeError):
        p.Image(img.blob[:-16])


def test_table_sizing():
    small = w.calculator_registry().image()
    assert small.header.n_slots == 8
    spec = w.stress_registry(300, 1)
    assert spec.image().header.n_slots == 600
    assert spec.image(n_slots=340).header.n_slots == 340
    with pytest.raises(p.PtypeError):
        spec.image(n_slots=320)         # above the ~90 % load 3-way cuckoo hashing can place


def test_stress_image_fits_shared_memory_budget():
    """C3's 10 000 services get a 12 288-slot (192 KiB) table: staged in shared memory, and small
    enough to leave the SM 60 KiB of L1 beside it.  10 444 services is the most that size takes."""
    spec = w.stress_registry(10_000, 1)
    img = spec.image()
    assert img.header.n_slots == 12_288 and img.header.n_services == 10_000
    assert w.stress_registry(10_444, 1).image().header.n_slots == 12_288
    assert w.stress_registry(10_445, 1).image().header.n_slots == 2 * 10_448    # probed in L2
    assert w.stress_registry(6_144, 1).image().header.n_slots == 12_288


def slot_candidates(key: np.ndarray, seed: int, n_slots: int) -> np.ndarray:
    """numpy restatement of ptype::slot_of (csrc/table_hash.h) for all three ways."""
    with np.errstate(over="ignore"):
        fold = ((key & np.uint64(0xFFFFFFFF)) ^ (key >> np.uint64(32))).astype(np.uint32) ^ np.uint32(seed)
        out = []
        for mul in (0x9E3779B1, 0x85EBCA77, 0xC2B2AE3D):
            h = fold * np.uint32(mul)
            out.append(((h.astype(np.uint64) * np.uint64(n_slots)) >> np.uint64(32)).astype(np.int64))
    return np.stack(out, axis=1)


@pytest.mark.parametrize("n_services,n_slots", [(10_000, 0), (500, 560), (3, 0), (1, 0)])
def test_every_key_sits_in_one_of_its_three_slots(n_services, n_slots):
    """3-way cuckoo placement: a device lookup is at most three probes, hit or miss."""
    img = w.stress_registry(n_services, 1).image(n_slots=n_slots)
    h = img.header
    slots = np.frombuffer(img.blob, dtype=[("key", "<u8"), ("base", "<u4"), ("nr", "<u4")],
                          count=h.n_slots, offset=h.off_slots)
    occ = np.flatnonzero(slots["base"] != 0xFFFFFFFF)
    assert len(occ) == n_services
    cand = slot_candidates(slots["key"][occ], h.table_seed, h.n_slots)
    assert np.all((cand == occ[:, None]).any(axis=1))
    keys = np.array([p.service_key(f"svc_{s:05d}", h.key_salt) for s in range(n_services)], dtype=np.uint64)
    assert set(keys.tolist()) == set(slots["key"][occ].tolist())


def test_method_table_is_a_perfect_hash():
    img = w.calculator_registry().image()
    h = img.header
    mt = np.frombuffer(img.blob, dtype=[("m", "<u4"), ("r", "<u4"), ("op", "<u4"), ("bit", "<u4")],
                       count=64, offset=h.off_mtable)
    assert h.method_mul % 2 == 1 and (mt["op"] != 0).sum() == h.n_methods == 3
    for name, op in (("Calculator.Multiply", 1), ("Calculator.Add", 2), ("Prime.Check", 3)):
        m, r = p.method_key(name)
        e = mt[((m * h.method_mul) & 0xFFFFFFFF) >> 26]
        assert (e["m"], e["r"], e["op"]) == (m, r, op)


def test_unregister_and_device_only_images(oracle):
    """A node's key goes away (lease expiry / delete, registry_test.go:179-191): the rebuilt image
    selects from the nodes left, like the oracle after the same delete.  Device-only images carry
    just what the kernels read."""
    spec = w.calculator_registry()
    b = p.ImageBuilder(spec.local_addr, spec.max_connections)
    for svc, node, addr, port in spec.nodes:
        b.register(svc, node, addr, port)
    b.expose("calculator", "Calculator")
    kv, snap = oracle.snapshot_from_spec(spec)
    assert b.finish(1).selected("calculator") == snap.selected("calculator")
    b.unregister("calculator", "calculator_node_04")
    assert kv.delete("services/calculator/calculator_node_04/") == 1
    snap2 = oracle.Snapshot(kv, spec.local_addr, spec.max_connections, [], 0, spec.expose)
    img2 = b.finish(2)
    assert img2.header.n_nodes == 15 and img2.selected("calculator") == snap2.selected("calculator")
    with pytest.raises(p.PtypeError):
        b.unregister("calculator", "calculator_node_04")
    small = b.set_device_only(True).finish(3)
    assert small.header.host_sections == 0 and small.header.n_nodes == 15
    assert len(small) < len(img2) and small.selected("calculator") is None
    assert small.blob[192:small.header.device_bytes] == img2.blob[192:img2.header.device_bytes]


def test_split_work_generator(oracle):
    for t in (-5, 0, 1, 9, 10, 11, 91, 100, 1000, 12345):
        buf = (C.c_int64 * (3 * 2000))()
        n = oracle.lib().po_split_work(t, buf, 2000)
        got = w.split_work(t)
        assert n == len(got)
        assert np.array_equal(np.array(buf[: 3 * n], dtype=np.int64).reshape(n, 3), got)


def test_optimus_stream_is_split_work():
    r = w.optimus_calls(5000, t_lo=50, t_hi=500)
    assert len(r) >= 5000
    # re-derive the stream target by target
    i = 0
    while i < len(r):
        t = int(r["a2"][i])
        chunks = w.split_work(t)
        seg = r[i: i + len(chunks)]
        assert np.array_equal(np.stack([seg["a0"], seg["a1"], seg["a2"]], axis=1), chunks)
        i += len(chunks)
    assert np.array_equal(r["seq"], np.arange(1, len(r) + 1, dtype=np.uint64))


def test_stress_stream_per_service_seq():
    r = w.stress_calls(20_000, 50, miss_rate=0.05)
    keys = {p.service_key(f"svc_{s:05d}"): s for s in range(50)}
    assert r["seq"].min() == 1
    hit = np.isin(r["svc_key"], np.array(list(keys), dtype=np.uint64))
    assert 0.93 < hit.mean() < 0.97


def test_sharding_partitions_stream():
    n, g = 100_000, 8
    owner = shard_of(np.arange(n, dtype=np.uint64), g)
    assert owner.min() == 0 and owner.max() == g - 1
    counts = np.bincount(owner, minlength=g)
    assert counts.sum() == n and counts.min() > n / g * 0.95
    parts = [shard_indices(n, g, r) for r in range(g)]
    allidx = np.sort(np.concatenate(parts))
    assert np.array_equal(allidx, np.arange(n))


def test_sharded_dispatch_equals_unsharded(oracle):
    """Sharding changes no reply: seq travels in the record (rpc.go:181)."""
    spec = w.calculator_registry()
    kv, snap = oracle.snapshot_from_spec(spec)
    r = w.calculator_calls(10_000)
    whole = snap.dispatch(r)
    out = np.empty_like(whole)
    for rank in range(4):
        idx = shard_indices(len(r), 4, rank)
        out[idx] = snap.dispatch(r[idx])
    assert out.tobytes() == whole.tobytes()


def test_random_registries_match_oracle(oracle):
    """Randomised registries with awkward names (characters that sort around '/', long and short
    names, re-registration, empty services): the image's per-service selection and global node
    numbering equal the oracle's for several (localAddr, MaxConnections) pairs."""
    rng = np.random.default_rng(2024)
    alphabet = list("abcXYZ019-._~+ !#$%&'()*,:;<=>?@[]^`{|}")
    for trial in range(25):
        n_services = int(rng.integers(1, 30))
        names = set()
        while len(names) < n_services:
            k = int(rng.integers(1, 12))
            s = "".join(rng.choice(alphabet, k))
            if s not in (".", ".."):
                names.add(s)
        nodes = {}
        for svc in names:
            for _ in range(int(rng.integers(0, 12))):
                nd = "".join(rng.choice(alphabet, int(rng.integers(1, 8))))
                if nd in (".", ".."):
                    continue
                nodes[(svc, nd)] = (svc, nd, f"h{int(rng.integers(0, 1000))}", int(rng.integers(1, 65536)))
        declared = [s for s in names if not any(k[0] == s for k in nodes)]
        local_addr = str(rng.choice(["", "10.0.0.5", "127.0.0.1", "192.168.1.77"]))
        max_conn = int(rng.choice([0, 1, 2, 3, 5, 300]))
        spec = w.RegistrySpec(list(nodes.values()), [], declared, local_addr, max_conn)
        img = spec.image()
        kv, snap = oracle.snapshot_from_spec(spec)
        assert img.header.n_nodes == len(nodes) == len(kv.range_prefix("services/"))
        for svc in names:
            assert img.selected(svc) == snap.selected(svc), (trial, svc)
        for nid in range(len(nodes)):
            a, port, svc, _ = img.node(nid)
            assert (a, port, svc) == snap.node(nid), (trial, nid)


@pytest.mark.parametrize("device_only", [False, True])
def test_builder_kept_alive_emits_the_same_bytes_as_a_fresh_one(device_only):
    """A builder that lives across live updates keeps its key order and table placement while the
    set of service names stands still (only nodes join and leave); every image it emits must be
    byte-identical to one built from scratch from the same registrations."""
    rng = np.random.default_rng(0xB111D)
    state = {}       # (service, node) -> (address, port)
    exposed = set()  # (service, receiver)
    declared = set()
    live = p.ImageBuilder("10.0.0.5", 3).set_device_only(device_only)

    def fresh(epoch):
        b = p.ImageBuilder("10.0.0.5", 3).set_device_only(device_only)
        for (svc, node), (addr, port) in state.items():
            b.register(svc, node, addr, port)
        for svc in declared:
            b.declare(svc)
        for svc, recv in exposed:
            b.expose(svc, recv)
        return b.finish(epoch)

    for step in range(120):
        kind = rng.integers(0, 10)
        svc = f"svc_{int(rng.integers(0, 40 if step > 30 else 12)):03d}"
        node = f"node_{int(rng.integers(0, 6)):02d}"
        if kind < 5:
            addr, port = f"10.1.{step}.{int(rng.integers(1, 250))}", int(rng.integers(1000, 60000))
            live.register(svc, node, addr, port)
            state[(svc, node)] = (addr, port)
            declared.add(svc)
        elif kind < 8:
            if (svc, node) in state:
                live.unregister(svc, node)
                del state[(svc, node)]
            elif svc not in declared:
                with pytest.raises(p.PtypeError):
                    live.unregister(svc, node)
        elif kind == 8:
            live.declare(svc)
            declared.add(svc)
        else:
            recv = "Calculator" if rng.integers(0, 2) else "Prime"
            live.expose(svc, recv)
            exposed.add((svc, recv))
            declared.add(svc)
        if step % 3 == 0:
            assert live.finish(step + 1).blob == fresh(step + 1).blob, step
    assert live.finish(999).blob == fresh(999).blob
