"""Pins the CPU oracle (oracle/ptype_oracle.c) to the results the reference's own tests
fix for the dispatch path, then to hand-derived known answers.  No GPU."""
import ctypes as C
import json
from pathlib import Path

import numpy as np
import pytest

GOLD = Path(__file__).parent / "golden"
REF = json.loads((GOLD / "reference_vectors.json").read_text())
DER = json.loads((GOLD / "derived_vectors.json").read_text())


def select(ob, addr, n, mx):
    out = (C.c_uint32 * 256)()
    k = ob.lib().po_select_nodes(addr.encode(), n, mx, out, 256)
    return list(out[:k])


@pytest.mark.parametrize("case", REF["select_nodes"], ids=lambda c: c["cite"])
def test_select_nodes_reference(oracle, case):
    assert select(oracle, case["local_addr"], case["n"], case["max"]) == case["want"]


@pytest.mark.parametrize("case", REF["round_robin"], ids=lambda c: c["cite"])
def test_round_robin_reference(oracle, case):
    assert oracle.lib().po_round_robin(case["seq"], case["n"]) == case["want"]


def test_round_robin_distinct_run(oracle):
    d = REF["distinct_run"]
    picks = [oracle.lib().po_round_robin(d["first_seq"] + i, d["n"]) for i in range(d["n"])]
    assert len(set(picks)) == d["n"] and min(picks) >= 0


def test_services_grouping_and_order(oracle):
    kv = oracle.KV()
    for k, v in REF["services"]["puts"]:
        kv.put(k, v)
    got = kv.services()
    want = {s: [tuple(n) for n in nodes] for s, nodes in REF["services"]["want"].items()}
    assert got == want


def test_nodes_prefix_case(oracle):
    kv = oracle.KV()
    for k, v in REF["nodes_prefix"]["puts"]:
        kv.put(k, v)
    assert kv.nodes(REF["nodes_prefix"]["service"]) == [tuple(n) for n in REF["nodes_prefix"]["want"]]


def test_register_json_shape(oracle):
    kv = oracle.KV()
    for svc, node, host, port in REF["register_json"]["registrations"]:
        kv.register(svc, node, host, port)
    for prefix, want in REF["register_json"]["want"].items():
        got = [json.loads(v) for _, v in kv.range_prefix(prefix)]
        assert got == want


def test_watch_sequence(oracle):
    kv = oracle.KV()
    for step in REF["watch_sequence"]["steps"]:
        if "put" in step:
            kv.put(*step["put"])
        else:
            assert kv.delete(step["delete"]) == 1
        assert kv.nodes("foo") == [tuple(n) for n in step["want"]]


def test_multiply_example(oracle):
    m = REF["multiply_example"]
    assert oracle.lib().po_multiply(m["a"], m["b"]) == m["want"]


# ---- derived known answers -------------------------------------------------

def test_fnv(oracle):
    for s, h in DER["fnv1a32"].items():
        assert oracle.lib().po_fnv1a32(s.encode(), len(s)) == int(h, 16)
    for s, h in DER["fnv1a64"].items():
        assert oracle.lib().po_fnv1a64_salted(s.encode(), len(s), 0) == int(h, 16)


@pytest.mark.parametrize("case", DER["select_nodes"], ids=lambda c: f"{c['local_addr']}/{c['n']}")
def test_select_nodes_derived(oracle, case):
    assert select(oracle, case["local_addr"], case["n"], case["max"]) == case["want"]


def test_round_robin_n3(oracle):
    assert [oracle.lib().po_round_robin(s, 3) for s in range(1, 7)] == DER["round_robin_n3"]


def test_wrapping_multiply(oracle):
    for a, b, want in DER["multiply"]:
        assert oracle.lib().po_multiply(int(a), int(b)) == int(want)


def test_check_chunks(oracle):
    import workloads as w
    for t, want in DER["check_chunks"].items():
        chunks = w.split_work(int(t))
        buf = (C.c_int64 * (3 * 64))()
        n = oracle.lib().po_split_work(int(t), buf, 64)
        assert n == len(want) == len(chunks)
        assert np.array_equal(np.array(buf[: 3 * n]).reshape(n, 3), chunks)
        assert [oracle.lib().po_check(*map(int, c)) for c in chunks] == want
    for lo, hi, t, want in DER["check"]:
        assert oracle.lib().po_check(lo, hi, t) == want


def test_go_integer_corners(oracle):
    L = oracle.lib()
    i64min, i64max = -(1 << 63), (1 << 63) - 1
    assert L.po_add(i64max, 1) == i64min
    assert L.po_check(-1, 5, 4) == -1          # 4 % -1 == 0
    assert L.po_check(-1, 5, i64min) == i64min  # bound = MinInt64: empty loop
    assert L.po_check(i64min, i64min + 4, i64min) == i64min
    assert L.po_round_robin(1 << 63, 3) == -2   # int(index) negative, remainder != 0: Go panics
    assert L.po_round_robin(1 << 63, 2) == 0
    assert L.po_select_nodes(b"", 4, -1, None, 0) == -1


def test_gob_values(oracle):
    buf = C.create_string_buffer(16)
    for v, hx in DER["gob_int"].items():
        n = oracle.lib().po_gob_int(int(v), buf)
        assert buf.raw[:n].hex() == hx
    for v, hx in DER["gob_uint"].items():
        n = oracle.lib().po_gob_uint(int(v), buf)
        assert buf.raw[:n].hex() == hx


def test_json_unmarshal_rules(oracle):
    nd = oracle.PoNode()
    L = oracle.lib()
    assert L.po_node_unmarshal(b'{"address":"a","port":1}', C.byref(nd)) == 0
    # a value missing a field keeps the previous one (the reused Node of registry.go:99,102)
    assert L.po_node_unmarshal(b'{"port":2}', C.byref(nd)) == 0
    assert (nd.address, nd.port) == (b"a", 2)
    assert L.po_node_unmarshal(b'{"Address":"b","PORT":3,"extra":[1,{"x":null}]}', C.byref(nd)) == 0
    assert (nd.address, nd.port) == (b"b", 3)
    assert L.po_node_unmarshal(b'{"port":3.5}', C.byref(nd)) != 0
    assert L.po_node_unmarshal(b'{"address":5}', C.byref(nd)) != 0
    assert L.po_node_unmarshal(b'{"address":"x\\u00e9\\n"}', C.byref(nd)) == 0
    assert nd.address == "xé\n".encode()


def test_client_call_statuses(oracle):
    import workloads as w
    spec = w.calculator_registry()
    spec.declare = ["ghost"]
    kv, snap = oracle.snapshot_from_spec(spec)
    assert snap.selected("calculator") == [3, 0, 9]
    assert snap.call("calculator", 1, "Calculator.Multiply", [7, 8]) == (56, 0, 0)
    assert snap.call("calculator", 3, "Calculator.Add", [7, 8]) == (15, 3, 0)
    assert snap.call("calculator", 2, "Calculator.Divide", [7, 8]) == (0, 9, 5)
    assert snap.call("calculator", 2, "Prime.Check", [2, 10, 91]) == (0, 9, 4)
    assert snap.call("calculator", 2, "nodot", [1, 2]) == (0, 9, 3)
    assert snap.call("ghost", 1, "Calculator.Multiply", [7, 8]) == (0, 0xFFFFFFFF, 2)
    assert snap.call("nobody", 1, "Calculator.Multiply", [7, 8]) == (0, 0xFFFFFFFF, 1)
    assert snap.call("calculator", (1 << 63) + 1, "Calculator.Multiply", [7, 8])[2] == 6


def test_golden_fixtures_match_oracle(oracle):
    """The committed fixture replies are what the oracle says today."""
    import workloads as w
    fx = np.load(GOLD / "dispatch_fixtures.npz")
    for name, spec in (("calc3", w.calculator_registry()), ("calcmesh", w.calculator_registry(mesh=True)),
                       ("optimus", w.optimus_registry())):
        kv, snap = oracle.snapshot_from_spec(spec)
        rep = snap.dispatch(fx[f"{name}_req"].view(oracle.REQ_DTYPE).reshape(-1), check_limit=1 << 16)
        assert rep.tobytes() == fx[f"{name}_rep"].tobytes(), name


# ---- BASELINE config 1: the reference's real path, restated over loopback TCP ---------

def test_loopback_wire_bytes(oracle):
    """gob messages of the calculator example: reply 56 is 03 04 00 70,
    Args{7,8} body is 01 0e 01 10 00, and rpc.Request's type definition names its two fields."""
    buf = C.create_string_buffer(128)
    n = oracle.lib().po_gob_reply_int_message(56, buf)
    assert buf.raw[:n].hex() == "03040070"
    a = (C.c_int64 * 3)(7, 8, 0)
    n = oracle.lib().po_gob_args_body(a, 2, buf)
    assert buf.raw[:n].hex() == "010e011000"
    a = (C.c_int64 * 3)(0, 8, 0)          # zero-valued fields are omitted: delta 2 reaches B
    n = oracle.lib().po_gob_args_body(a, 2, buf)
    assert buf.raw[:n].hex() == "021000"
    n = oracle.lib().po_gob_request_typedef(buf)
    raw = buf.raw[:n]
    assert raw[1:3] == b"\xff\x81" and b"Request" in raw and b"ServiceMethod" in raw and b"Seq" in raw


def test_loopback_calls_match_in_process_oracle(oracle):
    """Config 1 in miniature: sequential synchronous calls through TCP 127.0.0.1, the
    CONNECT /_goRPC_ handshake and gob framing give the replies of the in-process oracle, and
    the connection used is roundRobinSelect's (first call -> index 1 of 3, rpc_test.go:118-158)."""
    import workloads as w
    spec = w.calculator_registry()
    kv, snap = oracle.snapshot_from_spec(spec)
    reqs = w.calculator_calls(3000)
    rep, st = oracle.loopback_run(reqs, k=3, receivers=1)
    want = snap.dispatch(reqs)
    assert np.array_equal(rep["value"], want["value"]) and np.array_equal(rep["status"], want["status"])
    sel = np.array(snap.selected("calculator"))
    assert np.array_equal(sel[rep["node"]], want["node"]) and rep["node"][0] == 1
    assert st.calls == 3000 and 30 < st.bytes_client_to_server / st.calls < 120
    # a node that registered only Prime answers "can't find service" to Calculator calls
    rep, _ = oracle.loopback_run(reqs[:50], k=1, receivers=2)
    assert np.all(rep["status"] == 4) and np.all(rep["value"] == 0)
    # optimus chunks over the wire
    chk = w.optimus_calls(500, t_lo=50, t_hi=400)
    rep, _ = oracle.loopback_run(chk, k=2, receivers=2)
    kv2, snap2 = oracle.snapshot_from_spec(w.optimus_registry())
    want = snap2.dispatch(chk)
    assert np.array_equal(rep["value"], want["value"]) and np.array_equal(rep["node"], want["node"])


def test_assign_seq_is_n_sequential_calls(oracle):
    """po_assign_seq against the rule written out in numpy: a counter per Client, advanced only by
    calls that find a non-empty client list (rpc.go:177-182), values taken in batch order; and
    against the reference's own fixed point - a fresh Client's first call draws 1, so with two
    nodes it lands on index 1 (rpc_test.go:118-158)."""
    import workloads as w
    from ptype_b200 import service_key
    spec = w.RegistrySpec([("a", "n1", "h", 1), ("a", "n2", "h", 2), ("b", "n1", "h", 3)], [("a", "Calculator")],
                          ["empty"], "", 3)
    kv, snap = oracle.snapshot_from_spec(spec)
    names = ["a", "b", "empty", "ghost"]
    rng = np.random.default_rng(7)
    pick = rng.integers(0, 4, 5000)
    reqs = np.zeros(len(pick), dtype=oracle.REQ_DTYPE)
    reqs["svc_key"] = np.array([service_key(s) for s in names], dtype=np.uint64)[pick]
    reqs["seq"] = 0xDEAD
    counters = np.zeros(snap.n_clients(), dtype=np.uint64)
    counters[snap.client_index("b")] = 41
    snap.assign_seq(reqs[:3000], counters)
    snap.assign_seq(reqs[3000:], counters)      # counters persist across batches
    want = np.zeros(len(pick), dtype=np.uint64)
    for i, start in ((0, 0), (1, 41)):
        at = np.flatnonzero(pick == i)
        want[at] = start + 1 + np.arange(len(at), dtype=np.uint64)
    assert np.array_equal(reqs["seq"], want)
    assert counters[snap.client_index("a")] == (pick == 0).sum()
    assert counters[snap.client_index("b")] == 41 + (pick == 1).sum()
    assert counters[snap.client_index("empty")] == 0 and snap.client_index("ghost") == -1
    first = np.zeros(1, dtype=oracle.REQ_DTYPE)
    first["svc_key"] = service_key("a")
    fresh = np.zeros(snap.n_clients(), dtype=np.uint64)
    snap.assign_seq(first, fresh)
    assert first["seq"][0] == 1 and oracle.lib().po_round_robin(int(first["seq"][0]), 2) == 1
