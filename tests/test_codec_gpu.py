"""Device-side formats either side of the dispatch path, against the
oracle: name resolution like net/rpc's server, gob reply messages, gob Args bodies."""
import ctypes as C

import numpy as np
import pytest

import ptype_b200 as p
import workloads as w

pytestmark = pytest.mark.gpu


def test_resolve_names_like_net_rpc(ctx, oracle):
    names = ["Calculator.Multiply", "Calculator.Add", "Prime.Check", "nodot", "", ".", "a.b.c", "Calculator.",
             ".Multiply", "x" * 32, "y" * 31 + ".", "RPCTest.Call", "RPCRetryTest.Go", "a.b" + "c" * 29]
    names += [f"Svc{i}.Method{i * 7}" for i in range(5000)]
    got = ctx.resolve_names(names, stride=32)
    for j, s in enumerate(names):
        m, r = C.c_uint32(), C.c_uint32()
        oracle.lib().po_method_key(s.encode(), C.byref(m), C.byref(r))
        assert (int(got["method"][j]), int(got["recv"][j])) == (m.value, r.value), s
    assert got["recv"][3] == p.RECV_ILL_FORMED
    wide = ctx.resolve_names(["Some.Very.Long.Receiver.Name.With.Many.Dots.Method"], stride=64)
    assert (int(wide["method"][0]), int(wide["recv"][0])) == p.method_key(
        "Some.Very.Long.Receiver.Name.With.Many.Dots.Method")


def test_resolved_names_drive_dispatch(ctx, oracle):
    """Names hashed on the device and patched into the records give the oracle's replies."""
    spec = w.calculator_registry()
    ctx.publish(spec.image())
    kv, snap = oracle.snapshot_from_spec(spec)
    n = 100_000
    reqs = w.calculator_calls(n)
    names = np.where(reqs["method"] == w.MUL[0], "Calculator.Multiply", "Calculator.Add").tolist()
    blank = reqs.copy()
    blank["method"] = 0
    blank["recv"] = 0
    host = np.zeros((n, 32), dtype=np.uint8)
    for s in ("Calculator.Multiply", "Calculator.Add"):
        b = np.frombuffer(s.encode(), dtype=np.uint8)
        host[np.array(names) == s, : len(b)] = b
    d_names = ctx.device_buffer(host.nbytes).upload(host)
    d_req = ctx.device_buffer(n * 48).upload(blank)
    d_rep = ctx.device_buffer(n * 16)
    from ptype_b200._abi import lib, check
    check(lib.ptype_b200_resolve_names(ctx._h, 0, d_names.ptr, 32, n, None, d_req.ptr), "resolve_names")
    ctx.dispatch_device(d_req, n, d_rep)
    assert d_rep.download(p.REPLY_DTYPE, n).tobytes() == snap.dispatch(reqs).tobytes()
    for b in (d_names, d_req, d_rep):
        b.free()


def test_gob_reply_messages(ctx, oracle):
    """`LL 04 00 <gob int>` per OK reply, nothing for failed ones; bytes equal the oracle's
    encoder (known answers: 56 -> 03 04 00 70)."""
    vals = [56, 0, 1, -1, 63, 64, -64, -65, 127, 128, 255, 256, 65535, 65536, (1 << 31), -(1 << 31),
            (1 << 62), (1 << 63) - 1, -(1 << 63), -(1 << 63) + 1, 3037000500 * 3, -9223372036709301616]
    rng = np.random.default_rng(3)
    n = 200_003
    rep = np.zeros(n, dtype=p.REPLY_DTYPE)
    rep["value"] = rng.integers(-(1 << 63), (1 << 63) - 1, n, dtype=np.int64) >> rng.integers(0, 64, n)
    rep["value"][: len(vals)] = vals
    rep["status"] = (rng.integers(0, 50, n) == 0).astype(np.uint32) * 5
    rep["status"][: len(vals)] = 0
    d_rep = ctx.device_buffer(n * 16).upload(rep)
    off, data = ctx.gob_encode_replies(d_rep, n)
    d_rep.free()
    assert off[0] == 0 and off[-1] == len(data)
    assert data[:4].tobytes().hex() == "03040070"
    buf = C.create_string_buffer(16)
    check_idx = list(range(len(vals))) + rng.integers(0, n, 3000).tolist() + [n - 1]
    for j in check_idx:
        got = data[int(off[j]): int(off[j + 1])].tobytes()
        if rep["status"][j] != 0:
            assert got == b""
        else:
            k = oracle.lib().po_gob_reply_int_message(int(rep["value"][j]), buf)
            assert got == buf.raw[:k], (j, rep["value"][j])
    lens = np.diff(off.astype(np.int64))
    assert np.all(lens[rep["status"] != 0] == 0) and np.all((lens[rep["status"] == 0] >= 4) & (lens[rep["status"] == 0] <= 12))


def test_gob_args_bodies(ctx, oracle):
    """calculator.Args / prime.Args bodies as the reference's client would send them decode to
    the same int64 fields (zero-valued fields omitted on the wire)."""
    rng = np.random.default_rng(4)
    for k in (2, 3):
        n = 50_000
        args = rng.integers(-(1 << 63), (1 << 63) - 1, (n, k), dtype=np.int64) >> rng.integers(0, 64, (n, k))
        args[rng.integers(0, n, n // 5), rng.integers(0, k, n // 5)] = 0
        args[:3] = [[7, 8, 0][:k], [0, 0, 0][:k], [0, 8, 9][:k]]
        chunks, off = [], [0]
        buf = C.create_string_buffer(64)
        for row in args:
            a = (C.c_int64 * 3)(*(row.tolist() + [0])[:3])
            m = oracle.lib().po_gob_args_body(a, k, buf)
            chunks.append(buf.raw[:m])
            off.append(off[-1] + m)
        data = np.frombuffer(b"".join(chunks), dtype=np.uint8)
        got, st = ctx.gob_decode_args(data, np.array(off, dtype=np.uint64), k)
        assert np.array_equal(got, args) and not st.any()
    # truncated and oversized bodies are flagged, not read past
    bad = np.frombuffer(bytes.fromhex("01f8ffff") + bytes.fromhex("01f7" + "ff" * 9 + "00"), dtype=np.uint8)
    got, st = ctx.gob_decode_args(bad, np.array([0, 4, len(bad)], dtype=np.uint64), 2)
    assert st.tolist() == [1, 1]
