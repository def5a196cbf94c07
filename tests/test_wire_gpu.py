"""net/rpc's byte streams on the device: the client-to-server bytes of many
connections are decoded into request records, dispatched, and the replies encoded into the
server-to-client bytes.  Bar: byte-equal with the streams the CPU restatement of the reference's
loopback path puts on its TCP connections (oracle/loopback.c: po_wire_streams uses the encoders
po_loopback_run sends with)."""
import numpy as np
import pytest

import ptype_b200 as p
import workloads as w

pytestmark = pytest.mark.gpu


def build_streams(oracle, per_stream, receivers, check_limit=1 << 16):
    """Concatenate the wire streams of one fresh connection per request array."""
    c2s, s2c, c_off, s_off, call_off = [], [], [0], [0], [0]
    for reqs in per_stream:
        a, _, b, _ = oracle.wire_streams(reqs, receivers=receivers, check_limit=check_limit)
        c2s.append(a)
        s2c.append(b)
        c_off.append(c_off[-1] + len(a))
        s_off.append(s_off[-1] + len(b))
        call_off.append(call_off[-1] + len(reqs))
    cat = lambda xs: np.concatenate(xs) if xs else np.zeros(0, dtype=np.uint8)
    return cat(c2s), np.array(c_off, dtype=np.uint64), cat(s2c), np.array(s_off, dtype=np.uint64), \
        np.array(call_off, dtype=np.uint64)


def roundtrip(ctx, oracle, spec, service, per_stream, receivers):
    ctx.set_check_limit(1 << 16)
    ctx.publish(spec.image())
    kv, snap = oracle.snapshot_from_spec(spec)
    c2s, c_off, s2c, s_off, call_off = build_streams(oracle, per_stream, receivers)
    streams = np.zeros(len(per_stream), dtype=ctx.WIRE_STREAM_DTYPE)
    streams["svc_key"] = p.service_key(service)
    streams["seq0"] = 1 + np.arange(len(per_stream), dtype=np.uint64) * 1000
    streams["seq_step"] = 3
    reqs, calls, got_call_off, _ = ctx.wire_decode(c2s, c_off, streams)
    assert np.array_equal(got_call_off, call_off)
    want = np.concatenate(per_stream) if per_stream else np.zeros(0, dtype=p.REQ_DTYPE)
    assert len(reqs) == len(want)
    known = np.isin(want["method"], [w.MUL[0], w.ADD[0], w.CHECK[0]])
    # names the wire carried: the three known ones as they are, anything else as the stand-ins the
    # oracle's client sends ("Unknown.Method" / "nodot")
    unk, nodot = w.method_key("Unknown.Method"), w.method_key("nodot")
    want_m = np.where(known, want["method"], np.where(want["recv"] == p.RECV_ILL_FORMED, nodot[0], unk[0]))
    recv_of = {w.MUL[0]: w.MUL[1], w.ADD[0]: w.ADD[1], w.CHECK[0]: w.CHECK[1]}   # the name decides, not the record
    known_r = np.array([recv_of.get(int(m), 0) for m in want["method"]], dtype=np.uint64)
    want_r = np.where(known, known_r, np.where(want["recv"] == p.RECV_ILL_FORMED, nodot[1], unk[1]))
    assert np.array_equal(reqs["method"], want_m.astype(np.uint32)) and np.array_equal(reqs["recv"], want_r.astype(np.uint32))
    three = want["method"] == w.CHECK[0]
    assert np.array_equal(reqs["a0"], want["a0"]) and np.array_equal(reqs["a1"], want["a1"])
    assert np.array_equal(reqs["a2"], np.where(three, want["a2"], 0))
    assert not calls["flags"].any()
    for s in range(len(per_stream)):
        lo, hi = int(call_off[s]), int(call_off[s + 1])
        assert np.array_equal(calls["rpc_seq"][lo:hi], np.arange(hi - lo, dtype=np.uint64))
        assert np.array_equal(reqs["seq"][lo:hi], streams["seq0"][s] + 3 * np.arange(hi - lo, dtype=np.uint64))
    # the server's side of the call on the GPU, then its answer on the wire
    replies = ctx.dispatch(reqs)
    assert replies.tobytes() == snap.dispatch(reqs, check_limit=1 << 16).tobytes()
    out, out_off, state, _ = ctx.wire_encode(c2s, call_off, calls, replies, np.zeros(len(per_stream), dtype=np.uint8))
    assert np.array_equal(out_off, s_off)
    assert out.tobytes() == s2c.tobytes()
    return replies, state


def test_calculator_streams_byte_equal(ctx, oracle):
    spec = w.calculator_registry()
    rng = np.random.default_rng(3)
    per_stream = []
    for s in range(300):
        n = int(rng.integers(0, 200)) if s else 1
        per_stream.append(w.calculator_calls(n, seed=100 + s))
    per_stream.append(np.concatenate([w.calculator_calls(40, seed=7), w.edge_case_calls(["calculator"], n_random=600)]))
    per_stream.append(w.calculator_calls(5000, seed=8))
    replies, state = roundtrip(ctx, oracle, spec, "calculator", per_stream, receivers=1)
    assert (replies["status"] != 0).any() and state[-2] == 3 and state[-1] == 1 and state[1] in (0, 1)


def test_prime_streams_and_errors_byte_equal(ctx, oracle):
    spec = w.optimus_registry()
    per_stream = [w.optimus_calls(3000, t_lo=50, t_hi=5000)[:3000],
                  w.edge_case_calls(["prime_worker"], n_random=800),
                  np.concatenate([w.optimus_calls(10, t_lo=50, t_hi=500)[:10], w.calculator_calls(10)])]
    roundtrip(ctx, oracle, spec, "prime_worker", per_stream, receivers=2)


def test_streams_continue_across_batches(ctx, oracle):
    """A connection's bytes arrive in pieces: the second piece starts in mid-conversation (no type
    definitions, net/rpc's Seq continuing) and the encoder is told what the connection has seen."""
    spec = w.calculator_registry()
    ctx.set_check_limit(1 << 16)
    ctx.publish(spec.image())
    reqs = np.concatenate([w.calculator_calls(300, seed=5), w.edge_case_calls(["calculator"], n_random=100)[:200],
                           w.calculator_calls(100, seed=6)])
    c2s, coff, s2c, soff = oracle.wire_streams(reqs, receivers=1)
    cut = 250
    pieces = [(0, cut), (cut, len(reqs))]
    state = np.zeros(1, dtype=np.uint8)
    streams = np.zeros(1, dtype=ctx.WIRE_STREAM_DTYPE)
    streams["svc_key"], streams["seq_step"] = p.service_key("calculator"), 1
    got = b""
    for lo, hi in pieces:
        part = c2s[int(coff[lo]):int(coff[hi])]
        streams["seq0"] = lo + 1
        r, calls, call_off, _ = ctx.wire_decode(part, np.array([0, len(part)], dtype=np.uint64), streams)
        assert len(r) == hi - lo and calls["rpc_seq"][0] == lo
        replies = ctx.dispatch(r)
        out, _, state, _ = ctx.wire_encode(part, call_off, calls, replies, state)
        got += out.tobytes()
    assert got == s2c.tobytes()


def test_malformed_and_empty_input(ctx):
    streams = np.zeros(2, dtype=ctx.WIRE_STREAM_DTYPE)
    junk = np.frombuffer(bytes([5, 0xff, 0x82, 1, 200, 1]) + bytes([0x7f, 1, 2]), dtype=np.uint8)
    r, calls, call_off, _ = ctx.wire_decode(junk, np.array([0, 6, 9], dtype=np.uint64), streams)
    assert len(r) == 0 and call_off.tolist() == [0, 0, 0]
    r, calls, call_off, _ = ctx.wire_decode(np.zeros(0, dtype=np.uint8), np.array([0], dtype=np.uint64), streams[:0])
    assert len(r) == 0
