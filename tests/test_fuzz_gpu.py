"""Randomised registries x randomised call streams: the CUDA path against the oracle, byte for
byte, over awkward service/node names, every MaxConnections regime, sequence numbers of every
magnitude, known/unknown/ill-formed methods and adversarial Check ranges."""
import numpy as np
import pytest

import ptype_b200 as p
import workloads as w

pytestmark = pytest.mark.gpu
ALPHABET = list("abcXYZ019-._~+ !#$%&'()*,:;<=>?@[]^`{|}")
METHODS = ["Calculator.Multiply", "Calculator.Add", "Prime.Check", "Calculator.Sub", "Prime.Sieve",
           "Other.Method", "nodot", "", "Calculator.", ".Add", "Prime.Check.Check"]


def random_spec(rng):
    names = set()
    while len(names) < int(rng.integers(1, 40)):
        s = "".join(rng.choice(ALPHABET, int(rng.integers(1, 10))))
        if s not in (".", ".."):
            names.add(s)
    names = sorted(names)
    nodes, expose, declared = {}, [], []
    for svc in names:
        for _ in range(int(rng.integers(0, 9))):
            nd = "".join(rng.choice(ALPHABET, int(rng.integers(1, 6))))
            if nd not in (".", ".."):
                nodes[(svc, nd)] = (svc, nd, f"h{int(rng.integers(0, 99))}", int(rng.integers(1, 65536)))
        if not any(k[0] == svc for k in nodes):
            declared.append(svc)
        for recv in ("Calculator", "Prime"):
            if rng.integers(0, 3) > 0:
                expose.append((svc, recv))
    local_addr = str(rng.choice(["", "10.0.0.5", "127.0.0.1", "fe80::1"]))
    return w.RegistrySpec(list(nodes.values()), expose, declared, local_addr,
                          int(rng.choice([0, 1, 2, 3, 4, 7, 300]))), names


def random_calls(rng, names, n, salt):
    r = np.zeros(n, dtype=p.REQ_DTYPE)
    keys = np.array([p.service_key(s, salt) for s in names + ["no such service", "x"]], dtype=np.uint64)
    r["svc_key"] = keys[rng.integers(0, len(keys), n)]
    shift = rng.integers(0, 64, n).astype(np.uint64)
    r["seq"] = rng.integers(0, 1 << 63, n, dtype=np.uint64) * np.uint64(2) + rng.integers(0, 2, n).astype(np.uint64)
    r["seq"] >>= shift
    mk = np.array([p.method_key(m) for m in METHODS], dtype=np.uint32)
    which = rng.integers(0, len(METHODS), n)
    which[rng.integers(0, 4, n) > 0] %= 3                       # mostly the three real methods
    r["method"], r["recv"] = mk[which, 0], mk[which, 1]
    def ints():
        v = rng.integers(-(1 << 63), (1 << 63) - 1, n, dtype=np.int64)
        return v >> rng.integers(0, 64, n)
    r["a0"], r["a1"], r["a2"] = ints(), ints(), ints()
    chk = which == 2                                            # Check: windows of every width near a0
    width = rng.integers(-3, 70000, n)
    width[rng.integers(0, 3, n) > 0] %= 40
    with np.errstate(over="ignore"):
        r["a1"] = np.where(chk, r["a0"] + width, r["a1"])
    small = chk & (rng.integers(0, 2, n) == 0)                  # the optimus regime: small positive values
    r["a0"] = np.where(small, rng.integers(1, 5000, n), r["a0"])
    r["a1"] = np.where(small, r["a0"] + rng.integers(0, 30, n), r["a1"])
    r["a2"] = np.where(small, rng.integers(1, 1 << 23, n), r["a2"])
    return r


@pytest.mark.parametrize("seed", range(8))
def test_random_registry_and_calls(ctx, oracle, seed):
    rng = np.random.default_rng(9000 + seed)
    spec, names = random_spec(rng)
    img = spec.image()
    ctx.publish(img)
    kv, snap = oracle.snapshot_from_spec(spec, img.salt)
    n = 60_000
    reqs = random_calls(rng, names, n, img.salt)
    for limit in (1 << 16, 50):
        ctx.set_check_limit(limit)
        want = snap.dispatch(reqs, check_limit=limit, threads=4)
        got = ctx.dispatch(reqs)
        if got.tobytes() != want.tobytes():
            j = int(np.flatnonzero((got["value"] != want["value"]) | (got["node"] != want["node"]) |
                                   (got["status"] != want["status"]))[0])
            raise AssertionError(f"seed {seed} limit {limit}: first difference at {j}: got {got[j]} want {want[j]} "
                                 f"req {reqs[j]} max_conn {spec.max_connections}")
        d_req = ctx.device_buffer(n * 48).upload(reqs)
        d_rep = ctx.device_buffer(n * 16)
        for flags in (p.F_BULK_LOADS, p.F_GLOBAL_TABLE, p.F_GLOBAL_TABLE | p.F_LDG_LOADS):
            ctx.dispatch_device(d_req, n, d_rep, flags=flags)
            assert d_rep.download(p.REPLY_DTYPE, n).tobytes() == want.tobytes(), (seed, limit, flags)
        ctx.dispatch_batches(ctx.make_batches([(d_req, d_rep, n)]))
        assert d_rep.download(p.REPLY_DTYPE, n).tobytes() == want.tobytes()
        d_req.free()
        d_rep.free()
    ctx.set_check_limit(1 << 16)
    assert set(np.unique(want["status"]).tolist()) >= {0, 1}
