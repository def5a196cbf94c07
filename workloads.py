"""Synthetic registries and call streams for the BASELINE.json configs.

Pure data generation (numpy and pure-Python FNV); nothing here computes a reply and nothing
here loads libptype_b200.so or the oracle, so both arms of bench.py, the tests and the tools
build their inputs from the same code without depending on the thing they measure.  PRNG:
splitmix64, seed 0x5eed0000 + config number, little-endian records.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Tuple

import numpy as np

# Record layout of include/ptype_b200.h (ptype_req, 48 bytes).
REQ_DTYPE = np.dtype(
    [("svc_key", "<u8"), ("seq", "<u8"), ("method", "<u4"), ("recv", "<u4"),
     ("a0", "<i8"), ("a1", "<i8"), ("a2", "<i8")], align=False)
RECV_ILL_FORMED = 0xFFFFFFFF


def fnv1a32(data: bytes) -> int:
    """hash/fnv New32a [Go stdlib]; the reference's one explicit hash (cluster/rpc.go:267)."""
    h = 0x811C9DC5
    for b in data:
        h = ((h ^ b) * 0x01000193) & 0xFFFFFFFF
    return h


def service_key(name: str, salt: int = 0) -> int:
    """FNV-1a-64 of a cluster service name, offset basis xor-ed with the image's key salt
    (include/ptype_b200.h: ptype_b200_service_key; equality is pinned by tests/test_image_host.py)."""
    h = 0xCBF29CE484222325 ^ salt
    for b in name.encode():
        h = ((h ^ b) * 0x100000001B3) & 0xFFFFFFFFFFFFFFFF
    return h


def method_key(service_method: str) -> Tuple[int, int]:
    """(method, recv) hashes of "Receiver.Method", split at the last '.' like net/rpc's server."""
    raw = service_method.encode()
    dot = raw.rfind(b".")
    return fnv1a32(raw), (fnv1a32(raw[:dot]) if dot >= 0 else RECV_ILL_FORMED)

_GAMMA = np.uint64(0x9E3779B97F4A7C15)


def splitmix64(seed: int, n: int, stream: int = 0) -> np.ndarray:
    """n outputs of splitmix64 started at `seed` (+ a per-stream offset), vectorised."""
    with np.errstate(over="ignore"):
        base = np.uint64((seed + stream * 0x632BE59BD9B4E019) & 0xFFFFFFFFFFFFFFFF)
        z = base + (np.arange(1, n + 1, dtype=np.uint64) * _GAMMA)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


@dataclass
class RegistrySpec:
    """What was registered, and the Client-side selection inputs."""
    nodes: List[Tuple[str, str, str, int]]           # (service, node_name, address, port)
    expose: List[Tuple[str, str]] = field(default_factory=list)   # (service, receiver)
    declare: List[str] = field(default_factory=list)              # services with an empty list
    local_addr: str = ""
    max_connections: int = 3

    def image(self, epoch: int = 1, n_slots: int = 0):
        """The registry image of this spec (needs the product library: imported here only)."""
        from ptype_b200.snapshot import build_image
        return build_image(self.nodes, local_addr=self.local_addr,
                           max_connections=self.max_connections, expose=self.expose,
                           declare=self.declare, epoch=epoch, n_slots=n_slots)


MUL = method_key("Calculator.Multiply")
ADD = method_key("Calculator.Add")
CHECK = method_key("Prime.Check")


def calculator_registry(n_nodes: int = 16, *, mesh: bool = False,
                        local_addr: str = "10.0.0.5") -> RegistrySpec:
    """C2: 1 service `calculator` x 16 nodes calculator_node_01..16 @ 10.0.0.1..16:31234
    (node naming after example/calculator/server/calculator_server.yaml:1-3)."""
    nodes = [("calculator", f"calculator_node_{i:02d}", f"10.0.0.{i}", 31234)
             for i in range(1, n_nodes + 1)]
    return RegistrySpec(nodes, [("calculator", "Calculator")], [], local_addr, 0 if mesh else 3)


def calculator_calls(n: int, *, seed: int = 0x5EED0002, seq_base: int = 0, salt: int = 0) -> np.ndarray:
    """C2 stream: 50/50 Add/Multiply; half the args full-range int64 (wraparound), half in
    [-2^20, 2^20]; seq = seq_base+1 .. seq_base+n (one Client, N sequential Calls)."""
    r = np.zeros(n, dtype=REQ_DTYPE)
    bits = splitmix64(seed, n, 0)
    a = splitmix64(seed, n, 1).view(np.int64)
    b = splitmix64(seed, n, 2).view(np.int64)
    small = (bits >> np.uint64(1)) & np.uint64(1) == 1
    span = np.int64((1 << 21) + 1)
    a = np.where(small, (a.view(np.uint64) % np.uint64(span)).astype(np.int64) - np.int64(1 << 20), a)
    b = np.where(small, (b.view(np.uint64) % np.uint64(span)).astype(np.int64) - np.int64(1 << 20), b)
    is_mul = (bits & np.uint64(1)) == 1
    r["svc_key"] = service_key("calculator", salt)
    r["seq"] = np.arange(seq_base + 1, seq_base + n + 1, dtype=np.uint64)
    r["method"] = np.where(is_mul, MUL[0], ADD[0]).astype(np.uint32)
    r["recv"] = MUL[1]
    r["a0"], r["a1"] = a, b
    return r


def stress_registry(n_services: int = 10_000, n_nodes: int = 64, *, mesh: bool = False,
                    local_addr: str = "10.0.0.5") -> RegistrySpec:
    """C3: svc_00000..svc_09999 x node_00..63."""
    nodes = []
    for s in range(n_services):
        for k in range(n_nodes):
            nodes.append((f"svc_{s:05d}", f"node_{k:02d}", f"10.{(s >> 8) & 255}.{s & 255}.{k}", 9000 + k))
    expose = [(f"svc_{s:05d}", "Calculator") for s in range(n_services)]
    return RegistrySpec(nodes, expose, [], local_addr, 0 if mesh else 3)


def stress_calls(n: int, n_services: int = 10_000, *, seed: int = 0x5EED0003, miss_rate: float = 0.01,
                 zipf: bool = False, salt: int = 0) -> np.ndarray:
    """C3 stream: service uniform over the registered ones (or Zipf s=1), `miss_rate` of the
    calls name an unregistered service, seq = running per-service count, Multiply."""
    r = np.zeros(n, dtype=REQ_DTYPE)
    u = splitmix64(seed, n, 0)
    if zipf:
        w = 1.0 / np.arange(1, n_services + 1, dtype=np.float64)
        cdf = np.cumsum(w / w.sum())
        svc = np.searchsorted(cdf, (u >> np.uint64(11)).astype(np.float64) / float(1 << 53)).astype(np.int64)
        svc = np.minimum(svc, n_services - 1)
    else:
        svc = (u % np.uint64(n_services)).astype(np.int64)
    miss = (splitmix64(seed, n, 1) % np.uint64(1_000_000)) < np.uint64(int(miss_rate * 1_000_000))
    keys = np.array([service_key(f"svc_{s:05d}", salt) for s in range(n_services)], dtype=np.uint64)
    ghost = np.array([service_key(f"ghost_{s:05d}", salt) for s in range(1024)], dtype=np.uint64)
    r["svc_key"] = np.where(miss, ghost[svc % 1024], keys[svc])
    # running per-service count: rank of each call among the calls of its service, from 1
    order = np.argsort(svc, kind="stable")
    sorted_svc = svc[order]
    starts = np.flatnonzero(np.r_[True, sorted_svc[1:] != sorted_svc[:-1]])
    run = np.arange(n, dtype=np.int64) - np.repeat(starts, np.diff(np.r_[starts, n]))
    seq = np.empty(n, dtype=np.uint64)
    seq[order] = (run + 1).astype(np.uint64)
    r["seq"] = seq
    r["method"], r["recv"] = MUL
    r["a0"] = splitmix64(seed, n, 2).view(np.int64)
    r["a1"] = splitmix64(seed, n, 3).view(np.int64)
    return r


def optimus_registry() -> RegistrySpec:
    """C4: prime_worker x 2 nodes (example/optimus/worker/prime_worker.yaml:1-3,
    prime_worker2.yaml:1-3); 2 <= MaxConnections 3 -> identity selection."""
    nodes = [("prime_worker", "prime_worker_1", "127.0.0.1", 8000),
             ("prime_worker", "prime_worker_2", "127.0.0.1", 8002)]
    return RegistrySpec(nodes, [("prime_worker", "Prime")], [], "127.0.0.1", 3)


def split_work(target: int) -> np.ndarray:
    """splitWork (example/optimus/coordinator/coordinator.go:67-73) as (min,max,target) rows."""
    if target <= 0:
        return np.zeros((0, 3), dtype=np.int64)
    hi = np.arange(10, target + 10, 10, dtype=np.int64)
    lo = np.r_[np.int64(2), hi[:-1]]
    return np.stack([lo, hi, np.full_like(hi, target)], axis=1)


def optimus_calls(min_calls: int, *, seed: int = 0x5EED0004, t_lo: int = 1_000, t_hi: int = 1_000_000,
                  salt: int = 0) -> np.ndarray:
    """C4 stream: targets uniform in [t_lo, t_hi] until >= min_calls calls are emitted by the
    splitWork rule; the 250 ms sleep of prime.go:17 is dropped."""
    est = max(8, int(min_calls / ((t_lo + t_hi) / 20.0) * 1.5) + 8)
    while True:
        t = (splitmix64(seed, est, 0) % np.uint64(t_hi - t_lo + 1)).astype(np.int64) + t_lo
        per = (t + 9) // 10
        csum = np.cumsum(per)
        if csum[-1] >= min_calls:
            break
        est *= 2
    cut = int(np.searchsorted(csum, min_calls)) + 1
    t, per = t[:cut], per[:cut]
    n = int(per.sum())
    tgt = np.repeat(t, per)
    first = np.repeat(np.cumsum(per) - per, per)
    j = np.arange(n, dtype=np.int64) - first          # chunk number within its target
    hi = (j + 1) * 10
    lo = np.where(j == 0, 2, j * 10)
    r = np.zeros(n, dtype=REQ_DTYPE)
    r["svc_key"] = service_key("prime_worker", salt)
    r["seq"] = np.arange(1, n + 1, dtype=np.uint64)
    r["method"], r["recv"] = CHECK
    r["a0"], r["a1"], r["a2"] = lo, hi, tgt
    return r


def edge_case_calls(spec_services: List[str], *, salt: int = 0, seed: int = 0x5EED00EE,
                    n_random: int = 4096) -> np.ndarray:
    """A mixed batch that reaches every status and arithmetic corner: unknown service and
    method, ill-formed names, wrong receiver, seq at and beyond 2^63, int64 wraparound,
    zero / negative / MinInt64 Check arguments, ranges past the iteration limit."""
    I64_MIN, I64_MAX = -(1 << 63), (1 << 63) - 1
    rows = []

    def add(svc, seq, sm, a0=0, a1=0, a2=0):
        if isinstance(sm, tuple):
            m, rv = sm
        else:
            m, rv = method_key(sm)
        rows.append((service_key(svc, salt), seq & 0xFFFFFFFFFFFFFFFF, m, rv, a0, a1, a2))

    ints = [0, 1, -1, 2, -2, 7, 8, 10, 97, 100, 121, 1 << 31, (1 << 31) - 1, -(1 << 31), 1 << 32,
            (1 << 32) + 1, 3037000500, 1 << 62, I64_MAX, I64_MIN, I64_MIN + 1, 0xFFFFFFFF, -0xFFFFFFFF]
    seqs = [0, 1, 2, 3, 15, 16, 17, (1 << 32) - 1, 1 << 32, (1 << 32) + 5, (1 << 63) - 1, 1 << 63,
            (1 << 63) + 1, (1 << 63) + 6, (1 << 64) - 1, (1 << 64) - 64]
    for svc in spec_services + ["nobody_home"]:
        for i, s in enumerate(seqs):
            add(svc, s, "Calculator.Multiply", ints[i % len(ints)], ints[(i * 7 + 3) % len(ints)])
            add(svc, s, "Calculator.Add", ints[(i * 5) % len(ints)], ints[(i * 3 + 1) % len(ints)])
            add(svc, s, "Prime.Check", 2, 10, 91)
        for a in ints:
            for b in ints:
                add(svc, 1 + len(rows), "Calculator.Multiply", a, b)
                add(svc, 1 + len(rows), "Calculator.Add", a, b)
        for sm in ["Calculator.Divide", "Prime.Multiply", "Nope.Multiply", "nodot", "", ".", "Calculator.",
                   ".Multiply", "Calculator.Multiply.Multiply", "calculator.multiply"]:
            add(svc, 1 + len(rows), sm, 3, 4, 5)
        add(svc, 5, (MUL[0], RECV_ILL_FORMED), 3, 4)
        checks = [(2, 10, 91), (10, 20, 91), (2, 10, 97), (0, 3, 7), (-5, 10, 20), (-1, 5, 4),
                  (-1, 5, I64_MIN), (-10, 10, 0), (-10, 10, -12), (5, 5, 25), (6, 5, 25), (2, 100, 97),
                  (2, 1 << 40, 1 << 41), (I64_MIN, I64_MAX, I64_MAX), (I64_MAX - 5, I64_MAX, I64_MAX),
                  ((1 << 32) - 3, (1 << 32) + 4, ((1 << 32) + 1) * 3), (2, 50, (1 << 33) + 2),
                  (-3, 3, (1 << 40)), (I64_MIN, I64_MIN + 4, I64_MIN), (1, 2, 0), (0, 0, 0),
                  (3, 70000, 65537 * 3), (2, 65540, 4295098369)]
        for c in checks:
            add(svc, 1 + len(rows), "Prime.Check", *c)
    fixed = np.array(rows, dtype=[("svc_key", "<u8"), ("seq", "<u8"), ("method", "<u4"), ("recv", "<u4"),
                                  ("a0", "<i8"), ("a1", "<i8"), ("a2", "<i8")]).astype(REQ_DTYPE)
    # random tail: random service among the listed ones, random method, small Check ranges
    n = n_random
    rnd = np.zeros(n, dtype=REQ_DTYPE)
    u = splitmix64(seed, n, 0)
    names = spec_services + ["nobody_home"]
    keys = np.array([service_key(s, salt) for s in names], dtype=np.uint64)
    rnd["svc_key"] = keys[(u % np.uint64(len(names))).astype(np.int64)]
    rnd["seq"] = splitmix64(seed, n, 1) >> (splitmix64(seed, n, 2) % np.uint64(64))
    which = (splitmix64(seed, n, 3) % np.uint64(4)).astype(np.int64)
    meths = np.array([MUL[0], ADD[0], CHECK[0], 0xDEADBEEF], dtype=np.uint32)
    recvs = np.array([MUL[1], ADD[1], CHECK[1], MUL[1]], dtype=np.uint32)
    rnd["method"], rnd["recv"] = meths[which], recvs[which]
    a0 = splitmix64(seed, n, 4).view(np.int64) >> (splitmix64(seed, n, 5) % np.uint64(64)).astype(np.int64)
    rnd["a0"] = a0
    width = (splitmix64(seed, n, 6) % np.uint64(40)).astype(np.int64)
    is_check = which == 2
    rnd["a1"] = np.where(is_check, a0 + width, splitmix64(seed, n, 7).view(np.int64))
    tgt = splitmix64(seed, n, 8).view(np.int64) >> (splitmix64(seed, n, 9) % np.uint64(64)).astype(np.int64)
    rnd["a2"] = tgt
    return np.concatenate([fixed, rnd])


def as_client_batch(reqs: np.ndarray):
    """Compact form of a stream issued by ONE Client (same service, consecutive seq): what
    (*Client).CallBatch submits.  Returns (svc_key, seq_base, methods, method_sel, args) for
    Context.call_batch; method_sel is None when every call uses one method."""
    assert len(reqs) > 0
    key = int(reqs["svc_key"][0])
    assert np.all(reqs["svc_key"] == reqs["svc_key"][0]), "one Client names one service"
    seq0 = int(reqs["seq"][0])
    assert np.array_equal(reqs["seq"], np.arange(seq0, seq0 + len(reqs), dtype=np.uint64))
    pairs = np.stack([reqs["method"], reqs["recv"]], axis=1)
    uniq, inv = np.unique(pairs, axis=0, return_inverse=True)
    assert len(uniq) <= 4, "at most PTYPE_BATCH_METHODS methods per batch"
    methods = [(int(m), int(r)) for m, r in uniq]
    sel = None if len(uniq) == 1 else inv.reshape(-1).astype(np.uint8)
    three = bool(np.any(reqs["a2"] != 0))
    cols = [reqs["a0"], reqs["a1"]] + ([reqs["a2"]] if three else [])
    args = np.ascontiguousarray(np.stack(cols, axis=1), dtype=np.int64)
    return key, (seq0 - 1) & 0xFFFFFFFFFFFFFFFF, methods, sel, args
