"""ctypes binding of the CPU oracle (oracle/ptype_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product package never imports this.
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path
from typing import Iterable, List, Optional, Sequence, Tuple

import numpy as np

HERE = Path(__file__).resolve().parent
LIB = HERE / "_build" / "libptype_oracle.so"

REQ_DTYPE = np.dtype([("svc_key", "<u8"), ("seq", "<u8"), ("method", "<u4"), ("recv", "<u4"),
                      ("a0", "<i8"), ("a1", "<i8"), ("a2", "<i8")])
REPLY_DTYPE = np.dtype([("value", "<i8"), ("node", "<u4"), ("status", "<u4")])
U64_MAX = (1 << 64) - 1


def build(force: bool = False) -> Path:
    srcs = [HERE / "ptype_oracle.c", HERE / "ptype_oracle.h", HERE / "loopback.c", HERE / "Makefile"]
    newest = max(p.stat().st_mtime for p in srcs if p.exists())
    if force or not LIB.exists() or LIB.stat().st_mtime < newest:
        subprocess.run(["make", "-C", str(HERE), "all"], check=True, capture_output=True)
    return LIB


class PoNode(C.Structure):
    _fields_ = [("address", C.c_char * 256), ("port", C.c_int64)]


class PoServiceList(C.Structure):
    _fields_ = [("name", C.c_char * 256), ("nodes", C.POINTER(PoNode)), ("n", C.c_size_t)]


_lib: Optional[C.CDLL] = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(str(LIB))
        vp = C.c_void_p
        sigs = {
            "po_fnv1a32": (C.c_uint32, [C.c_char_p, C.c_size_t]),
            "po_fnv1a64_salted": (C.c_uint64, [C.c_char_p, C.c_size_t, C.c_uint64]),
            "po_itoa": (C.c_int, [C.c_int64, C.c_char_p]),
            "po_etcd_key": (C.c_int, [C.c_char_p, C.c_size_t, C.POINTER(C.c_char_p), C.c_int]),
            "po_kv_new": (vp, []), "po_kv_free": (None, [vp]),
            "po_kv_put": (None, [vp, C.c_char_p, C.c_char_p]),
            "po_kv_delete": (C.c_int, [vp, C.c_char_p]),
            "po_kv_range_prefix": (C.c_size_t, [vp, C.c_char_p, C.POINTER(C.c_size_t)]),
            "po_kv_key_at": (C.c_char_p, [vp, C.c_size_t]),
            "po_kv_value_at": (C.c_char_p, [vp, C.c_size_t]),
            "po_kv_len": (C.c_size_t, [vp]),
            "po_node_marshal": (C.c_int, [C.POINTER(PoNode), C.c_char_p, C.c_size_t]),
            "po_node_unmarshal": (C.c_int, [C.c_char_p, C.POINTER(PoNode)]),
            "po_register": (C.c_int, [vp, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int64]),
            "po_services": (C.c_long, [vp, C.POINTER(C.POINTER(PoServiceList))]),
            "po_services_free": (None, [C.POINTER(PoServiceList), C.c_size_t]),
            "po_nodes": (C.c_long, [vp, C.c_char_p, C.POINTER(C.POINTER(PoNode))]),
            "po_hash_connection_index": (C.c_int64, [C.c_char_p, C.c_int64, C.c_int64]),
            "po_select_nodes": (C.c_long, [C.c_char_p, C.c_size_t, C.c_int64, C.POINTER(C.c_uint32), C.c_size_t]),
            "po_round_robin": (C.c_int64, [C.c_uint64, C.c_size_t]),
            "po_multiply": (C.c_int64, [C.c_int64, C.c_int64]),
            "po_add": (C.c_int64, [C.c_int64, C.c_int64]),
            "po_check": (C.c_int64, [C.c_int64, C.c_int64, C.c_int64]),
            "po_check_width": (C.c_uint64, [C.c_int64, C.c_int64, C.c_int64]),
            "po_split_work": (C.c_size_t, [C.c_int64, vp, C.c_size_t]),
            "po_snapshot_build": (vp, [vp, C.c_char_p, C.c_int64, C.POINTER(C.c_char_p), C.c_size_t, C.c_uint64]),
            "po_snapshot_free": (None, [vp]),
            "po_snapshot_expose": (C.c_int, [vp, C.c_char_p, C.c_char_p]),
            "po_snapshot_n_nodes": (C.c_size_t, [vp]),
            "po_snapshot_selected": (C.c_long, [vp, C.c_char_p, C.POINTER(C.c_uint32), C.c_size_t]),
            "po_snapshot_node": (C.c_int, [vp, C.c_uint32, C.POINTER(PoNode), C.c_char_p, C.c_size_t]),
            "po_call": (None, [vp, C.c_char_p, C.c_uint64, C.c_char_p, C.POINTER(C.c_int64), C.c_uint64, vp]),
            "po_dispatch": (None, [vp, vp, C.c_size_t, vp, C.c_uint64]),
            "po_dispatch_mt": (None, [vp, vp, C.c_size_t, vp, C.c_uint64, C.c_int]),
            "po_assign_seq": (None, [vp, vp, C.c_size_t, vp]),
            "po_snapshot_n_clients": (C.c_size_t, [vp]),
            "po_snapshot_client_index": (C.c_long, [vp, C.c_char_p]),
            "po_reply_checksum": (C.c_uint64, [vp, C.c_size_t, C.c_uint64]),
            "po_method_key": (None, [C.c_char_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
            "po_gob_uint": (C.c_int, [C.c_uint64, C.c_char_p]),
            "po_gob_int": (C.c_int, [C.c_int64, C.c_char_p]),
            "po_loopback_run": (C.c_int, [vp, C.c_size_t, vp, C.c_int, C.c_uint32, C.c_uint64,
                                          C.POINTER(LoopbackStats)]),
            "po_wire_streams": (C.c_int, [vp, C.c_size_t, C.c_uint32, C.c_uint64, vp, C.c_size_t, vp, vp, C.c_size_t, vp]),
            "po_gob_reply_int_message": (C.c_int, [C.c_int64, C.c_char_p]),
            "po_gob_args_body": (C.c_int, [C.POINTER(C.c_int64), C.c_int, C.c_char_p]),
            "po_gob_request_typedef": (C.c_int, [C.c_char_p]),
        }
        for name, (res, args) in sigs.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


class LoopbackStats(C.Structure):
    _fields_ = [("seconds", C.c_double), ("calls", C.c_uint64), ("bytes_client_to_server", C.c_uint64),
                ("bytes_server_to_client", C.c_uint64)]


def loopback_run(reqs: np.ndarray, k: int = 1, receivers: int = 1, check_limit: int = 1 << 16):
    """BASELINE config 1: len(reqs) sequential synchronous Client.Call()s over loopback TCP with
    the net/rpc handshake and gob framing restated in C (oracle/loopback.c).  Returns
    (replies, stats); replies[j].node is the index of the connection the call went out on."""
    reqs = np.ascontiguousarray(reqs, dtype=REQ_DTYPE)
    rep = np.zeros(len(reqs), dtype=REPLY_DTYPE)
    st = LoopbackStats()
    rc = lib().po_loopback_run(reqs.ctypes.data, len(reqs), rep.ctypes.data, k, receivers, check_limit,
                               C.byref(st))
    if rc != 0:
        raise RuntimeError(f"po_loopback_run failed: {rc}")
    return rep, st


class KV:
    """The ordered KV standing in for etcd."""

    def __init__(self):
        self.h = lib().po_kv_new()

    def put(self, key: str, value: str) -> None:
        lib().po_kv_put(self.h, key.encode(), value.encode())

    def delete(self, key: str) -> int:
        return lib().po_kv_delete(self.h, key.encode())

    def register(self, service: str, node: str, host: str, port: int) -> None:
        if lib().po_register(self.h, service.encode(), node.encode(), host.encode(), port) != 0:
            raise ValueError("po_register failed")

    def range_prefix(self, prefix: str) -> List[Tuple[str, str]]:
        first = C.c_size_t()
        n = lib().po_kv_range_prefix(self.h, prefix.encode(), C.byref(first))
        return [(lib().po_kv_key_at(self.h, first.value + i).decode(),
                 lib().po_kv_value_at(self.h, first.value + i).decode()) for i in range(n)]

    def services(self):
        out = C.POINTER(PoServiceList)()
        n = lib().po_services(self.h, C.byref(out))
        if n < 0:
            raise ValueError("po_services: unmarshal error")
        res = {}
        for i in range(n):
            sl = out[i]
            res[sl.name.decode()] = [(sl.nodes[j].address.decode(), sl.nodes[j].port) for j in range(sl.n)]
        lib().po_services_free(out, n)
        return res

    def nodes(self, service: str):
        out = C.POINTER(PoNode)()
        n = lib().po_nodes(self.h, service.encode(), C.byref(out))
        if n < 0:
            raise ValueError("po_nodes: unmarshal error")
        res = [(out[j].address.decode(), out[j].port) for j in range(n)]
        C.CDLL(None).free(out)
        return res

    def close(self):
        if self.h:
            lib().po_kv_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Snapshot:
    def __init__(self, kv: KV, local_addr: str, max_connections: int, declared: Sequence[str] = (),
                 salt: int = 0, expose: Iterable[Tuple[str, str]] = ()):
        arr = (C.c_char_p * max(len(declared), 1))(*[d.encode() for d in declared])
        self.h = lib().po_snapshot_build(kv.h, local_addr.encode(), max_connections, arr, len(declared), salt)
        if not self.h:
            raise ValueError("po_snapshot_build failed")
        for svc, recv in expose:
            if lib().po_snapshot_expose(self.h, svc.encode(), recv.encode()) != 0:
                raise ValueError(f"expose({svc},{recv}) failed")

    def selected(self, service: str) -> Optional[List[int]]:
        out = (C.c_uint32 * 4096)()
        k = lib().po_snapshot_selected(self.h, service.encode(), out, 4096)
        if k < 0:
            return None
        if k > 4096:
            out = (C.c_uint32 * k)()
            lib().po_snapshot_selected(self.h, service.encode(), out, k)
        return list(out[:k])

    def node(self, node_id: int):
        nd, svc = PoNode(), C.create_string_buffer(256)
        if lib().po_snapshot_node(self.h, node_id, C.byref(nd), svc, 256) != 0:
            raise IndexError(node_id)
        return nd.address.decode(), nd.port, svc.value.decode()

    def call(self, service: str, seq: int, service_method: str, args: Sequence[int],
             check_limit: int = U64_MAX):
        a = (C.c_int64 * 3)(*(list(args) + [0, 0, 0])[:3])
        rep = np.zeros(1, dtype=REPLY_DTYPE)
        lib().po_call(self.h, service.encode(), seq, service_method.encode(), a, check_limit, rep.ctypes.data)
        return int(rep["value"][0]), int(rep["node"][0]), int(rep["status"][0])

    def dispatch(self, reqs: np.ndarray, check_limit: int = U64_MAX, threads: int = 1) -> np.ndarray:
        reqs = np.ascontiguousarray(reqs, dtype=REQ_DTYPE)
        rep = np.empty(len(reqs), dtype=REPLY_DTYPE)
        if threads > 1:
            lib().po_dispatch_mt(self.h, reqs.ctypes.data, len(reqs), rep.ctypes.data, check_limit, threads)
        else:
            lib().po_dispatch(self.h, reqs.ctypes.data, len(reqs), rep.ctypes.data, check_limit)
        return rep

    def n_clients(self) -> int:
        return lib().po_snapshot_n_clients(self.h)

    def client_index(self, service: str) -> int:
        return lib().po_snapshot_client_index(self.h, service.encode())

    def assign_seq(self, reqs: np.ndarray, counters: np.ndarray) -> None:
        """Fills reqs['seq'] in place and advances counters (uint64[n_clients()], key order)."""
        assert reqs.dtype == REQ_DTYPE and reqs.flags.c_contiguous
        assert counters.dtype == np.uint64 and len(counters) == self.n_clients()
        lib().po_assign_seq(self.h, reqs.ctypes.data, len(reqs), counters.ctypes.data)

    def dispatch_into(self, reqs: np.ndarray, rep: np.ndarray, check_limit: int, threads: int) -> None:
        lib().po_dispatch_mt(self.h, reqs.ctypes.data, len(reqs), rep.ctypes.data, check_limit, threads)

    def close(self):
        if self.h:
            lib().po_snapshot_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def wire_streams(reqs: np.ndarray, receivers: int = 1, check_limit: int = 1 << 16):
    """The net/rpc byte streams of ONE fresh connection carrying the calls `reqs` (after the CONNECT
    handshake), from the encoders the loopback run sends over TCP: returns
    (client_to_server bytes, offsets[n+1], server_to_client bytes, offsets[n+1])."""
    reqs = np.ascontiguousarray(reqs, dtype=REQ_DTYPE)
    n = len(reqs)
    cap = 256 + 160 * n
    c2s, s2c = np.zeros(cap, dtype=np.uint8), np.zeros(cap + 64 * n, dtype=np.uint8)
    co, so = np.zeros(n + 1, dtype=np.uint64), np.zeros(n + 1, dtype=np.uint64)
    rc = lib().po_wire_streams(reqs.ctypes.data, n, receivers, check_limit, c2s.ctypes.data, len(c2s),
                               co.ctypes.data, s2c.ctypes.data, len(s2c), so.ctypes.data)
    if rc != 0:
        raise RuntimeError("po_wire_streams: buffer too small")
    return c2s[: int(co[n])].copy(), co, s2c[: int(so[n])].copy(), so


def reply_checksum(rep: np.ndarray, index_base: int = 0) -> int:
    rep = np.ascontiguousarray(rep, dtype=REPLY_DTYPE)
    return lib().po_reply_checksum(rep.ctypes.data, len(rep), index_base)


def snapshot_from_spec(spec, salt: int = 0):
    """Feed a ptype_b200.workloads.RegistrySpec's registrations to the oracle."""
    kv = KV()
    for svc, node, addr, port in spec.nodes:
        kv.register(svc, node, addr, port)
    snap = Snapshot(kv, spec.local_addr, spec.max_connections, list(spec.declare), salt, spec.expose)
    return kv, snap
