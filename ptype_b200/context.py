"""Device contexts: install snapshots, dispatch batches (host or device buffers).

Mirrors the call path of cluster.(*Client).Call (cluster/rpc.go:59-67) for whole
batches.  Everything goes through the C ABI; nothing here computes a reply.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple

import numpy as np

from . import _abi
from ._abi import lib, check, REQ_DTYPE, REPLY_DTYPE, F_DEVICE_PTRS, F_ASYNC, F_GLOBAL_TABLE
from .snapshot import Image


class PinnedArray:
    """A numpy view over page-locked host memory from ptype_b200_alloc_pinned."""

    def __init__(self, count: int, dtype: np.dtype):
        self.nbytes = int(count) * np.dtype(dtype).itemsize
        self.ptr = lib.ptype_b200_alloc_pinned(max(self.nbytes, 1))
        if not self.ptr:
            raise MemoryError("ptype_b200_alloc_pinned failed")
        buf = (C.c_char * max(self.nbytes, 1)).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=dtype, count=count)

    def free(self) -> None:
        if self.ptr:
            self.array = None
            lib.ptype_b200_free_pinned(self.ptr)
            self.ptr = None


class DeviceBuffer:
    def __init__(self, ctx: "Context", dev: int, nbytes: int):
        self.ctx, self.dev, self.nbytes = ctx, dev, nbytes
        p = C.c_void_p()
        check(lib.ptype_b200_device_alloc(ctx._h, dev, nbytes, C.byref(p)), "device_alloc")
        self.ptr = p.value

    def upload(self, arr: np.ndarray) -> "DeviceBuffer":
        arr = np.ascontiguousarray(arr)
        assert arr.nbytes <= self.nbytes
        check(lib.ptype_b200_memcpy_h2d(self.ctx._h, self.dev, self.ptr, arr.ctypes.data, arr.nbytes),
              "memcpy_h2d")
        return self

    def download(self, dtype: np.dtype, count: int) -> np.ndarray:
        out = np.empty(count, dtype=dtype)
        assert out.nbytes <= self.nbytes
        check(lib.ptype_b200_memcpy_d2h(self.ctx._h, self.dev, out.ctypes.data, self.ptr, out.nbytes),
              "memcpy_d2h")
        return out

    def free(self) -> None:
        if self.ptr:
            lib.ptype_b200_device_free(self.ctx._h, self.dev, self.ptr)
            self.ptr = None


def _ptr(x) -> int:
    if isinstance(x, np.ndarray):
        return x.ctypes.data
    if isinstance(x, DeviceBuffer):
        return x.ptr
    return int(x)


class Context:
    def __init__(self, devices: Sequence[int] = (0,)):
        arr = (C.c_int * len(devices))(*devices)
        self._h = C.c_void_p()
        check(lib.ptype_b200_open(arr, len(devices), C.byref(self._h)), "open")
        self.devices = list(devices)
        self.image: Optional[Image] = None

    def close(self) -> None:
        if self._h:
            lib.ptype_b200_close(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # ---- snapshots -------------------------------------------------------
    def publish(self, image: Image) -> None:
        check(lib.ptype_b200_publish_image(self._h, image._buf, len(image)), "publish_image")
        self.image = image

    def update_service(self, service: str, delta: int) -> None:
        """A node joined (+1) or left (-1) `service`: patch the installed snapshot on the device."""
        check(lib.ptype_b200_update_service(self._h, service.encode(), delta), "update_service")

    # ---- the balancer's counters on the device (rpc.go:129,172-183) --------
    def assign_seq(self, reqs, n: int, *, dev: int = 0, sync: bool = True) -> None:
        """Fill the seq field of n device-resident records as n sequential Client.Call()s would."""
        check(lib.ptype_b200_assign_seq(self._h, dev, _ptr(reqs), n, 0 if sync else F_ASYNC), "assign_seq")

    def assign_seq_array(self, reqs, n: int, seq, *, dev: int = 0, sync: bool = True) -> None:
        """The same draw with the numbers written to the device array `seq` (n x uint64); records untouched."""
        check(lib.ptype_b200_assign_seq_array(self._h, dev, _ptr(reqs), n, _ptr(seq), 0 if sync else F_ASYNC),
              "assign_seq_array")

    def dispatch_seq(self, reqs, seq, n: int, replies, *, dev: int = 0, flags: int = 0, sync: bool = True) -> None:
        """Dispatch device-resident records whose sequence numbers lie in the device array `seq`."""
        check(lib.ptype_b200_dispatch_seq(self._h, dev, _ptr(reqs), _ptr(seq), n, _ptr(replies),
                                          flags | (0 if sync else F_ASYNC)), "dispatch_seq")

    def seq_get(self, service: str, dev: int = 0) -> int:
        v = C.c_uint64()
        check(lib.ptype_b200_seq_get(self._h, dev, service.encode(), C.byref(v)), "seq_get")
        return v.value

    def seq_set(self, service: str, value: int, dev: int = 0) -> None:
        check(lib.ptype_b200_seq_set(self._h, dev, service.encode(), value), "seq_set")

    def epoch(self, dev: int = 0) -> int:
        e = C.c_uint64()
        check(lib.ptype_b200_epoch(self._h, dev, C.byref(e)), "epoch")
        return e.value

    def current_image(self) -> Image:
        n = C.c_size_t(0)
        rc = lib.ptype_b200_current_image(self._h, None, C.byref(n))
        if rc not in (_abi.E_OK, _abi.E_LIMIT):
            check(rc, "current_image")
        buf = C.create_string_buffer(n.value)
        check(lib.ptype_b200_current_image(self._h, buf, C.byref(n)), "current_image")
        return Image(buf.raw[: n.value])

    def set_check_limit(self, iterations: int) -> None:
        check(lib.ptype_b200_set_check_limit(self._h, iterations), "set_check_limit")

    # ---- multi-process snapshot broadcast --------------------------------
    @staticmethod
    def nccl_unique_id() -> bytes:
        buf = C.create_string_buffer(128)
        check(lib.ptype_b200_nccl_unique_id(buf), "nccl_unique_id")
        return buf.raw

    def comm_init(self, unique_id: bytes, rank: int, nranks: int) -> None:
        buf = C.create_string_buffer(unique_id, 128)
        check(lib.ptype_b200_comm_init(self._h, buf, rank, nranks), "comm_init")

    def broadcast_image(self, image: Optional[Image], root: int = 0) -> Image:
        if image is not None:
            check(lib.ptype_b200_broadcast_image(self._h, image._buf, len(image), root),
                  "broadcast_image")
        else:
            check(lib.ptype_b200_broadcast_image(self._h, None, 0, root), "broadcast_image")
        self.image = self.current_image()
        return self.image

    # ---- dispatch ---------------------------------------------------------
    def dispatch(self, reqs: np.ndarray, replies: Optional[np.ndarray] = None, *, dev: int = 0,
                 flags: int = 0) -> np.ndarray:
        """Host buffers in, host buffers out (copies are inside the call)."""
        reqs = np.ascontiguousarray(reqs, dtype=REQ_DTYPE)
        if replies is None:
            replies = np.empty(len(reqs), dtype=REPLY_DTYPE)
        check(lib.ptype_b200_dispatch(self._h, dev, reqs.ctypes.data, len(reqs),
                                      replies.ctypes.data, flags), "dispatch")
        return replies

    def dispatch_ex(self, reqs: np.ndarray, failed_cap: int, *, dev: int = 0,
                    flags: int = 0) -> Tuple[np.ndarray, np.ndarray, _abi.BatchStats]:
        reqs = np.ascontiguousarray(reqs, dtype=REQ_DTYPE)
        replies = np.empty(len(reqs), dtype=REPLY_DTYPE)
        failed = np.zeros(max(failed_cap, 1), dtype=np.uint32)
        stats = _abi.BatchStats()
        check(lib.ptype_b200_dispatch_ex(self._h, dev, reqs.ctypes.data, len(reqs),
                                         replies.ctypes.data, failed.ctypes.data, failed_cap,
                                         C.addressof(stats), flags), "dispatch_ex")
        return replies, failed[: min(failed_cap, stats.failed_count)], stats

    def dispatch_device(self, reqs, n: int, replies, *, dev: int = 0, flags: int = 0,
                        failed=None, failed_cap: int = 0, stats=None, sync: bool = True) -> None:
        """Batch already resident in HBM: one kernel launch on the context stream."""
        f = flags | F_DEVICE_PTRS | (0 if sync else F_ASYNC)
        if stats is None:
            check(lib.ptype_b200_dispatch(self._h, dev, _ptr(reqs), n, _ptr(replies), f), "dispatch")
        else:
            check(lib.ptype_b200_dispatch_ex(self._h, dev, _ptr(reqs), n, _ptr(replies),
                                             _ptr(failed) if failed is not None else None,
                                             failed_cap, _ptr(stats), f), "dispatch_ex")

    # The two batch calls below are what a per-request host loop makes, so their Python cost is
    # kept off the path: method tables and the failed-call buffer are built once and reused, and
    # arrays that already have the right layout are passed through untouched.
    def _method_refs(self, methods):
        key = tuple(methods)
        cache = self.__dict__.setdefault("_refs_cache", {})
        refs = cache.get(key)
        if refs is None:
            refs = cache[key] = (_abi.MethodRef * len(key))(*[_abi.MethodRef(m, r) for m, r in key])
        return refs

    @staticmethod
    def _as(arr: np.ndarray, dtype) -> np.ndarray:
        return arr if arr.dtype == dtype and arr.flags.c_contiguous else np.ascontiguousarray(arr, dtype=dtype)

    def call_batch(self, service_key: int, seq_base: int, methods, args, method_sel=None, replies=None,
                   *, dev: int = 0, flags: int = 0, n: Optional[int] = None, args_per_call: int = 2):
        """One Client's batch in compact form (ptype_b200_call_batch).  methods: list of
        (method, recv) hash pairs; args: int64 array [n, 2|3] (host) or a device buffer with
        n/args_per_call given; method_sel: uint8 [n] or None."""
        refs = self._method_refs(methods)
        if isinstance(args, np.ndarray):
            args = self._as(args, np.int64)
            n, args_per_call = args.shape
            if method_sel is not None:
                method_sel = self._as(method_sel, np.uint8)
            if replies is None:
                replies = np.empty(n, dtype=REPLY_DTYPE)
        else:
            flags |= F_DEVICE_PTRS
        check(lib.ptype_b200_call_batch(self._h, dev, service_key, seq_base, refs, len(refs),
                                        _ptr(method_sel) if method_sel is not None else None,
                                        _ptr(args), args_per_call, n, _ptr(replies), flags), "call_batch")
        return replies

    def call_batch_values(self, service_key: int, seq_base: int, methods, args, method_sel=None,
                          values=None, failed_cap: int = 1024, *, dev: int = 0, flags: int = 0,
                          n: Optional[int] = None, args_per_call: int = 2):
        """ptype_b200_call_batch_values: returns (values, failed[(index, status)], failed_count)."""
        refs = self._method_refs(methods)
        if isinstance(args, np.ndarray):
            args = self._as(args, np.int64)
            n, args_per_call = args.shape
            if method_sel is not None:
                method_sel = self._as(method_sel, np.uint8)
            if values is None:
                values = np.empty(n, dtype=np.int64)
        else:
            flags |= F_DEVICE_PTRS
        failed = self.__dict__.get("_failed_buf")
        if failed is None or len(failed) < max(failed_cap, 1):
            failed = self._failed_buf = np.zeros(max(failed_cap, 1), dtype=[("index", "<u4"), ("status", "<u4")])
            self._failed_ptr = failed.ctypes.data
            self._failed_count = C.c_uint64()
        count = self._failed_count
        check(lib.ptype_b200_call_batch_values(self._h, dev, service_key, seq_base, refs, len(refs),
                                               _ptr(method_sel) if method_sel is not None else None,
                                               _ptr(args), args_per_call, n, _ptr(values),
                                               self._failed_ptr, failed_cap, C.byref(count), flags),
              "call_batch_values")
        taken = min(failed_cap, count.value)
        return values, (failed[:taken].copy() if taken else failed[:0]), count.value

    def make_batches(self, items) -> "C.Array":
        """items: iterable of (device reqs, device replies, n) -> a ptype_batch array."""
        items = list(items)
        arr = (_abi.Batch * len(items))()
        for i, (q, r, n) in enumerate(items):
            arr[i].reqs, arr[i].replies, arr[i].n = _ptr(q), _ptr(r), n
        return arr

    def dispatch_batches(self, batches, *, dev: int = 0, flags: int = 0, sync: bool = True) -> None:
        """Enqueue several device-resident batches with one C call (see make_batches)."""
        f = flags | F_DEVICE_PTRS | (0 if sync else F_ASYNC)
        check(lib.ptype_b200_dispatch_batches(self._h, dev, batches, len(batches), f),
              "dispatch_batches")

    # ---- formats either side of the path (device buffers) ----------------------
    def resolve_names(self, names: Sequence[str], stride: int = 32, *, dev: int = 0) -> np.ndarray:
        """"Receiver.Method" strings -> [(method, recv)] hashes computed on the device."""
        n = len(names)
        host = np.zeros((n, stride), dtype=np.uint8)
        for j, s in enumerate(names):
            b = s.encode()[:stride]
            host[j, : len(b)] = np.frombuffer(b, dtype=np.uint8)
        d_names = self.device_buffer(max(host.nbytes, 16), dev).upload(host)
        d_out = self.device_buffer(max(n, 1) * 8, dev)
        check(lib.ptype_b200_resolve_names(self._h, dev, d_names.ptr, stride, n, d_out.ptr, None),
              "resolve_names")
        out = d_out.download(np.dtype([("method", "<u4"), ("recv", "<u4")]), n)
        d_names.free()
        d_out.free()
        return out

    def gob_encode_replies(self, d_replies, n: int, *, dev: int = 0):
        """Reply values as gob messages: returns (offsets[n+1], bytes)."""
        d_off = self.device_buffer((n + 1) * 8, dev)
        total = C.c_uint64()
        check(lib.ptype_b200_gob_encode_replies(self._h, dev, _ptr(d_replies), n, d_off.ptr, None, 0,
                                                C.byref(total)), "gob_encode_replies")
        d_out = self.device_buffer(max(total.value, 1), dev)
        check(lib.ptype_b200_gob_encode_replies(self._h, dev, _ptr(d_replies), n, d_off.ptr, d_out.ptr,
                                                total.value, C.byref(total)), "gob_encode_replies")
        off = d_off.download(np.uint64, n + 1)
        data = d_out.download(np.uint8, total.value)
        d_off.free()
        d_out.free()
        return off, data

    def gob_decode_args(self, data: np.ndarray, offsets: np.ndarray, args_per_call: int, *, dev: int = 0):
        """gob Args bodies -> (args[n, args_per_call], status[n])."""
        n = len(offsets) - 1
        d_bytes = self.device_buffer(max(len(data), 1), dev).upload(np.ascontiguousarray(data, dtype=np.uint8))
        d_off = self.device_buffer((n + 1) * 8, dev).upload(np.ascontiguousarray(offsets, dtype=np.uint64))
        d_args = self.device_buffer(max(n, 1) * 8 * args_per_call, dev)
        d_st = self.device_buffer(max(n, 1), dev)
        check(lib.ptype_b200_gob_decode_args(self._h, dev, d_bytes.ptr, d_off.ptr, n, d_args.ptr,
                                             args_per_call, d_st.ptr), "gob_decode_args")
        args = d_args.download(np.int64, n * args_per_call).reshape(n, args_per_call)
        st = d_st.download(np.uint8, n)
        for b in (d_bytes, d_off, d_args, d_st):
            b.free()
        return args, st

    # ---- net/rpc byte streams (device buffers inside; host arrays in and out) ----
    WIRE_STREAM_DTYPE = np.dtype([("svc_key", "<u8"), ("seq0", "<u8"), ("seq_step", "<u8")])
    WIRE_CALL_DTYPE = np.dtype([("rpc_seq", "<u8"), ("sm_off", "<u8"), ("sm_len", "<u4"), ("flags", "<u4")])

    def wire_decode(self, data: np.ndarray, stream_off: np.ndarray, streams: np.ndarray, *, dev: int = 0,
                    repeat: int = 1):
        """Client-to-server net/rpc bytes of many connections -> (reqs, calls, call_off, seconds per pass)."""
        import time
        n_streams = len(stream_off) - 1
        data = np.ascontiguousarray(data, dtype=np.uint8)
        d_bytes = self.device_buffer(max(len(data), 1), dev).upload(data)
        d_soff = self.device_buffer((n_streams + 1) * 8, dev).upload(np.ascontiguousarray(stream_off, dtype=np.uint64))
        d_streams = self.device_buffer(max(n_streams, 1) * 24, dev).upload(
            np.ascontiguousarray(streams, dtype=self.WIRE_STREAM_DTYPE))
        d_coff = self.device_buffer((n_streams + 1) * 8, dev)
        total = C.c_uint64()
        rc = lib.ptype_b200_wire_decode(self._h, dev, d_bytes.ptr, d_soff.ptr, n_streams, d_streams.ptr, d_coff.ptr,
                                        None, None, 0, C.byref(total))
        if rc not in (_abi.E_OK, _abi.E_LIMIT):
            check(rc, "wire_decode")
        n = total.value
        d_reqs = self.device_buffer(max(n, 1) * 48, dev)
        d_calls = self.device_buffer(max(n, 1) * 24, dev)
        t0 = time.perf_counter()
        for _ in range(repeat):
            check(lib.ptype_b200_wire_decode(self._h, dev, d_bytes.ptr, d_soff.ptr, n_streams, d_streams.ptr,
                                             d_coff.ptr, d_reqs.ptr, d_calls.ptr, n, C.byref(total)), "wire_decode")
        seconds = (time.perf_counter() - t0) / repeat
        out = (d_reqs.download(REQ_DTYPE, n), d_calls.download(self.WIRE_CALL_DTYPE, n),
               d_coff.download(np.uint64, n_streams + 1), seconds)
        for b in (d_bytes, d_soff, d_streams, d_coff, d_reqs, d_calls):
            b.free()
        return out

    def wire_encode(self, request_bytes: np.ndarray, call_off: np.ndarray, calls: np.ndarray, replies: np.ndarray,
                    state: np.ndarray, *, response_id: int = 66, empty_id: int = 69, dev: int = 0, repeat: int = 1):
        """Reply records -> server-to-client net/rpc bytes: (bytes, out_off, new state, seconds per pass)."""
        import time
        n_streams = len(call_off) - 1
        n = len(calls)
        d_bytes = self.device_buffer(max(len(request_bytes), 1), dev).upload(
            np.ascontiguousarray(request_bytes, dtype=np.uint8))
        d_coff = self.device_buffer((n_streams + 1) * 8, dev).upload(np.ascontiguousarray(call_off, dtype=np.uint64))
        d_calls = self.device_buffer(max(n, 1) * 24, dev).upload(np.ascontiguousarray(calls, dtype=self.WIRE_CALL_DTYPE))
        d_rep = self.device_buffer(max(n, 1) * 16, dev).upload(np.ascontiguousarray(replies, dtype=REPLY_DTYPE))
        state0 = np.ascontiguousarray(state, dtype=np.uint8)
        d_state = self.device_buffer(max(n_streams, 1), dev).upload(state0)
        d_ooff = self.device_buffer((n_streams + 1) * 8, dev)
        total = C.c_uint64()
        rc = lib.ptype_b200_wire_encode(self._h, dev, d_bytes.ptr, d_coff.ptr, n_streams, d_calls.ptr, d_rep.ptr,
                                        response_id, empty_id, d_state.ptr, d_ooff.ptr, None, 0, C.byref(total))
        if rc not in (_abi.E_OK, _abi.E_LIMIT):
            check(rc, "wire_encode")
        d_out = self.device_buffer(max(total.value, 1), dev)
        seconds = 0.0
        for _ in range(repeat):
            d_state.upload(state0)
            t0 = time.perf_counter()
            check(lib.ptype_b200_wire_encode(self._h, dev, d_bytes.ptr, d_coff.ptr, n_streams, d_calls.ptr, d_rep.ptr,
                                             response_id, empty_id, d_state.ptr, d_ooff.ptr, d_out.ptr, total.value,
                                             C.byref(total)), "wire_encode")
            seconds += (time.perf_counter() - t0) / repeat
        out = (d_out.download(np.uint8, total.value), d_ooff.download(np.uint64, n_streams + 1),
               d_state.download(np.uint8, n_streams), seconds)
        for b in (d_bytes, d_coff, d_calls, d_rep, d_state, d_ooff, d_out):
            b.free()
        return out

    def sync(self, dev: int = 0) -> None:
        check(lib.ptype_b200_sync(self._h, dev), "sync")

    def stream(self, dev: int = 0) -> int:
        s = C.c_void_p()
        check(lib.ptype_b200_stream(self._h, dev, C.byref(s)), "stream")
        return s.value or 0

    def reply_checksum(self, replies, n: Optional[int] = None, *, index_base: int = 0, dev: int = 0,
                       device: bool = False) -> int:
        out = C.c_uint64()
        if not device:
            replies = np.ascontiguousarray(replies, dtype=REPLY_DTYPE)
            n = len(replies)
        check(lib.ptype_b200_reply_checksum(self._h, dev, _ptr(replies), n, index_base, C.byref(out),
                                            F_DEVICE_PTRS if device else 0), "reply_checksum")
        return out.value

    def launch_count(self) -> int:
        v = C.c_uint64()
        check(lib.ptype_b200_launch_count(self._h, C.byref(v)), "launch_count")
        return v.value

    def device_buffer(self, nbytes: int, dev: int = 0) -> DeviceBuffer:
        return DeviceBuffer(self, dev, nbytes)


def device_count() -> int:
    n = C.c_int()
    check(lib.ptype_b200_device_count(C.byref(n)), "device_count")
    return n.value


__all__ = ["Context", "DeviceBuffer", "PinnedArray", "device_count", "F_DEVICE_PTRS", "F_ASYNC",
           "F_GLOBAL_TABLE"]
