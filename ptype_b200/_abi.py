"""ctypes binding of the C ABI declared in include/ptype_b200.h.

This is the same boundary a Go host reaches through cgo; the
Python side only marshals numpy buffers.  There is no fallback: if the CUDA
library was not built, importing this module raises.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
LIB_PATH = _HERE / "libptype_b200.so"

# Record layouts (include/ptype_b200.h: ptype_req 48 B, ptype_reply 16 B).
REQ_DTYPE = np.dtype(
    [("svc_key", "<u8"), ("seq", "<u8"), ("method", "<u4"), ("recv", "<u4"),
     ("a0", "<i8"), ("a1", "<i8"), ("a2", "<i8")], align=False)
REPLY_DTYPE = np.dtype([("value", "<i8"), ("node", "<u4"), ("status", "<u4")], align=False)
assert REQ_DTYPE.itemsize == 48 and REPLY_DTYPE.itemsize == 16

ST_OK, ST_NO_SERVICE, ST_NO_CLIENT, ST_ILL_FORMED = 0, 1, 2, 3
ST_CANT_FIND_SERVICE, ST_CANT_FIND_METHOD, ST_BAD_INDEX, ST_RANGE_LIMIT = 4, 5, 6, 7
ST_COUNT = 8
NO_NODE = 0xFFFFFFFF
RECV_ILL_FORMED = 0xFFFFFFFF

F_DEVICE_PTRS, F_ASYNC, F_GLOBAL_TABLE, F_LDG_LOADS, F_BULK_LOADS, F_ONE_LAUNCH_PER_BATCH = 1, 2, 4, 8, 16, 32
F_AFTER_PRODUCER = 64

E_OK, E_INVAL, E_NOMEM, E_CUDA, E_NO_IMAGE, E_BAD_IMAGE = 0, 1, 2, 3, 4, 5
E_COLLISION, E_NCCL, E_LIMIT, E_NOT_FOUND, E_BUSY = 6, 7, 8, 9, 10


class BatchStats(C.Structure):
    _fields_ = [("calls", C.c_uint64), ("by_status", C.c_uint64 * ST_COUNT),
                ("failed_count", C.c_uint64)]


class MethodRef(C.Structure):
    _fields_ = [("method", C.c_uint32), ("recv", C.c_uint32)]


class Batch(C.Structure):
    _fields_ = [("reqs", C.c_void_p), ("replies", C.c_void_p), ("n", C.c_size_t)]


class ImageHeader(C.Structure):
    _fields_ = [
        ("magic", C.c_uint32), ("version", C.c_uint32), ("epoch", C.c_uint64),
        ("total_bytes", C.c_uint64), ("body_checksum", C.c_uint64), ("key_salt", C.c_uint64),
        ("n_services", C.c_uint32), ("n_slots", C.c_uint32), ("n_nodes", C.c_uint32),
        ("max_connections", C.c_uint32), ("n_conn_hash", C.c_uint32), ("n_methods", C.c_uint32),
        ("n_receivers", C.c_uint32), ("off_slots", C.c_uint32), ("off_conn_hash", C.c_uint32),
        ("off_methods", C.c_uint32), ("off_nodes", C.c_uint32), ("off_services", C.c_uint32),
        ("off_strings", C.c_uint32), ("strings_bytes", C.c_uint32), ("device_bytes", C.c_uint32),
        ("table_seed", C.c_uint32), ("method_mul", C.c_uint32), ("off_mtable", C.c_uint32),
        ("host_sections", C.c_uint32), ("off_order", C.c_uint32), ("reserved", C.c_uint32 * 2), ("local_addr", C.c_char * 64),
    ]


assert C.sizeof(ImageHeader) == 192

# Every symbol include/ptype_b200.h declares: (name, restype, argtypes).
_vp, _u64p, _u32p = C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint32)
SIGNATURES = [
    ("ptype_b200_strerror", C.c_char_p, [C.c_int]),
    ("ptype_b200_last_error", C.c_char_p, []),
    ("ptype_b200_abi_version", C.c_int, []),
    ("ptype_b200_fnv1a32", C.c_uint32, [_vp, C.c_size_t]),
    ("ptype_b200_fnv1a64", C.c_uint64, [_vp, C.c_size_t]),
    ("ptype_b200_service_key", C.c_uint64, [C.c_char_p, C.c_uint64]),
    ("ptype_b200_method_key", None, [C.c_char_p, _u32p, _u32p]),
    ("ptype_b200_conn_hash", C.c_uint32, [C.c_char_p, C.c_int]),
    ("ptype_b200_builder_new", C.c_int, [C.c_char_p, C.c_int, C.POINTER(_vp)]),
    ("ptype_b200_builder_register", C.c_int, [_vp, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int]),
    ("ptype_b200_builder_unregister", C.c_int, [_vp, C.c_char_p, C.c_char_p]),
    ("ptype_b200_builder_set_device_only", C.c_int, [_vp, C.c_int]),
    ("ptype_b200_builder_declare", C.c_int, [_vp, C.c_char_p]),
    ("ptype_b200_builder_expose", C.c_int, [_vp, C.c_char_p, C.c_char_p]),
    ("ptype_b200_builder_set_slots", C.c_int, [_vp, C.c_uint32]),
    ("ptype_b200_builder_finish", C.c_int, [_vp, C.c_uint64, C.POINTER(_vp), C.POINTER(C.c_size_t)]),
    ("ptype_b200_builder_free", None, [_vp]),
    ("ptype_b200_image_free", None, [_vp]),
    ("ptype_b200_image_validate", C.c_int, [_vp, C.c_size_t]),
    ("ptype_b200_image_selected", C.c_int, [_vp, C.c_char_p, _u32p, C.c_size_t, _u32p]),
    ("ptype_b200_image_node", C.c_int,
     [_vp, C.c_uint32, C.c_char_p, C.c_size_t, C.POINTER(C.c_int), C.c_char_p, C.c_size_t,
      C.c_char_p, C.c_size_t]),
    ("ptype_b200_open", C.c_int, [C.POINTER(C.c_int), C.c_int, C.POINTER(_vp)]),
    ("ptype_b200_close", None, [_vp]),
    ("ptype_b200_device_count", C.c_int, [C.POINTER(C.c_int)]),
    ("ptype_b200_alloc_pinned", _vp, [C.c_size_t]),
    ("ptype_b200_free_pinned", None, [_vp]),
    ("ptype_b200_device_alloc", C.c_int, [_vp, C.c_int, C.c_size_t, C.POINTER(_vp)]),
    ("ptype_b200_device_free", C.c_int, [_vp, C.c_int, _vp]),
    ("ptype_b200_memcpy_h2d", C.c_int, [_vp, C.c_int, _vp, _vp, C.c_size_t]),
    ("ptype_b200_memcpy_d2h", C.c_int, [_vp, C.c_int, _vp, _vp, C.c_size_t]),
    ("ptype_b200_stream", C.c_int, [_vp, C.c_int, C.POINTER(_vp)]),
    ("ptype_b200_publish_image", C.c_int, [_vp, _vp, C.c_size_t]),
    ("ptype_b200_update_service", C.c_int, [_vp, C.c_char_p, C.c_int]),
    ("ptype_b200_epoch", C.c_int, [_vp, C.c_int, _u64p]),
    ("ptype_b200_current_image", C.c_int, [_vp, _vp, C.POINTER(C.c_size_t)]),
    ("ptype_b200_nccl_unique_id", C.c_int, [_vp]),
    ("ptype_b200_comm_init", C.c_int, [_vp, _vp, C.c_int, C.c_int]),
    ("ptype_b200_broadcast_image", C.c_int, [_vp, _vp, C.c_size_t, C.c_int]),
    ("ptype_b200_dispatch", C.c_int, [_vp, C.c_int, _vp, C.c_size_t, _vp, C.c_uint32]),
    ("ptype_b200_dispatch_ex", C.c_int,
     [_vp, C.c_int, _vp, C.c_size_t, _vp, _vp, C.c_size_t, _vp, C.c_uint32]),
    ("ptype_b200_dispatch_batches", C.c_int, [_vp, C.c_int, _vp, C.c_size_t, C.c_uint32]),
    ("ptype_b200_call_batch", C.c_int,
     [_vp, C.c_int, C.c_uint64, C.c_uint64, _vp, C.c_int, _vp, _vp, C.c_int, C.c_size_t, _vp, C.c_uint32]),
    ("ptype_b200_call_batch_values", C.c_int,
     [_vp, C.c_int, C.c_uint64, C.c_uint64, _vp, C.c_int, _vp, _vp, C.c_int, C.c_size_t, _vp, _vp,
      C.c_size_t, _u64p, C.c_uint32]),
    ("ptype_b200_sync", C.c_int, [_vp, C.c_int]),
    ("ptype_b200_set_check_limit", C.c_int, [_vp, C.c_uint64]),
    ("ptype_b200_reply_checksum", C.c_int,
     [_vp, C.c_int, _vp, C.c_size_t, C.c_uint64, _u64p, C.c_uint32]),
    ("ptype_b200_reply_checksum_noalloc", C.c_int,
     [C.c_int, _vp, C.c_size_t, C.c_uint64, _vp, _vp]),
    ("ptype_b200_resolve_names", C.c_int, [_vp, C.c_int, _vp, C.c_uint32, C.c_size_t, _vp, _vp]),
    ("ptype_b200_gob_encode_replies", C.c_int, [_vp, C.c_int, _vp, C.c_size_t, _vp, _vp, C.c_size_t, _u64p]),
    ("ptype_b200_gob_decode_args", C.c_int, [_vp, C.c_int, _vp, _vp, C.c_size_t, _vp, C.c_int, _vp]),
    ("ptype_b200_wire_decode", C.c_int, [_vp, C.c_int, _vp, _vp, C.c_size_t, _vp, _vp, _vp, _vp, C.c_size_t, _u64p]),
    ("ptype_b200_wire_encode", C.c_int,
     [_vp, C.c_int, _vp, _vp, C.c_size_t, _vp, _vp, C.c_uint32, C.c_uint32, _vp, _vp, _vp, C.c_size_t, _u64p]),
    ("ptype_b200_assign_seq", C.c_int, [_vp, C.c_int, _vp, C.c_size_t, C.c_uint32]),
    ("ptype_b200_assign_seq_array", C.c_int, [_vp, C.c_int, _vp, C.c_size_t, _vp, C.c_uint32]),
    ("ptype_b200_dispatch_seq", C.c_int, [_vp, C.c_int, _vp, _vp, C.c_size_t, _vp, C.c_uint32]),
    ("ptype_b200_seq_get", C.c_int, [_vp, C.c_int, C.c_char_p, _u64p]),
    ("ptype_b200_seq_set", C.c_int, [_vp, C.c_int, C.c_char_p, C.c_uint64]),
    ("ptype_b200_launch_count", C.c_int, [_vp, _u64p]),
]


class PtypeError(RuntimeError):
    def __init__(self, code: int, where: str, detail: str = ""):
        self.code = code
        super().__init__(f"{where}: error {code}" + (f" ({detail})" if detail else ""))


def _load() -> C.CDLL:
    path = Path(os.environ.get("PTYPE_B200_LIB", LIB_PATH))
    if not path.exists():
        raise ImportError(
            f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C ptype_b200/csrc`. There is no CPU fallback for the dispatch path.")
    lib = C.CDLL(str(path), mode=C.RTLD_GLOBAL)
    for name, restype, argtypes in SIGNATURES:
        fn = getattr(lib, name)  # AttributeError here = header and library disagree
        fn.restype = restype
        fn.argtypes = argtypes
    return lib


lib = _load()


def check(code: int, where: str) -> None:
    if code != E_OK:
        text = lib.ptype_b200_strerror(code).decode()
        last = lib.ptype_b200_last_error().decode()
        raise PtypeError(code, where, text + (": " + last if last and code in (E_CUDA, E_NCCL) else ""))
