"""Registry snapshot images (host side, no GPU needed).

`ImageBuilder` collects what `Registry.Register` wrote (cluster/registry.go:51-86)
and emits the blob the kernels read; `Image` answers the host-side questions the
reference answers with `getSelectedNodes()` (cluster/rpc.go:293-297) and
`Registry.Services()` (cluster/registry.go:93-117).
"""
from __future__ import annotations

import ctypes as C
from typing import Iterable, List, Optional, Tuple

from . import _abi
from ._abi import lib, check


def service_key(name: str, salt: int = 0) -> int:
    return lib.ptype_b200_service_key(name.encode(), salt)


def method_key(service_method: str) -> Tuple[int, int]:
    """(method, recv) hashes of a "Receiver.Method" string, split at the last '.'."""
    m, r = C.c_uint32(), C.c_uint32()
    lib.ptype_b200_method_key(service_method.encode(), C.byref(m), C.byref(r))
    return m.value, r.value


def conn_hash(local_addr: str, i: int) -> int:
    """FNV-1a-32(localAddr + strconv.Itoa(i)) (cluster/rpc.go:266-270)."""
    return lib.ptype_b200_conn_hash(local_addr.encode(), i)


class Image:
    """An immutable registry image blob."""

    def __init__(self, blob: bytes):
        self.blob = bytes(blob)
        self._buf = C.create_string_buffer(self.blob, len(self.blob))
        check(lib.ptype_b200_image_validate(self._buf, len(self.blob)), "image_validate")
        self.header = _abi.ImageHeader.from_buffer_copy(self.blob[:192])

    def __len__(self) -> int:
        return len(self.blob)

    @property
    def epoch(self) -> int:
        return self.header.epoch

    @property
    def salt(self) -> int:
        return self.header.key_salt

    def key(self, service: str) -> int:
        return service_key(service, self.salt)

    def selected(self, service: str) -> Optional[List[int]]:
        """Global node ids a Client of `service` holds; None if not in the snapshot."""
        k = C.c_uint32()
        cap = 1 << 12
        while True:
            out = (C.c_uint32 * cap)()
            rc = lib.ptype_b200_image_selected(self._buf, service.encode(), out, cap, C.byref(k))
            if rc == _abi.E_NOT_FOUND:
                return None
            check(rc, "image_selected")
            if k.value <= cap:
                return list(out[: k.value])
            cap = k.value

    def node(self, node_id: int) -> Tuple[str, int, str, str]:
        """(address, port, service, node_name) of a global node id."""
        a, s, nm = C.create_string_buffer(256), C.create_string_buffer(256), C.create_string_buffer(256)
        port = C.c_int()
        check(lib.ptype_b200_image_node(self._buf, node_id, a, 256, C.byref(port), s, 256, nm, 256),
              "image_node")
        return a.value.decode(), port.value, s.value.decode(), nm.value.decode()


class ImageBuilder:
    """local_addr / max_connections: the Client host's address string and
    ConnConfig.MaxConnections (cluster/rpc.go:19-38; default 3, 0 = mesh)."""

    def __init__(self, local_addr: str = "", max_connections: int = 3):
        self._b = C.c_void_p()
        check(lib.ptype_b200_builder_new(local_addr.encode(), max_connections, C.byref(self._b)),
              "builder_new")

    def register(self, service: str, node_name: str, address: str, port: int) -> "ImageBuilder":
        check(lib.ptype_b200_builder_register(self._b, service.encode(), node_name.encode(),
                                              address.encode(), port), "builder_register")
        return self

    def unregister(self, service: str, node_name: str) -> "ImageBuilder":
        check(lib.ptype_b200_builder_unregister(self._b, service.encode(), node_name.encode()),
              "builder_unregister")
        return self

    def set_device_only(self, on: bool = True) -> "ImageBuilder":
        check(lib.ptype_b200_builder_set_device_only(self._b, 1 if on else 0), "builder_set_device_only")
        return self

    def declare(self, service: str) -> "ImageBuilder":
        check(lib.ptype_b200_builder_declare(self._b, service.encode()), "builder_declare")
        return self

    def expose(self, service: str, receiver: str) -> "ImageBuilder":
        check(lib.ptype_b200_builder_expose(self._b, service.encode(), receiver.encode()),
              "builder_expose")
        return self

    def set_slots(self, n_slots: int) -> "ImageBuilder":
        check(lib.ptype_b200_builder_set_slots(self._b, n_slots), "builder_set_slots")
        return self

    def finish(self, epoch: int = 1) -> Image:
        blob, n = C.c_void_p(), C.c_size_t()
        check(lib.ptype_b200_builder_finish(self._b, epoch, C.byref(blob), C.byref(n)),
              "builder_finish")
        try:
            return Image(C.string_at(blob, n.value))
        finally:
            lib.ptype_b200_image_free(blob)

    def close(self) -> None:
        if self._b:
            lib.ptype_b200_builder_free(self._b)
            self._b = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def build_image(nodes: Iterable[Tuple[str, str, str, int]], *, local_addr: str = "",
                max_connections: int = 3, expose: Iterable[Tuple[str, str]] = (),
                declare: Iterable[str] = (), epoch: int = 1, n_slots: int = 0) -> Image:
    """nodes: (service, node_name, address, port) tuples, as passed to Register."""
    b = ImageBuilder(local_addr, max_connections)
    try:
        for svc, name, addr, port in nodes:
            b.register(svc, name, addr, port)
        for svc in declare:
            b.declare(svc)
        for svc, recv in expose:
            b.expose(svc, recv)
        if n_slots:
            b.set_slots(n_slots)
        return b.finish(epoch)
    finally:
        b.close()
