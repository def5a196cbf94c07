"""ptype_b200 — B200-native batched RPC dispatch for edegens/ptype (one hot path).

Host-side mirror of the reference's `cluster` package surface for that path, over the
C ABI in include/ptype_b200.h.  Importing this package loads libptype_b200.so and
raises if it has not been built: there is no CPU fallback.
"""
from ._abi import (REQ_DTYPE, REPLY_DTYPE, PtypeError, BatchStats, LIB_PATH,
                   ST_OK, ST_NO_SERVICE, ST_NO_CLIENT, ST_ILL_FORMED, ST_CANT_FIND_SERVICE,
                   ST_CANT_FIND_METHOD, ST_BAD_INDEX, ST_RANGE_LIMIT, NO_NODE, RECV_ILL_FORMED,
                   F_DEVICE_PTRS, F_ASYNC, F_GLOBAL_TABLE, F_LDG_LOADS, F_BULK_LOADS, F_ONE_LAUNCH_PER_BATCH,
                   F_AFTER_PRODUCER)
from .snapshot import Image, ImageBuilder, build_image, service_key, method_key, conn_hash
from .context import Context, DeviceBuffer, PinnedArray, device_count

__all__ = [n for n in dir() if not n.startswith("_")]
