"""Request-id-hash sharding of a call stream over G GPUs.

Calls are independent units: a call reads its own record and the read-only
registry image, so the path partitions with no data-path collective.  Because the
round-robin sequence number travels in the record (cluster/rpc.go:181), sharding
changes no routing decision.
"""
from __future__ import annotations

import numpy as np


def mix64(x: np.ndarray) -> np.ndarray:
    """murmur3 fmix64, the same finaliser the kernels use (csrc/table_hash.h)."""
    x = np.asarray(x, dtype=np.uint64).copy()
    with np.errstate(over="ignore"):
        x ^= x >> np.uint64(33)
        x *= np.uint64(0xFF51AFD7ED558CCD)
        x ^= x >> np.uint64(33)
        x *= np.uint64(0xC4CEB9FE1A85EC53)
        x ^= x >> np.uint64(33)
    return x


def shard_of(request_ids: np.ndarray, n_gpus: int) -> np.ndarray:
    """gpu = mix64(request_id) % G."""
    return (mix64(request_ids) % np.uint64(n_gpus)).astype(np.int64)


def shard_indices(n: int, n_gpus: int, rank: int, id_base: int = 0) -> np.ndarray:
    """Positions (in stream order) of the calls rank `rank` owns among ids id_base..id_base+n-1."""
    ids = np.arange(id_base, id_base + n, dtype=np.uint64)
    return np.flatnonzero(shard_of(ids, n_gpus) == rank)
