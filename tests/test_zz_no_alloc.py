"""ptype_b200_reply_checksum_noalloc: the checksum for hosts that may not allocate.

Some ptype hosts cannot let a device allocation happen once they are serving: cudaMalloc
and cudaFree hold up all streams of the GPU, and a process that owns no spare HBM (its
own pool took it, or a quota caps it) may be refused outright.  This entry point keeps
its partial sums in an array built into the library instead of allocating them.

The file name sorts last on purpose: with `pytest -x` every other test has run before a
failure here can stop the session.  It also passes when run alone in a fresh process.
"""
import ctypes as C

import numpy as np
import pytest

import ptype_b200 as p
from ptype_b200 import _abi

pytestmark = pytest.mark.gpu

MAX_REPLIES = 16384 * 4096  # stated maximum of the built-in scratch array


def random_replies(n: int, seed: int) -> np.ndarray:
    rng = np.random.default_rng(seed)
    rep = np.zeros(n, dtype=p.REPLY_DTYPE)
    rep["value"] = rng.integers(-(1 << 63), (1 << 63) - 1, n, dtype=np.int64)
    rep["node"] = rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32)
    rep["status"] = rng.integers(0, 8, n, dtype=np.uint32)
    return rep


def noalloc_checksum(ctx, d_rep, n, index_base, d_out, device=0) -> int:
    """Call the entry point through the C ABI on the context's stream, then read the result."""
    rc = _abi.lib.ptype_b200_reply_checksum_noalloc(device, d_rep.ptr, n, index_base, d_out.ptr, ctx.stream())
    assert rc == 0, _abi.lib.ptype_b200_last_error().decode()
    ctx.sync()
    return int(d_out.download(np.uint64, 1)[0])


def test_noalloc_matches_allocating_entry_point_twice(oracle):
    """Two calls in a row — the second smaller than the first, so stale partial sums from
    the first call sit in the shared array — must both equal ptype_b200_reply_checksum
    exactly (integer result: the tolerance is zero, as in every other test here), and
    both must equal the checksum the CPU oracle computes for the same records."""
    ctx = p.Context([0])
    try:
        d_out = ctx.device_buffer(8)
        for n, seed, base in ((3_000_003, 1, 0), (70_001, 2, (1 << 63) + 9), (0, 3, 5), (4096, 4, 77)):
            rep = random_replies(n, seed)
            d_rep = ctx.device_buffer(max(n, 1) * 16).upload(rep) if n else ctx.device_buffer(16)
            want = ctx.reply_checksum(d_rep, n, index_base=base, device=True)
            assert want == oracle.reply_checksum(rep, base)
            assert noalloc_checksum(ctx, d_rep, n, base, d_out) == want
            assert noalloc_checksum(ctx, d_rep, n, base, d_out) == want   # and again, same array
            d_rep.free()
        d_out.free()
    finally:
        ctx.close()


def test_noalloc_rejects_sizes_beyond_the_scratch_array():
    ctx = p.Context([0])
    try:
        d = ctx.device_buffer(16)
        rc = _abi.lib.ptype_b200_reply_checksum_noalloc(0, d.ptr, MAX_REPLIES + 1, 0, d.ptr, None)
        assert rc == _abi.E_LIMIT
        msg = _abi.lib.ptype_b200_last_error().decode()
        assert "beyond" in msg and str(MAX_REPLIES) in msg
        assert _abi.lib.ptype_b200_reply_checksum_noalloc(0, d.ptr, 16, 0, None, None) == _abi.E_INVAL
        assert _abi.lib.ptype_b200_reply_checksum_noalloc(-1, d.ptr, 16, 0, d.ptr, None) == _abi.E_INVAL
        d.free()
    finally:
        ctx.close()


def test_noalloc_on_a_gpu_that_is_not_the_current_one(oracle):
    """Each GPU has its own copy of the built-in array, looked up on that GPU's first call.
    Asking for GPU 1 while GPU 0 is the thread's current device (and the other way round) must
    use the right copy and leave the caller's current device as it was.  Needs two GPUs."""
    import torch
    if p.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    rep = random_replies(1_234_567, 9)
    want = oracle.reply_checksum(rep, 42)
    ctxs = {dev: p.Context([dev]) for dev in (0, 1)}
    try:
        for current, target in ((0, 1), (1, 0), (1, 1), (0, 0)):
            torch.cuda.set_device(current)
            ctx = ctxs[target]
            d_rep = ctx.device_buffer(len(rep) * 16).upload(rep)
            d_out = ctx.device_buffer(8)
            torch.cuda.set_device(current)   # the context calls above may have moved it
            for _ in range(2):
                assert noalloc_checksum(ctx, d_rep, len(rep), 42, d_out, device=target) == want
            d_rep.free()
            d_out.free()
        torch.cuda.set_device(0)
        before = torch.cuda.current_device()
        d = ctxs[1].device_buffer(16)
        torch.cuda.set_device(0)
        assert _abi.lib.ptype_b200_reply_checksum_noalloc(1, d.ptr, 0, 0, d.ptr, None) == 0
        assert torch.cuda.current_device() == before == 0
        d.free()
    finally:
        for c in ctxs.values():
            c.close()
