"""World-size-2 run of the N>1 host logic on CPU (gloo): rank 0 builds the registry image and
ships the blob, every rank validates it, takes its request-id-hash shard of the stream and
answers it (with the oracle here - there is no GPU), and the gathered replies and the summed
order-independent checksums equal the unsharded result."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, out_dir: str):
    sys.path.insert(0, str(ROOT))
    import torch
    import torch.distributed as dist
    import ptype_b200 as p
    import workloads as w
    from ptype_b200.sharding import shard_indices
    from oracle import binding as ob

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    spec = w.calculator_registry()
    # snapshot: rank 0 builds, everyone receives the same bytes (NCCL does this on GPUs)
    if rank == 0:
        blob = spec.image(epoch=42).blob
        size = torch.tensor([len(blob)], dtype=torch.int64)
    else:
        size = torch.zeros(1, dtype=torch.int64)
    dist.broadcast(size, src=0)
    buf = torch.frombuffer(bytearray(blob), dtype=torch.uint8) if rank == 0 else \
        torch.zeros(int(size.item()), dtype=torch.uint8)
    dist.broadcast(buf, src=0)
    image = p.Image(buf.numpy().tobytes())          # validates magic, bounds, checksum
    assert image.epoch == 42 and image.selected("calculator") == [3, 0, 9]

    n = 50_000
    stream = w.calculator_calls(n, salt=image.salt)
    idx = shard_indices(n, world, rank)
    kv, snap = ob.snapshot_from_spec(spec, image.salt)
    mine = snap.dispatch(stream[idx])
    # per-shard checksum under global stream positions
    part = 0
    runs = np.flatnonzero(np.diff(idx) != 1) + 1
    for lo, hi in zip(np.r_[0, runs], np.r_[runs, len(idx)]):
        part = (part + ob.reply_checksum(mine[lo:hi], int(idx[lo]))) & ((1 << 64) - 1)
    gathered = [None] * world
    dist.all_gather_object(gathered, (idx, mine.tobytes(), part))
    if rank == 0:
        whole = snap.dispatch(stream)
        out = np.empty_like(whole)
        total = 0
        seen = 0
        for gidx, raw, cks in gathered:
            out[gidx] = np.frombuffer(raw, dtype=ob.REPLY_DTYPE)
            total = (total + cks) & ((1 << 64) - 1)
            seen += len(gidx)
        assert seen == n
        assert out.tobytes() == whole.tobytes()
        assert total == ob.reply_checksum(whole)
        Path(out_dir, "ok").write_text("ok")
    dist.barrier()
    dist.destroy_process_group()


def test_world_size_2_sharded_dispatch(tmp_path):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok").read_text() == "ok"
