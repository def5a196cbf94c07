"""Two GPUs, one process each: the registry snapshot travels by ncclBroadcast through the
C ABI (ptype_b200_comm_init / ptype_b200_broadcast_image), each rank dispatches its
request-id-hash shard, and the gathered replies equal the oracle's - before and after a
live snapshot update (BASELINE config 5 in miniature).  Skipped on single-GPU boxes."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
pytestmark = pytest.mark.gpu


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, out_dir: str):
    sys.path.insert(0, str(ROOT))
    import torch.distributed as dist
    import ptype_b200 as p
    from ptype_b200 import workloads as w
    from ptype_b200.sharding import shard_indices
    from oracle import binding as ob

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)   # host-side plumbing only
    ctx = p.Context([rank])
    ids = [p.Context.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ids, src=0)
    ctx.comm_init(ids[0], rank, world)

    n = 400_000
    specs = [w.stress_registry(300, 8), w.stress_registry(300, 5)]   # epoch 2: nodes 5..7 gone
    for epoch, spec in enumerate(specs, start=1):
        image = ctx.broadcast_image(spec.image(epoch=epoch) if rank == 0 else None, root=0)
        assert ctx.epoch() == epoch and image.epoch == epoch
        if rank == 0:
            assert image.blob == spec.image(epoch=epoch).blob
        stream = w.stress_calls(n, 300, salt=image.salt, seed=77 + epoch)
        idx = shard_indices(n, world, rank)
        mine = ctx.dispatch(stream[idx])
        gathered = [None] * world
        dist.all_gather_object(gathered, (idx, mine.tobytes(), image.blob))
        if rank == 0:
            kv, snap = ob.snapshot_from_spec(spec, image.salt)
            whole = snap.dispatch(stream)
            out = np.empty_like(whole)
            for gidx, raw, blob in gathered:
                out[gidx] = np.frombuffer(raw, dtype=ob.REPLY_DTYPE)
                assert blob == image.blob            # every GPU holds the same snapshot bytes
            assert out.tobytes() == whole.tobytes()
    ctx.close()
    if rank == 0:
        Path(out_dir, "ok").write_text("ok")
    dist.barrier()
    dist.destroy_process_group()


def test_two_gpu_broadcast_and_sharded_dispatch(tmp_path):
    import ptype_b200 as p
    if p.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok").read_text() == "ok"
