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
    import workloads as w
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


def test_one_process_two_devices(oracle=None):
    """A single host process driving two GPUs through one context (how a Go host with several
    GPUs would use the ABI): one publish installs the snapshot on both, dispatches carry the
    device index, and both devices answer like the oracle."""
    import ptype_b200 as p
    import workloads as w
    from oracle import binding as ob
    if p.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    spec = w.stress_registry(500, 4)
    kv, snap = ob.snapshot_from_spec(spec)
    with p.Context([0, 1]) as ctx:
        ctx.publish(spec.image(epoch=3))
        assert ctx.epoch(0) == 3 and ctx.epoch(1) == 3
        for dev in (0, 1):
            reqs = w.stress_calls(200_000, 500, seed=40 + dev)
            want = snap.dispatch(reqs, threads=4)
            assert ctx.dispatch(reqs, dev=dev).tobytes() == want.tobytes()
            d_req = ctx.device_buffer(len(reqs) * 48, dev).upload(reqs)
            d_rep = ctx.device_buffer(len(reqs) * 16, dev)
            ctx.dispatch_batches(ctx.make_batches([(d_req, d_rep, len(reqs))]), dev=dev)
            assert d_rep.download(p.REPLY_DTYPE, len(reqs)).tobytes() == want.tobytes()
            assert ctx.reply_checksum(d_rep, len(reqs), dev=dev, device=True) == ob.reply_checksum(want)
        ctx.update_service("svc_00007", -1)
        assert ctx.epoch(0) == 4 and ctx.epoch(1) == 4
