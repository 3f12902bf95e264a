"""Host logic of the multi-GPU path on CPU: world-size-2 gloo run of the image sharding + record gather."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ovmono3d_b200 import shard


def test_local_image_range_matches_inference_sampler():
    # detectron2 InferenceSampler: contiguous, first `n % world` ranks one longer, union = everything, no overlap
    for n in (0, 1, 7, 16, 20000):
        for world in (1, 2, 3, 8):
            seen = []
            for r in range(world):
                rg = shard.local_image_range(n, r, world)
                seen.extend(rg)
                assert len(rg) in (n // world, n // world + 1)
            assert seen == list(range(n))
    with pytest.raises(ValueError):
        shard.local_image_range(4, 2, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, num_images, boxes_per_image, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(5)
        full = torch.randn(sum(boxes_per_image), shard.RECORD_WIDTH, generator=g)     # same on every rank
        image_of = torch.repeat_interleave(torch.arange(num_images), torch.tensor(boxes_per_image))
        mine = shard.local_image_range(num_images, rank, world)
        mask = (image_of >= mine.start) & (image_of < mine.stop) if len(mine) else torch.zeros_like(image_of, dtype=torch.bool)
        rec, counts, idx = shard.gather_records(full[mask], image_of[mask])
        per_image = shard.split_by_image(rec, idx, num_images)
        ok = torch.equal(rec, full) and torch.equal(idx, image_of) and int(counts.sum()) == full.shape[0]
        ok = ok and [len(p) for p in per_image] == list(boxes_per_image)
        # without image ids
        rec2, counts2, idx2 = shard.gather_records(full[mask])
        ok = ok and torch.equal(rec2, full) and idx2 is None and torch.equal(counts, counts2)
        q.put((rank, bool(ok), counts.tolist()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("boxes_per_image", [(3, 0, 5, 2, 7), (0, 0, 4), (2,)])
def test_gather_records_world_size_2_gloo(boxes_per_image):
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, len(boxes_per_image), boxes_per_image, q))
             for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in results), results


def test_gather_records_single_process_is_identity():
    rec = torch.randn(5, shard.RECORD_WIDTH)
    out, counts, idx = shard.gather_records(rec)
    assert out is rec and counts.tolist() == [5] and idx is None
    with pytest.raises(ValueError):
        shard.gather_records(torch.zeros(3, 7))


def _exchange_worker(rank, world, port, counts, cap, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ex = shard.RecordExchange(cap, torch.device("cpu"), transport="nccl")     # "nccl" = the one-collective path (gloo here)
        g = torch.Generator().manual_seed(100)
        full = [torch.randn(c, shard.RECORD_WIDTH, generator=g) for c in counts]   # same on every rank
        for f in full:
            f[:, shard.RECORD_WIDTH - 1] = 0                                       # the pad column of a record is zero
        ok = ex.peer_ptrs(0) is None and ex.local.shape == (cap, shard.RECORD_WIDTH)
        ok = ok and ex.local.data_ptr() == ex.blocks[rank].data_ptr()              # written in place: no staging copy
        ex.local[: counts[rank]] = full[rank]
        blocks, cnt = ex.finish(counts[rank])
        ok = ok and cnt.tolist() == list(counts) and blocks.shape == (world, cap, shard.RECORD_WIDTH)
        dense = shard.RecordExchange.compact(blocks, cnt)
        ok = ok and torch.equal(dense, torch.cat(full))
        q.put((rank, bool(ok), cnt.tolist()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("counts,cap", [((5, 3), 8), ((0, 4), 4), ((6, 6), 6)])
def test_record_exchange_world_size_2_gloo(counts, cap):
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_exchange_worker, args=(r, world, port, counts, cap, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in results), results


def test_record_exchange_single_process():
    ex = shard.RecordExchange(4, torch.device("cpu"))
    ex.local[:3] = 1.0
    blocks, cnt = ex.finish(3)
    assert cnt.tolist() == [3] and shard.RecordExchange.compact(blocks, cnt).shape == (3, shard.RECORD_WIDTH)
    with pytest.raises(ValueError):
        ex.finish(5)
