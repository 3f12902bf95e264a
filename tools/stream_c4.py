"""BASELINE configs[3] (C4): an Omni3D-test-scale synthetic stream -- 20 000 images x 100 boxes in batches of 16 per
rank, images sharded contiguously by rank (shard.local_image_range = the reference's InferenceSampler), detections
all-gathered over NCCL either once at the end (what the reference does, omni3d_evaluation.py:717-720) or after every
batch (on a side stream, so that the exchange of batch i overlaps the lift of batch i + 1).  Device-resident inputs rotating over three 352 MB pyramids (larger than L2), no flush, whole loop between one
CUDA-event pair per rank, max over ranks.  Rank 0 prints one JSON line.

    python tools/stream_c4.py                                   # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/stream_c4.py
"""
import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402
from ovmono3d_b200 import ops, shard  # noqa: E402


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--images", type=int, default=20000)
    ap.add_argument("--batch", type=int, default=16)
    a = ap.parse_args()
    import torch.distributed as dist

    rank, world, local = bench._dist_env()
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    ops.load_library()
    head, _ = bench.build_head(device)
    dbs = [bench.DeviceBatch(bench.make_batch(4235 + 1000 * rank + i), device) for i in range(3)]
    n = dbs[0].n
    eng = head._get_engine()
    eng.prepare(device, head.priors_dims_per_cat, head.priors_z_scales, getattr(head, "priors_z_stats", None))
    mine = shard.local_image_range(a.images, rank, world)
    steps = len(mine) // a.batch                     # whole batches only (20 000 / 16 / world)
    records_all = torch.empty((steps, n, ops.RECORD_WIDTH), device=device, dtype=torch.float32)
    gathered_step = [torch.empty((world * n, ops.RECORD_WIDTH), device=device, dtype=torch.float32) for _ in range(2)]
    side = torch.cuda.Stream(device=device)          # the per-batch exchange overlaps the next batch's lift

    def lift(i, out):
        db = dbs[i % 3]
        return eng.forward(db.features, db.boxes, db.classes, db.scores, db.batch_idx, db.cams, out=out)[0]

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_ms(x):
        return bench._max_over_ranks(x, world, device)

    for i in range(5):
        lift(i, records_all[i])
    if world > 1:   # communicator, and the caching allocator's blocks for the full-size gather, before the timed loops
        shard.gather_records(records_all.view(-1, ops.RECORD_WIDTH))
        dist.all_gather_into_tensor(gathered_step[0], records_all[0])
    res = {}
    for mode in ("gather_at_end", "gather_every_batch"):
        sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for i in range(steps):
            lift(i, records_all[i])
            if mode == "gather_every_batch" and world > 1:
                side.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(side):
                    dist.all_gather_into_tensor(gathered_step[i & 1], records_all[i])
        torch.cuda.current_stream().wait_stream(side)
        if mode == "gather_at_end" and world > 1:
            out, counts, _ = shard.gather_records(records_all.view(-1, ops.RECORD_WIDTH))
            assert out.shape[0] == world * steps * n
        e1.record()
        sync()
        wall = (time.perf_counter() - t0) * 1e3
        ms = max_ms(e0.elapsed_time(e1))
        boxes = world * steps * n
        res[mode] = {"ms": ms, "wall_ms_rank0": wall, "boxes_per_s": boxes / ms * 1e3,
                     "images_per_s": world * steps * a.batch / ms * 1e3}
    if rank == 0:
        print(json.dumps({"workload": "C4: %d images x 100 boxes, batches of %d per rank, C2 shapes" % (a.images, a.batch),
                          "n_gpus": world, "steps_per_rank": steps, "images_processed": world * steps * a.batch,
                          "record_bytes_per_image": 100 * ops.RECORD_WIDTH * 4, **res}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
