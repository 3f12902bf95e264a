"""The cross-rank exchange of the detection records on N GPUs (torchrun), three ways, timed and checked:

  legacy   shard.gather_records: two all-gathers, a .item() host sync, a zero-filled block, a mask compaction (round 1)
  nccl     shard.RecordExchange("nccl"): ONE all_gather_into_tensor of fixed-capacity blocks, no host sync
  peer     shard.RecordExchange("peer"): no data collective -- the lift's decode stores every record into every rank's
           symmetric-memory block over NVLink while it runs (ovm_lift_forward_peers); finish = one signal-pad barrier

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tools/gather_probe.py [--steps 20] > gpurun_out/gather_probe_n2.json

Per variant: K lift steps on the C2 batch writing into the exchange, then the exchange; device time of the exchange
alone (events, max over ranks) and of lift + exchange; the gathered records are compared with each rank's own.
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from ovmono3d_b200 import ops, shard  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    args = ap.parse_args()
    K = args.steps
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    world, rank = dist.get_world_size(), dist.get_rank()
    head, _ = bench.build_head(dev)
    db = bench.DeviceBatch(bench.make_batch(1235 + 1000 * rank), dev)
    n = db.n
    eng = head._get_engine()
    eng.prepare(dev, head.priors_dims_per_cat, head.priors_z_scales, getattr(head, "priors_z_stats", None))

    def maxr(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    def run(variant):
        ex = None if variant == "legacy" else shard.RecordExchange(K * n, dev, transport=variant)
        local_buf = torch.empty(K * n, ops.RECORD_WIDTH, device=dev) if ex is None else ex.local
        res = {}
        for rep in range(3):       # two warm-ups (communicator / symmetric-memory set-up), one timed
            dist.barrier()
            torch.cuda.synchronize()
            e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            e[0].record()
            for i in range(K):
                out = local_buf[i * n:(i + 1) * n]
                eng.forward(db.features, db.boxes, db.classes, db.scores, db.batch_idx, db.cams, out=out,
                            peer_ptrs=ex.peer_ptrs(i * n) if ex is not None else None)
            e[1].record()
            if ex is None:
                dense, counts, _ = shard.gather_records(local_buf)
            else:
                blocks, counts = ex.finish(K * n)
            e[2].record()
            torch.cuda.synchronize()
            res = {"lift_ms": maxr(e[0].elapsed_time(e[1])), "exchange_ms": maxr(e[1].elapsed_time(e[2])),
                   "total_ms": maxr(e[0].elapsed_time(e[2]))}
        # correctness: every rank's block, as every rank sees it, equals that rank's own records
        if ex is not None:
            dense = shard.RecordExchange.compact(blocks, counts)
        mine = local_buf.clone()
        mine[:, ops.RECORD_WIDTH - 1] = 0
        allmine = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allmine, mine)
        ok = counts.tolist() == [K * n] * world and torch.equal(dense, torch.cat(allmine))
        res["correct"] = bool(ok)
        res["bytes_per_rank"] = K * n * ops.RECORD_WIDTH * 4
        return res

    out = {"world": world, "steps": K, "boxes_per_step": n}
    for variant in ("legacy", "nccl", "peer"):
        try:
            out[variant] = run(variant)
        except Exception as ex:      # e.g. symmetric memory unavailable on this box
            out[variant] = {"error": "%s: %s" % (type(ex).__name__, ex)}
        dist.barrier()
    if rank == 0:
        print(json.dumps(out))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
