"""Time the pieces of shard.gather_records on N GPUs (torchrun): the two all-gathers, the host sync, the compaction."""
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ovmono3d_b200 import shard  # noqa: E402


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    world = dist.get_world_size()
    for n in (1600, 48000):
        rec = torch.randn(n, 48, device=dev)
        out = torch.empty(world * n, 48, device=dev)
        for _ in range(3):
            dist.all_gather_into_tensor(out, rec)
            shard.gather_records(rec)
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        t0 = time.perf_counter()
        e[0].record()
        for _ in range(10):
            dist.all_gather_into_tensor(out, rec)
        e[1].record()
        for _ in range(10):
            shard.gather_records(rec)
        e[2].record()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        if dist.get_rank() == 0:
            print("n=%d world=%d: raw all_gather %.3f ms each, gather_records %.3f ms each (wall %.1f ms for 20 calls)" % (
                n, world, e[0].elapsed_time(e[1]) / 10, e[1].elapsed_time(e[2]) / 10, (t1 - t0) * 1e3))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
