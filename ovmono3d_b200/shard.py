"""Sharding of the lift across ranks: images are split by rank, detections are gathered once.

The reference runs inference data-parallel: `InferenceSampler` gives rank r a contiguous slice of the image list
(cubercnn/data/build.py:319-322, detectron2 `InferenceSampler._get_local_indices`) and the per-image results are
gathered once at the end (cubercnn/evaluation/omni3d_evaluation.py:717-720, a pickled-object gather over Gloo).
Here the result of the lift is the fixed-width fp32 record of `ovm_head_decode` (48 floats per box), so the gather is
one `all_gather_into_tensor` of a padded [max_boxes, 48] block plus the per-rank counts -- NCCL over NVLink for CUDA
tensors, gloo for CPU tensors (tests).  There is no collective inside the lift itself.
"""
from __future__ import annotations

import os
from typing import List, Optional, Tuple

import torch
import torch.distributed as dist

RECORD_WIDTH = 48


def local_image_range(num_images: int, rank: int, world_size: int) -> range:
    """Contiguous shard of rank `rank`: the first `num_images % world_size` ranks get one image more
    (detectron2 InferenceSampler._get_local_indices)."""
    if world_size <= 0 or not 0 <= rank < world_size:
        raise ValueError("rank %d not in a world of %d" % (rank, world_size))
    shard, left = divmod(num_images, world_size)
    sizes = [shard + (r < left) for r in range(world_size)]
    begin = sum(sizes[:rank])
    return range(begin, min(begin + sizes[rank], num_images))


def gather_records(records: torch.Tensor, image_index: Optional[torch.Tensor] = None,
                   group: Optional[dist.ProcessGroup] = None) -> Tuple[torch.Tensor, torch.Tensor, Optional[torch.Tensor]]:
    """All-gather the detection records of every rank.

    records [n_local, 48] fp32 (CUDA with NCCL, CPU with gloo); image_index [n_local] int64 = global image id of
    each record (optional).  Returns (all_records [n_total, 48] in rank order, counts [world] int64, all_image_index).
    Every rank gets the full result (the reference gathers to rank 0 only; an all-gather has the same cost on
    NVSwitch and needs no second code path)."""
    if records.dim() != 2 or records.shape[1] != RECORD_WIDTH:
        raise ValueError("records must be [n, %d], got %s" % (RECORD_WIDTH, tuple(records.shape)))
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return records, torch.tensor([records.shape[0]], dtype=torch.int64, device=records.device), image_index
    world = dist.get_world_size(group)
    device = records.device
    dbg = os.environ.get("OVM_GATHER_DEBUG") and device.type == "cuda"
    marks = []

    def mark(name):
        if dbg:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            marks.append((name, e))

    mark("start")
    n = torch.tensor([records.shape[0]], dtype=torch.int64, device=device)
    counts = torch.empty(world, dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(counts, n, group=group)
    mark("counts")
    cap = int(counts.max().item())          # the one host sync of the exchange; payload is ~19 KB per image
    width = RECORD_WIDTH + (1 if image_index is not None else 0)
    block = torch.zeros((cap, width), dtype=torch.float32, device=device)
    block[: records.shape[0], :RECORD_WIDTH] = records
    if image_index is not None:
        # image ids ride along as exact fp32 integers (< 2^24 images)
        block[: records.shape[0], RECORD_WIDTH] = image_index.to(torch.float32)
    mark("block")
    out = torch.empty((world * cap, width), dtype=torch.float32, device=device)
    dist.all_gather_into_tensor(out, block, group=group)
    mark("all_gather")
    out = out.view(world, cap, width)
    keep = torch.arange(cap, device=device).unsqueeze(0) < counts.unsqueeze(1)
    flat = out[keep]
    idx = flat[:, RECORD_WIDTH].to(torch.int64) if image_index is not None else None
    if dbg:
        mark("compact")
        torch.cuda.synchronize()
        import sys
        sys.stderr.write("gather_records: " + ", ".join("%s %.3f ms" % (b[0], a[1].elapsed_time(b[1]))
                                                         for a, b in zip(marks, marks[1:])) + "\n")
    return flat[:, :RECORD_WIDTH].contiguous(), counts, idx


def split_by_image(records: torch.Tensor, image_index: torch.Tensor, num_images: int) -> List[torch.Tensor]:
    """Per-image record lists in global image order (what the reference's evaluator consumes per image)."""
    order = torch.argsort(image_index, stable=True)
    rec, idx = records[order], image_index[order]
    counts = torch.bincount(idx, minlength=num_images).tolist()
    return list(torch.split(rec, counts))
