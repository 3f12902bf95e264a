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


class RecordExchange:
    """Fixed-capacity all-gather of detection records without a host round trip.

    Every rank owns a block of `cap` rows x 48 floats; `local` is where its lift calls write (`out=` of
    `LiftEngine.forward`, row by row as the batches come).  `finish(count)` completes the exchange and returns
    (`blocks` [world, cap, 48], `counts` [world] int64 on the device) -- rows past `counts[r]` of block r are undefined.
    No `.item()`, no zero-filled staging block, no compaction pass: `compact()` does that lazily where a consumer
    wants one dense [n_total, 48] tensor.  Two transports:

      * "nccl"  (and gloo for CPU tensors): ONE `all_gather_into_tensor` of the [cap, 48] blocks; the count rides in the
        pad column (47) of the block's first row.
      * "peer": no data collective at all.  The blocks live in symmetric memory (torch.distributed._symmetric_memory:
        every rank's buffer is mapped into every other rank's address space over NVLink); `peer_ptrs(row)` gives
        `LiftEngine.forward(peer_ptrs=...)` the addresses of this rank's slot in all the OTHER ranks' buffers, so the
        decode epilogue stores each 192-byte record straight into its final place on every GPU while the lift is still
        running (`ovm_lift_forward_peers`); `finish` is one barrier on the symmetric-memory signal pads.

    The reference's exchange is `comm.gather` of pickled per-image lists over a Gloo side group, once per dataset
    (cubercnn/evaluation/omni3d_evaluation.py:717-720)."""

    def __init__(self, cap: int, device: torch.device, transport: str = "nccl", group: Optional[dist.ProcessGroup] = None):
        self.cap, self.device, self.group = int(cap), torch.device(device), group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.transport = transport if self.world > 1 else "local"
        self._hdl = None
        if self.transport == "peer":
            import torch.distributed._symmetric_memory as symm_mem

            if self.device.type != "cuda":
                raise ValueError("the peer transport needs CUDA tensors")
            # rows: [world, cap + 1, 48]; the extra last row of every block carries that rank's count
            self._buf = symm_mem.empty((self.world, self.cap + 1, RECORD_WIDTH), dtype=torch.float32, device=self.device)
            self._hdl = symm_mem.rendezvous(self._buf, dist.group.WORLD if group is None else group)
            self.blocks = self._buf[:, : self.cap]
            self.local = self._buf[self.rank, : self.cap]
            self._slot_bytes = (self.cap + 1) * RECORD_WIDTH * 4
            self._count_ptrs = [int(p) + self.rank * self._slot_bytes + self.cap * RECORD_WIDTH * 4
                                for p in self._hdl.buffer_ptrs]
        else:
            self.blocks = torch.empty((self.world, self.cap, RECORD_WIDTH), dtype=torch.float32, device=self.device)
            # the send block is the rank's own row of the result: no copy before the collective, none after
            self.local = self.blocks[self.rank]

    def peer_ptrs(self, row: int) -> Optional[List[int]]:
        """Addresses (valid on this device) of row `row` of this rank's block in every other rank's buffer."""
        if self.transport != "peer":
            return None
        off = self.rank * self._slot_bytes + row * RECORD_WIDTH * 4
        return [int(p) + off for r, p in enumerate(self._hdl.buffer_ptrs) if r != self.rank]

    def finish(self, count: int) -> Tuple[torch.Tensor, torch.Tensor]:
        if count > self.cap:
            raise ValueError("%d records exceed the exchange capacity %d" % (count, self.cap))
        if self.transport == "local":
            return self.blocks, torch.tensor([count], dtype=torch.int64, device=self.device)
        if self.transport == "peer":
            from . import ops

            # this rank's count into every rank's buffer (4-byte posted writes, one launch), then one barrier on the
            # signal pads: every rank's records and counts have landed everywhere
            ops.peer_store_f32(self._count_ptrs, float(count))
            self._hdl.barrier(channel=0)
            counts = self._buf[:, self.cap, 0].to(torch.int64)
            return self.blocks, counts
        if self.cap:
            self.local[0, RECORD_WIDTH - 1].fill_(float(count))     # the record's pad column carries the count
        dist.all_gather_into_tensor(self.blocks.view(self.world * self.cap, RECORD_WIDTH), self.local, group=self.group)
        counts = self.blocks[:, 0, RECORD_WIDTH - 1].to(torch.int64) if self.cap else torch.zeros(
            self.world, dtype=torch.int64, device=self.device)
        return self.blocks, counts

    @staticmethod
    def compact(blocks: torch.Tensor, counts: torch.Tensor) -> torch.Tensor:
        """[world, cap, 48] + counts -> dense [n_total, 48] in rank order (this is where the host learns the sizes)."""
        world, cap, _ = blocks.shape
        keep = torch.arange(cap, device=blocks.device).unsqueeze(0) < counts.unsqueeze(1)
        out = blocks[keep].clone()
        if cap:     # the pad column is zero in every record; row 0 of a block carried the count on the nccl transport
            out[:, RECORD_WIDTH - 1] = 0
        return out


def split_by_image(records: torch.Tensor, image_index: torch.Tensor, num_images: int) -> List[torch.Tensor]:
    """Per-image record lists in global image order (what the reference's evaluator consumes per image)."""
    order = torch.argsort(image_index, stable=True)
    rec, idx = records[order], image_index[order]
    counts = torch.bincount(idx, minlength=num_images).tolist()
    return list(torch.split(rec, counts))
