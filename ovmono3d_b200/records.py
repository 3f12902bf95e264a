"""Detection records -> the reference's result schema.

After the lift every detection is one fixed-width fp32 record (`ops.RECORD_WIDTH` = 48 floats: corners 8x3 | centre 3 |
centre-2D 2 | W,H,L 3 | R 9 | 2-D box xyxy 4 | score | class | 0) -- the unit that is exchanged across ranks
(`shard.RecordExchange`).  This module turns a batch of records into what the reference writes to disk:

  * `GeneralizedRCNN._postprocess` / d2 `detector_postprocess` (reference cubercnn/modeling/meta_arch/rcnn3d.py:113-115):
    2-D boxes rescaled from the network resolution to the original image, clipped, empty boxes dropped;
  * `Omni3DEvaluator.instances_to_coco_json` (cubercnn/evaluation/omni3d_evaluation.py:1200-1252): one dict per
    detection with `image_id, category_id, bbox` (XYWH, original-resolution pixels), `score, bbox3D, center_cam,
    center_2D, dimensions, pose, depth`.

The reference does this per image on the host: a `.to('cpu')` of every field, `.numpy()` / `.tolist()` per field and a
Python loop per box.  Here the arithmetic (scale, clip, non-empty filter, XYXY -> XYWH) and the compaction run in one
kernel over the whole batch (`ovm_records_postprocess`, csrc/records.cu, bit-identical fp32 arithmetic), followed by ONE
device -> host copy of the surviving rows; what is left on the host is `tolist()` and assembling the dicts the schema
asks for.  There is no host implementation of the arithmetic: records on the CPU are an error.
"""
from __future__ import annotations

from typing import Any, Dict, List, Sequence, Tuple

import numpy as np
import torch

from . import ops

KEYS = ("image_id", "category_id", "bbox", "score", "bbox3D", "center_cam", "center_2D", "dimensions", "pose", "depth")


def postprocess_geometry(image_sizes: Sequence[Tuple[int, int]], output_sizes: Sequence[Tuple[int, int]]) -> np.ndarray:
    """[N, 4] fp32 (scale_x, scale_y, out_w, out_h) for `ovm_records_postprocess`.  `image_sizes[i]` = (H, W) the detector
    ran at (`Instances.image_size`), `output_sizes[i]` = (height, width) of the original image.  The factors are computed
    in double and rounded to fp32, which is what `Boxes.scale` does with its Python-float arguments."""
    g = np.empty((len(image_sizes), 4), dtype=np.float32)
    for i, ((h, w), (oh, ow)) in enumerate(zip(image_sizes, output_sizes)):
        g[i] = (np.float32(ow / w), np.float32(oh / h), ow, oh)
    return g


def rows_to_coco_json(rows: np.ndarray, counts: Sequence[int], offsets: Sequence[int], image_ids: Sequence[Any]
                      ) -> List[List[Dict[str, Any]]]:
    """Post-processed rows (host copy of `ovm_records_postprocess`'s output: XYWH boxes in columns 41..44) -> per image,
    the list `instances_to_coco_json` returns.  Formatting only: `tolist()` per field for the whole batch, then the
    dicts."""
    out: List[List[Dict[str, Any]]] = []
    for i, k in enumerate(counts):
        k = int(k)
        if k == 0:
            out.append([])
            continue
        r = rows[int(offsets[i]):int(offsets[i]) + k]
        center_cam = r[:, ops.REC_CENTER_CAM].tolist()
        cols = ([image_ids[i]] * k, r[:, ops.REC_CLASS].astype(np.int64).tolist(), r[:, ops.REC_BOX2D].tolist(),
                r[:, ops.REC_SCORE].tolist(), r[:, ops.REC_BBOX3D].reshape(k, 8, 3).tolist(), center_cam,
                r[:, ops.REC_CENTER_2D].tolist(), r[:, ops.REC_DIMENSIONS].tolist(),
                r[:, ops.REC_POSE].reshape(k, 3, 3).tolist(), [c[2] for c in center_cam])
        out.append([dict(zip(KEYS, vals)) for vals in zip(*cols)])
    return out


def records_to_coco_json(records: torch.Tensor, num_per_image: Sequence[int], image_ids: Sequence[Any],
                         image_sizes: Sequence[Tuple[int, int]], output_sizes: Sequence[Tuple[int, int]]
                         ) -> List[List[Dict[str, Any]]]:
    """[n, 48] CUDA records of a batch (image-major) -> per image, the list `instances_to_coco_json` returns.

    `image_ids[i]` is used as given (the reference maps file names to integer ids before this point,
    omni3d_evaluation.py:1207-1217)."""
    if not isinstance(records, torch.Tensor) or not records.is_cuda:
        raise ops.LiftError("records must be a CUDA tensor: the post-processing arithmetic runs in ovm_records_postprocess "
                            "(there is no host implementation)")
    if records.dim() != 2 or records.shape[1] != ops.RECORD_WIDTH:
        raise ops.LiftError("records must be [n, %d]" % ops.RECORD_WIDTH)
    if sum(num_per_image) != records.shape[0]:
        raise ops.LiftError("num_per_image sums to %d but there are %d records" % (sum(num_per_image), records.shape[0]))
    N = len(num_per_image)
    if N == 0:
        return []
    offsets = np.zeros(N + 1, dtype=np.int32)
    np.cumsum(np.asarray(num_per_image, dtype=np.int64), out=offsets[1:])
    if records.shape[0] == 0:
        return [[] for _ in range(N)]
    dev = records.device
    rows, counts = ops.records_postprocess(records.float().contiguous(), torch.from_numpy(offsets).to(dev),
                                           torch.from_numpy(postprocess_geometry(image_sizes, output_sizes)).to(dev))
    return rows_to_coco_json(rows.cpu().numpy(), counts.cpu().tolist(), offsets[:-1], image_ids)
