"""Detection records -> the reference's result schema.

After the lift every detection is one fixed-width fp32 record (`ops.RECORD_WIDTH` = 48 floats: corners 8x3 | centre 3 |
centre-2D 2 | W,H,L 3 | R 9 | 2-D box xyxy 4 | score | class | 0) -- the unit that is all-gathered across ranks
(`shard.gather_records`).  This module turns a batch of records into what the reference writes to disk:

  * `GeneralizedRCNN._postprocess` / d2 `detector_postprocess` (reference cubercnn/modeling/meta_arch/rcnn3d.py:113-115):
    2-D boxes rescaled from the network resolution to the original image, clipped, empty boxes dropped;
  * `Omni3DEvaluator.instances_to_coco_json` (cubercnn/evaluation/omni3d_evaluation.py:1200-1252): one dict per
    detection with `image_id, category_id, bbox` (XYWH, original-resolution pixels), `score, bbox3D, center_cam,
    center_2D, dimensions, pose, depth`.

The reference does this with one `.numpy()` / `.tolist()` per field and per image plus a Python loop per box; here it is
one device->host copy of the records of the whole batch and vectorised fp32 numpy arithmetic in the reference's order
of operations, so the numbers in the JSON are bit-identical.
"""
from __future__ import annotations

from typing import Any, Dict, List, Sequence, Tuple, Union

import numpy as np
import torch

from . import ops


def postprocess_boxes(boxes_xyxy: np.ndarray, image_size: Tuple[int, int], output_size: Tuple[int, int]
                      ) -> Tuple[np.ndarray, np.ndarray]:
    """d2 `detector_postprocess` on the 2-D boxes: scale (network -> original resolution), clip, non-empty mask.

    `image_size` = (H, W) the detector ran at (`Instances.image_size`), `output_size` = (height, width) of the original
    image (`batched_inputs[i]["height"/"width"]`).  fp32 throughout, like `Boxes.scale` / `Boxes.clip`."""
    out_h, out_w = output_size
    scale_x = np.float32(out_w / image_size[1])
    scale_y = np.float32(out_h / image_size[0])
    b = boxes_xyxy.astype(np.float32, copy=True)
    b[:, 0::2] *= scale_x
    b[:, 1::2] *= scale_y
    b[:, 0] = np.clip(b[:, 0], 0, out_w)
    b[:, 1] = np.clip(b[:, 1], 0, out_h)
    b[:, 2] = np.clip(b[:, 2], 0, out_w)
    b[:, 3] = np.clip(b[:, 3], 0, out_h)
    keep = ((b[:, 2] - b[:, 0]) > 0) & ((b[:, 3] - b[:, 1]) > 0)
    return b, keep


def records_to_coco_json(records: Union[torch.Tensor, np.ndarray], num_per_image: Sequence[int],
                         image_ids: Sequence[Any], image_sizes: Sequence[Tuple[int, int]],
                         output_sizes: Sequence[Tuple[int, int]]) -> List[List[Dict[str, Any]]]:
    """[n, 48] records of a batch (image-major) -> per image, the list `instances_to_coco_json` returns.

    One device->host copy for the whole batch; `image_ids[i]` is used as given (the reference maps file names to
    integer ids before this point, omni3d_evaluation.py:1207-1217)."""
    if isinstance(records, torch.Tensor):
        records = records.detach().to("cpu", torch.float32).numpy()
    records = np.ascontiguousarray(records, dtype=np.float32)
    if records.ndim != 2 or records.shape[1] != ops.RECORD_WIDTH:
        raise ops.LiftError("records must be [n, %d]" % ops.RECORD_WIDTH)
    if sum(num_per_image) != records.shape[0]:
        raise ops.LiftError("num_per_image sums to %d but there are %d records" % (sum(num_per_image), records.shape[0]))
    out: List[List[Dict[str, Any]]] = []
    start = 0
    for i, k in enumerate(num_per_image):
        rec = records[start:start + k]
        start += k
        if k == 0:
            out.append([])
            continue
        boxes, keep = postprocess_boxes(rec[:, ops.REC_BOX2D], image_sizes[i], output_sizes[i])
        rec, boxes = rec[keep], boxes[keep]
        if len(rec) == 0:
            out.append([])
            continue
        boxes[:, 2] -= boxes[:, 0]      # BoxMode.convert(XYXY_ABS -> XYWH_ABS)
        boxes[:, 3] -= boxes[:, 1]
        bbox = boxes.tolist()
        score = rec[:, ops.REC_SCORE].tolist()
        cls = rec[:, ops.REC_CLASS].astype(np.int64).tolist()
        bbox3d = rec[:, ops.REC_BBOX3D].reshape(-1, 8, 3).tolist()
        center_cam = rec[:, ops.REC_CENTER_CAM].tolist()
        center_2d = rec[:, ops.REC_CENTER_2D].tolist()
        dims = rec[:, ops.REC_DIMENSIONS].tolist()
        pose = rec[:, ops.REC_POSE].reshape(-1, 3, 3).tolist()
        img = []
        for j in range(len(rec)):
            img.append({"image_id": image_ids[i], "category_id": cls[j], "bbox": bbox[j], "score": score[j],
                        "bbox3D": bbox3d[j], "center_cam": center_cam[j], "center_2D": center_2d[j],
                        "dimensions": dims[j], "pose": pose[j], "depth": center_cam[j][2]})
        out.append(img)
    return out
