"""ORACLE (test infrastructure only) -- CPU restatement of what the reference does with the lifted Instances after
`_forward_cube`: d2 `detector_postprocess` (called by `GeneralizedRCNN._postprocess`, reference
cubercnn/modeling/meta_arch/rcnn3d.py:113-115) and `Omni3DEvaluator.instances_to_coco_json`
(cubercnn/evaluation/omni3d_evaluation.py:1200-1252), written with the same torch / numpy calls in the same order.

detectron2 is an un-vendored submodule (fork pinned in setup.sh:2); `Boxes.scale`, `Boxes.clip`, `Boxes.nonempty` and
`BoxMode.convert(XYXY_ABS -> XYWH_ABS)` are restated from its published source.  PARITY: unpinned by the reference's own
tests (it has none for this path).  Only `tests/` may import this module.
"""
from typing import Any, Dict, List

import torch


def detector_postprocess(fields: Dict[str, torch.Tensor], image_size, output_height: int, output_width: int):
    """d2 modeling/postprocessing.py: scale pred_boxes to the output resolution, clip, keep the non-empty ones."""
    scale_x, scale_y = output_width / image_size[1], output_height / image_size[0]
    boxes = fields["pred_boxes"].clone()
    boxes[:, 0::2] *= scale_x          # Boxes.scale
    boxes[:, 1::2] *= scale_y
    h, w = output_height, output_width  # Boxes.clip
    x1 = boxes[:, 0].clamp(min=0, max=w)
    y1 = boxes[:, 1].clamp(min=0, max=h)
    x2 = boxes[:, 2].clamp(min=0, max=w)
    y2 = boxes[:, 3].clamp(min=0, max=h)
    boxes = torch.stack((x1, y1, x2, y2), dim=-1)
    keep = ((boxes[:, 2] - boxes[:, 0]) > 0) & ((boxes[:, 3] - boxes[:, 1]) > 0)   # Boxes.nonempty(threshold=0)
    out = {k: v[keep] for k, v in fields.items()}
    out["pred_boxes"] = boxes[keep]
    return out


def instances_to_coco_json(fields: Dict[str, torch.Tensor], img_id: Any) -> List[Dict[str, Any]]:
    """omni3d_evaluation.py:1200-1252 (integer image ids)."""
    num_instance = len(fields["pred_boxes"])
    if num_instance == 0:
        return []
    boxes = fields["pred_boxes"].numpy().copy()
    boxes[:, 2] -= boxes[:, 0]         # BoxMode.convert(boxes, XYXY_ABS, XYWH_ABS)
    boxes[:, 3] -= boxes[:, 1]
    boxes = boxes.tolist()
    scores = fields["scores"].tolist()
    classes = fields["pred_classes"].tolist()
    bbox3D = fields["pred_bbox3D"].numpy()
    center_cam = fields["pred_center_cam"].numpy()
    center_2D = fields["pred_center_2D"].numpy()
    dimensions = fields["pred_dimensions"].numpy()
    pose = fields["pred_pose"].numpy()
    results = []
    for k in range(num_instance):
        result = {"image_id": img_id, "category_id": classes[k], "bbox": boxes[k], "score": scores[k]}
        result["bbox3D"] = bbox3D[k].tolist()
        result["center_cam"] = center_cam[k].tolist()
        result["center_2D"] = center_2D[k].tolist()
        result["dimensions"] = dimensions[k].tolist()
        result["pose"] = pose[k].tolist()
        result["depth"] = float(center_cam[k][2])
        results.append(result)
    return results
