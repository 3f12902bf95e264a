"""ORACLE (test infrastructure only) -- CPU restatement of the inference path of the reference's 2-D box head
(SURVEY 8f rank 2), with the same torch / torchvision calls in the same order:

  ROIHeads3D._forward_box, eval branch         cubercnn/modeling/roi_heads/roi_heads.py:252-296
  FastRCNNOutputs.inference                    cubercnn/modeling/roi_heads/fast_rcnn.py:121-143
  fast_rcnn_inference(_single_image)           fast_rcnn.py:16-116
  d2 FastRCNNConvFCHead (flatten, fc+ReLU x NUM_FC), FastRCNNOutputLayers.forward / predict_boxes / predict_probs,
  Box2BoxTransform.apply_deltas, Boxes.clip, layers.batched_nms -- detectron2 is an un-vendored submodule (fork
  pinned at setup.sh:2); these are restated from its published source; module shapes from nohup.out:645-661.

PARITY: `fast_rcnn_inference` is pinned to `tests/golden/box_inference_*.npz`, written by the reference's own
function (tests/golden/make_golden.py); the d2 pieces above it (softmax, apply_deltas) are restatements, unpinned.
Only `tests/`, `__graft_entry__.smoke()` and the CPU legs of `bench.py` may import this module.
"""
import math
from typing import Dict, List, Sequence, Tuple

import torch
from torch.nn import functional as F
from torchvision.ops import batched_nms as tv_batched_nms

from . import thirdparty as tp

SCALE_CLAMP = math.log(1000.0 / 16)      # d2 box_regression.py _DEFAULT_SCALE_CLAMP


def apply_deltas(deltas: torch.Tensor, boxes: torch.Tensor, weights: Sequence[float],
                 scale_clamp: float = SCALE_CLAMP) -> torch.Tensor:
    """d2 `Box2BoxTransform.apply_deltas`: deltas [R, 4k], boxes [R, 4] -> [R, 4k]."""
    deltas = deltas.float()
    boxes = boxes.to(deltas.dtype)
    widths = boxes[:, 2] - boxes[:, 0]
    heights = boxes[:, 3] - boxes[:, 1]
    ctr_x = boxes[:, 0] + 0.5 * widths
    ctr_y = boxes[:, 1] + 0.5 * heights
    wx, wy, ww, wh = weights
    dx = deltas[:, 0::4] / wx
    dy = deltas[:, 1::4] / wy
    dw = deltas[:, 2::4] / ww
    dh = deltas[:, 3::4] / wh
    dw = torch.clamp(dw, max=scale_clamp)
    dh = torch.clamp(dh, max=scale_clamp)
    pred_ctr_x = dx * widths[:, None] + ctr_x[:, None]
    pred_ctr_y = dy * heights[:, None] + ctr_y[:, None]
    pred_w = torch.exp(dw) * widths[:, None]
    pred_h = torch.exp(dh) * heights[:, None]
    x1 = pred_ctr_x - 0.5 * pred_w
    y1 = pred_ctr_y - 0.5 * pred_h
    x2 = pred_ctr_x + 0.5 * pred_w
    y2 = pred_ctr_y + 0.5 * pred_h
    pred_boxes = torch.stack((x1, y1, x2, y2), dim=-1)
    return pred_boxes.reshape(deltas.shape)


def predict(logits: torch.Tensor, deltas: torch.Tensor, proposals: Sequence[torch.Tensor],
            weights: Sequence[float]) -> Tuple[List[torch.Tensor], List[torch.Tensor]]:
    """d2 `FastRCNNOutputLayers.predict_boxes` + `predict_probs`: per-image splits of boxes [Ri, 4k], probs [Ri, K+1]."""
    num = [len(p) for p in proposals]
    boxes = apply_deltas(deltas, torch.cat(list(proposals), dim=0), weights)
    probs = F.softmax(logits, dim=-1)
    return list(boxes.split(num)), list(probs.split(num))


def fast_rcnn_inference_single_image(boxes, scores, image_shape, score_thresh, nms_thresh, topk_per_image):
    """fast_rcnn.py:57-116.  Returns (pred_boxes, scores, scores_full, pred_classes, kept proposal rows)."""
    valid_mask = torch.isfinite(boxes).all(dim=1) & torch.isfinite(scores).all(dim=1)
    if not valid_mask.all():
        boxes = boxes[valid_mask]
        scores = scores[valid_mask]
    scores = scores[:, :-1]
    num_bbox_reg_classes = boxes.shape[1] // 4
    boxes = boxes.reshape(-1, 4)
    h, w = image_shape                                   # Boxes.clip
    boxes = torch.stack((boxes[:, 0].clamp(min=0, max=w), boxes[:, 1].clamp(min=0, max=h),
                         boxes[:, 2].clamp(min=0, max=w), boxes[:, 3].clamp(min=0, max=h)), dim=-1)
    boxes = boxes.view(-1, num_bbox_reg_classes, 4)
    filter_mask = scores > score_thresh
    filter_inds = filter_mask.nonzero()
    if num_bbox_reg_classes == 1:
        boxes = boxes[filter_inds[:, 0], 0]
    else:
        boxes = boxes[filter_mask]
    scores_full = scores[filter_inds[:, 0]]
    scores = scores[filter_mask]
    keep = tv_batched_nms(boxes.float(), scores, filter_inds[:, 1], nms_thresh)   # d2 layers.batched_nms
    if topk_per_image >= 0:
        keep = keep[:topk_per_image]
    boxes, scores, filter_inds, scores_full = boxes[keep], scores[keep], filter_inds[keep], scores_full[keep]
    return boxes, scores, scores_full, filter_inds[:, 1], filter_inds[:, 0]


def fast_rcnn_inference(boxes, scores, image_shapes, score_thresh, nms_thresh, topk_per_image):
    """fast_rcnn.py:16-55."""
    return [fast_rcnn_inference_single_image(b, s, shp, score_thresh, nms_thresh, topk_per_image)
            for s, b, shp in zip(scores, boxes, image_shapes)]


def box_head_features(features: List[torch.Tensor], proposals: Sequence[torch.Tensor], scales: Sequence[float],
                      weights: Dict[str, torch.Tensor], pooler_resolution: int = 7, sampling_ratio: int = 0,
                      bf16_operands: bool = False) -> torch.Tensor:
    """box_pooler -> FastRCNNConvFCHead (roi_heads.py:274-276): [R, fc_dim]."""
    pooler = tp.ROIPoolerOracle(pooler_resolution, scales, sampling_ratio, "ROIAlignV2")
    x = pooler(features, list(proposals)).flatten(1)
    k = 1
    while "box_head.fc%d.weight" % k in weights:
        w, b = weights["box_head.fc%d.weight" % k], weights["box_head.fc%d.bias" % k]
        if bf16_operands:
            x = F.relu(F.linear(x.bfloat16().float(), w.bfloat16().float(), b))
        else:
            x = F.relu(F.linear(x, w, b))
        k += 1
    return x


def box_predictor(x: torch.Tensor, weights: Dict[str, torch.Tensor], bf16_operands: bool = False):
    """d2 `FastRCNNOutputLayers.forward`: (scores [R, K+1], proposal_deltas [R, 4K or 4])."""
    if bf16_operands:
        lin = lambda n: F.linear(x.bfloat16().float(), weights["box_predictor.%s.weight" % n].bfloat16().float(),
                                 weights["box_predictor.%s.bias" % n])
    else:
        lin = lambda n: F.linear(x, weights["box_predictor.%s.weight" % n], weights["box_predictor.%s.bias" % n])
    return lin("cls_score"), lin("bbox_pred")
