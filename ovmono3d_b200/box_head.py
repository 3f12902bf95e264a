"""The 2-D box branch of `ROIHeads3D` at inference (reference roi_heads.py:252-296, fast_rcnn.py:16-143) on the
lift's kernels: box_pooler -> FastRCNNConvFCHead -> FastRCNNOutputs -> fast_rcnn_inference.

Module and parameter names are the reference's (`box_head.fc1`, `box_head.fc2`, `box_predictor.cls_score`,
`box_predictor.bbox_pred`; shapes nohup.out:645-661), so its checkpoints load unchanged.  The compute is
`ovm_roi_align` (bf16 pooled features) -> `ovm_gemm_bf16` per FC layer (tcgen05, bf16 operands, fp32 accumulate) ->
one `ovm_gemm_bf16` for both predictor layers (their weights stacked: K+1 class rows, then the 4K / 4 box rows) ->
`ovm_box_inference` (softmax, apply_deltas, clip, threshold, per-class NMS, top-k: three launches).  There is no
CPU or PyTorch fallback.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import torch
from torch import nn

from . import ops
from .compat import Boxes, Instances, ShapeSpec


class FastRCNNConvFCHead(nn.Sequential):
    """d2 `FastRCNNConvFCHead` with NUM_CONV = 0 (the only shipped form, Base.yaml:67-70): flatten, (fc, ReLU) x NUM_FC."""

    def __init__(self, input_shape: ShapeSpec, num_fc: int, fc_dim: int):
        super().__init__()
        self.add_module("flatten", nn.Flatten())
        dim = input_shape.channels * (input_shape.height or 1) * (input_shape.width or 1)
        for k in range(num_fc):
            fc = nn.Linear(dim, fc_dim)
            nn.init.kaiming_uniform_(fc.weight, a=1)          # c2_xavier_fill
            nn.init.constant_(fc.bias, 0)
            self.add_module("fc%d" % (k + 1), fc)
            self.add_module("fc_relu%d" % (k + 1), nn.ReLU())
            dim = fc_dim
        self.output_dim = dim


class FastRCNNOutputs(nn.Module):
    """`FastRCNNOutputs(FastRCNNOutputLayers)` (fast_rcnn.py:119): the two linear layers and the test-time settings."""

    def __init__(self, input_dim: int, num_classes: int, cls_agnostic_bbox_reg: bool, box2box_weights: Sequence[float],
                 test_score_thresh: float, test_nms_thresh: float, test_topk_per_image: int):
        super().__init__()
        self.num_classes = num_classes
        self.cls_agnostic_bbox_reg = bool(cls_agnostic_bbox_reg)
        self.cls_score = nn.Linear(input_dim, num_classes + 1)
        self.bbox_pred = nn.Linear(input_dim, 4 if cls_agnostic_bbox_reg else 4 * num_classes)
        nn.init.normal_(self.cls_score.weight, std=0.01)
        nn.init.normal_(self.bbox_pred.weight, std=0.001)
        for layer in (self.cls_score, self.bbox_pred):
            nn.init.constant_(layer.bias, 0)
        self.box2box_weights = tuple(float(w) for w in box2box_weights)
        self.scale_clamp = math.log(1000.0 / 16)
        self.test_score_thresh = float(test_score_thresh)
        self.test_nms_thresh = float(test_nms_thresh)
        self.test_topk_per_image = int(test_topk_per_image)


class BoxBranchB200:
    """Runs `_forward_box` for a head that owns `box_pooler`, `box_head`, `box_predictor`."""

    # torchvision's batched_nms switches from the coordinate trick to a class-by-class loop above 4000 numbers on the
    # CPU (1000 boxes) and 20000 on CUDA (5000 boxes).  The reference's box branch runs on the GPU, so the CUDA rule is
    # the default; the parity tests against the CPU oracle set 1000 (between the two thresholds the two arithmetics can
    # differ in the last bit of a near-threshold IoU).
    COORD_TRICK_MAX = 5000

    def __init__(self, pooler, box_head: FastRCNNConvFCHead, predictor: FastRCNNOutputs):
        self.pooler, self.box_head, self.predictor = pooler, box_head, predictor
        self._key = None
        self._fc: List[Tuple[torch.Tensor, torch.Tensor]] = []
        self._pred: Optional[Tuple[torch.Tensor, torch.Tensor]] = None
        self.last_logits = self.last_deltas = None

    def _prepare(self, device: torch.device) -> None:
        params = list(self.box_head.parameters()) + list(self.predictor.parameters())
        key = tuple((p.data_ptr(), p._version) for p in params)
        if key == self._key:
            return
        self._fc = []
        for m in self.box_head:
            if isinstance(m, nn.Linear):
                if m.weight.device != device:
                    raise ops.LiftError("box_head lives on %s but the features are on %s" % (m.weight.device, device))
                K = m.weight.shape[1]
                self._fc.append((ops.cast_bf16(m.weight.detach().float().contiguous(), (K + 7) // 8 * 8)[:, :K],
                                 m.bias.detach().float().contiguous()))
        p = self.predictor
        w = torch.cat([p.cls_score.weight.detach(), p.bbox_pred.weight.detach()]).float().contiguous()
        b = torch.cat([p.cls_score.bias.detach(), p.bbox_pred.bias.detach()]).float().contiguous()
        K = w.shape[1]
        self._pred = (ops.cast_bf16(w, (K + 7) // 8 * 8)[:, :K], b)
        self._key = key

    def predictor_outputs(self, features: List[torch.Tensor], proposal_boxes: List[torch.Tensor]):
        """(scores [R, K+1], proposal_deltas [R, 4K or 4]) -- `box_predictor(box_head(box_pooler(...)))`, fp32 views of
        one [R, ld] buffer."""
        device = features[0].device
        self._prepare(device)
        boxes = torch.cat(proposal_boxes).to(device=device, dtype=torch.float32).contiguous()
        counts = [len(b) for b in proposal_boxes]
        batch_idx = torch.repeat_interleave(torch.arange(len(counts), device=device, dtype=torch.int32),
                                            torch.tensor(counts, device=device))
        x, _ = ops.roi_align(features, self.pooler.scales, boxes, batch_idx, self.pooler.output_size,
                             self.pooler.sampling_ratio, out_dtype=torch.bfloat16)
        x = x.flatten(1)                                  # (c, ph, pw)-major columns = nn.Flatten (roi_heads.py:275)
        for w, b in self._fc:
            # no bias / ReLU in the GEMM call lets AUTO cut K across the SMs for the long-K first layer
            acc, _ = ops.gemm_bf16(x, w, bias=None, relu=False, out_dtype=torch.float32)
            x = ops.bias_act(acc, b, True, torch.bfloat16)
        w, b = self._pred
        out, _ = ops.gemm_bf16(x, w, bias=b, relu=False, out_dtype=torch.float32, mode=ops.GEMM_DIRECT)
        K1 = self.predictor.num_classes + 1
        return out[:, :K1], out[:, K1:], boxes, counts

    def forward_box(self, features: List[torch.Tensor], proposals: List[Instances]) -> List[Instances]:
        """roi_heads.py:273-276,294-296 + fast_rcnn.py:121-143."""
        p = self.predictor
        device = features[0].device
        if sum(len(x) for x in proposals) == 0:
            res = []
            for x in proposals:
                inst = Instances(x.image_size)
                inst.pred_boxes = Boxes(torch.zeros((0, 4), device=device))
                inst.scores = torch.zeros((0,), device=device)
                inst.scores_full = torch.zeros((0, p.num_classes), device=device)
                inst.pred_classes = torch.zeros((0,), dtype=torch.int64, device=device)
                res.append(inst)
            self.last_index = [torch.zeros((0,), dtype=torch.int64, device=device) for _ in proposals]
            return res
        logits, deltas, boxes, counts = self.predictor_outputs(features, [x.proposal_boxes.tensor for x in proposals])
        self.last_logits, self.last_deltas = logits, deltas
        return self.inference(logits, deltas, boxes, counts, [x.image_size for x in proposals])

    def inference(self, logits, deltas, boxes, counts, image_sizes) -> List[Instances]:
        p = self.predictor
        device = boxes.device
        off = torch.tensor([0] + list(torch.tensor(counts).cumsum(0).tolist()), dtype=torch.int32, device=device)
        sizes = torch.tensor([list(s) for s in image_sizes], dtype=torch.int32, device=device)
        ob, osc, oc, oi, ofull, cnt = ops.box_inference(
            logits, deltas, boxes, off, sizes, max(counts), p.num_classes, p.box2box_weights, p.test_score_thresh,
            p.test_nms_thresh, p.test_topk_per_image, p.cls_agnostic_bbox_reg, self.COORD_TRICK_MAX,
            want_scores_full=True, scale_clamp=p.scale_clamp)
        res, kept = [], []
        for i, k in enumerate(cnt.tolist()):           # the one host read of the branch
            inst = Instances(tuple(image_sizes[i]))
            inst.pred_boxes = Boxes(ob[i, :k])
            inst.scores = osc[i, :k]
            inst.scores_full = ofull[i, :k]
            inst.pred_classes = oc[i, :k]
            res.append(inst)
            kept.append(oi[i, :k].long())
        self.last_index = kept
        return res


def build_box_branch(cfg, input_shape: Dict[str, ShapeSpec], pooler_cls) -> Dict[str, nn.Module]:
    """d2 `StandardROIHeads._init_box_head` for the shipped configuration (Base.yaml:60-70; module tree
    nohup.out:645-661): returns {"box_pooler", "box_head", "box_predictor"}."""
    in_features = cfg.MODEL.ROI_HEADS.IN_FEATURES
    bh = cfg.MODEL.ROI_BOX_HEAD
    res = bh.POOLER_RESOLUTION
    scales = tuple(1.0 / input_shape[k].stride for k in in_features)
    in_channels = [input_shape[f].channels for f in in_features]
    assert len(set(in_channels)) == 1, in_channels
    pooler = pooler_cls(output_size=res, scales=scales, sampling_ratio=bh.POOLER_SAMPLING_RATIO,
                        pooler_type=bh.POOLER_TYPE)
    if bh.get("NUM_CONV", 0) != 0:
        raise ops.LiftError("ROI_BOX_HEAD.NUM_CONV > 0 is not part of the shipped configurations")
    head = FastRCNNConvFCHead(ShapeSpec(channels=in_channels[0], height=res, width=res), bh.NUM_FC, bh.FC_DIM)
    predictor = FastRCNNOutputs(head.output_dim, cfg.MODEL.ROI_HEADS.NUM_CLASSES, bh.CLS_AGNOSTIC_BBOX_REG,
                                bh.BBOX_REG_WEIGHTS, cfg.MODEL.ROI_HEADS.SCORE_THRESH_TEST,
                                cfg.MODEL.ROI_HEADS.NMS_THRESH_TEST, cfg.TEST.DETECTIONS_PER_IMAGE)
    return {"box_pooler": pooler, "box_head": head, "box_predictor": predictor}
