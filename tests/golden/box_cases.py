"""Seeded inputs for the 2-D box head's inference tail, shared by `make_golden.py` (which runs
the reference's own `fast_rcnn_inference`, fast_rcnn.py:16-116) and the tests."""
from __future__ import annotations

import hashlib
import os
from dataclasses import dataclass
from typing import List, Tuple

import torch

HERE = os.path.dirname(os.path.abspath(__file__))


@dataclass(frozen=True)
class BoxCase:
    name: str
    seed: int
    proposals_per_image: Tuple[int, ...]
    num_classes: int
    image_sizes: Tuple[Tuple[int, int], ...]
    cls_agnostic: bool = False
    score_thresh: float = 0.01            # Base.yaml:65
    nms_thresh: float = 0.5               # d2 default MODEL.ROI_HEADS.NMS_THRESH_TEST
    topk: int = 100                       # config.py:220
    logit_scale: float = 3.0
    nonfinite_rows: int = 0


CASES = [
    # few candidates per image: torchvision's coordinate-trick path (<= 1000 candidates)
    BoxCase("box_inference_small", 21, (60, 0, 35, 1), 50, ((512, 683), (512, 512), (384, 512), (600, 800)),
            logit_scale=6.0),
    # shipped sizes: 1000 proposals, 50 classes -> thousands of candidates: torchvision's class-by-class path
    BoxCase("box_inference_1000x50", 22, (1000, 777), 50, ((896, 1184), (896, 896)), logit_scale=4.0),
    BoxCase("box_inference_agnostic", 23, (300, 120), 7, ((512, 640), (480, 640)), cls_agnostic=True,
            score_thresh=0.05, nms_thresh=0.3, topk=40, logit_scale=4.0),
    BoxCase("box_inference_nonfinite_keepall", 24, (80, 50), 5, ((512, 512), (512, 512)), topk=-1, nonfinite_rows=4,
            logit_scale=5.0),
]
CASE_BY_NAME = {c.name: c for c in CASES}
REG_WEIGHTS = (10.0, 10.0, 5.0, 5.0)      # d2 default MODEL.ROI_BOX_HEAD.BBOX_REG_WEIGHTS


def golden_path(case: BoxCase) -> str:
    return os.path.join(HERE, case.name + ".npz")


def build_case(case: BoxCase):
    """-> (logits [R, K+1], deltas [R, 4K or 4], proposals: list of [Ri, 4])."""
    g = torch.Generator().manual_seed(case.seed)
    K = case.num_classes
    props: List[torch.Tensor] = []
    for n, (h, w) in zip(case.proposals_per_image, case.image_sizes):
        clusters = max(1, n // 10)
        cen = torch.rand(clusters, 2, generator=g) * torch.tensor([w, h], dtype=torch.float32)
        size = torch.exp(torch.rand(clusters, 2, generator=g) * 2.5 + 2.5)
        which = torch.randint(0, clusters, (n,), generator=g)
        c = cen[which] + torch.randn(n, 2, generator=g) * 8.0
        s = size[which] * torch.exp(torch.randn(n, 2, generator=g) * 0.2)
        b = torch.cat([c - s / 2, c + s / 2], 1)
        b[:, 0::2] = b[:, 0::2].clamp(0, w)
        b[:, 1::2] = b[:, 1::2].clamp(0, h)
        props.append(b.contiguous())
    R = sum(case.proposals_per_image)
    logits = torch.randn(R, K + 1, generator=g) * case.logit_scale
    deltas = torch.randn(R, 4 if case.cls_agnostic else 4 * K, generator=g) * 1.5
    deltas[::17, 2::4] = 40.0             # exercises the scale clamp
    if R > 6:
        logits[5] = logits[3]             # exact score ties between two proposals
    for i in range(case.nonfinite_rows):
        r = 7 + 11 * i
        if i % 2 == 0:
            logits[r, i % (K + 1)] = float("nan")
        else:
            deltas[r, 2] = float("inf")
    return logits, deltas, props


def input_checksum(logits, deltas, props) -> str:
    h = hashlib.sha256()
    for t in [logits, deltas] + list(props):
        h.update(t.numpy().tobytes())
    return h.hexdigest()
