"""Golden-vector case definitions shared by `make_golden.py` (reference side) and the tests.

Inputs are regenerated from seeds (`ovmono3d_b200.synth`), so the committed `.npz` files hold only
the reference's outputs plus an input checksum that detects RNG drift.
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional, Sequence

import numpy as np
import torch

from ovmono3d_b200 import synth

GOLDEN_DIR = os.path.dirname(os.path.abspath(__file__))


@dataclass
class Case:
    name: str
    seed: int
    strides: Sequence[int] = (7, 14, 28)
    canvas: int = 896
    content: Optional[int] = None
    boxes_per_image: Sequence[int] = (12, 12)
    channels: int = 16
    fc_dim: int = 64
    num_classes: int = 50
    with_scores: bool = True
    head_std: float = 0.05
    bias_std: float = 0.0
    cube: Dict[str, Any] = field(default_factory=dict)     # MODEL.ROI_CUBE_HEAD overrides
    mutate: Optional[str] = None                            # name of an edge-case mutator below
    pooled_stride: int = 1                                  # subsample step for the stored pooled tensor
    use_priors: bool = False

    @property
    def in_features(self) -> List[str]:
        return ["p%d" % (i + 2) for i in range(len(self.strides))]


SHIPPED = dict(DIMS_PRIORS_ENABLED=False)   # Base.yaml:81 -- the class-agnostic head the repo ships

CASES: List[Case] = [
    Case("dinov2_agnostic", 11, cube=dict(SHIPPED)),
    Case("dinov2_no_scores", 12, cube=dict(SHIPPED), with_scores=False),
    Case("priors_exp_zsigmoid", 13, cube=dict(DIMS_PRIORS_ENABLED=True, DIMS_PRIORS_FUNC="exp", Z_TYPE="sigmoid"),
         num_classes=7, use_priors=True),
    Case("priors_sigmoid_clusters_quat", 14, num_classes=5, use_priors=True,
         cube=dict(DIMS_PRIORS_ENABLED=True, DIMS_PRIORS_FUNC="sigmoid", Z_TYPE="clusters", CLUSTER_BINS=3,
                   POSE_TYPE="quaternion")),
    # (class-agnostic + CLUSTER_BINS>1 crashes inside the reference itself, roi_heads.py:449 -- not a case)
    Case("priors_bins_zlog", 15, num_classes=5, use_priors=True,
         cube=dict(DIMS_PRIORS_ENABLED=True, CLUSTER_BINS=4, Z_TYPE="log")),
    Case("euler_log_ego_real", 16, cube=dict(DIMS_PRIORS_ENABLED=False, POSE_TYPE="euler", Z_TYPE="log",
                                             ALLOCENTRIC_POSE=False, VIRTUAL_DEPTH=False)),
    Case("no_confidence", 17, cube=dict(DIMS_PRIORS_ENABLED=False, USE_CONFIDENCE=0.0)),
    Case("clip_4level", 18, strides=(4, 8, 16, 32), canvas=512, boxes_per_image=(20, 3, 9), cube=dict(SHIPPED)),
    Case("dla_5level", 19, strides=(4, 8, 16, 32, 64), canvas=1024, channels=8, boxes_per_image=(25,),
         cube=dict(SHIPPED)),
    Case("single_level", 20, strides=(16,), canvas=256, boxes_per_image=(6, 6), cube=dict(SHIPPED)),
    Case("non_shared_fc", 21, cube=dict(DIMS_PRIORS_ENABLED=False, SHARED_FC=False)),
    Case("scale_roi_boxes", 22, cube=dict(DIMS_PRIORS_ENABLED=False, SCALE_ROI_BOXES=1.5)),
    Case("sampling_ratio_2", 23, cube=dict(DIMS_PRIORS_ENABLED=False, POOLER_SAMPLING_RATIO=2)),
    Case("one_box_total", 24, boxes_per_image=(0, 1), cube=dict(SHIPPED)),
    Case("ragged_with_empty", 25, boxes_per_image=(5, 0, 9, 1), cube=dict(SHIPPED)),
    Case("edge_boxes", 26, boxes_per_image=(16, 16), cube=dict(SHIPPED), mutate="edge_boxes"),
    Case("edge_clips", 27, boxes_per_image=(10, 10), cube=dict(SHIPPED), mutate="edge_clips", head_std=0.2,
         bias_std=2.0),
    Case("level_thresholds", 28, boxes_per_image=(40, 40), cube=dict(SHIPPED), mutate="level_thresholds"),
    Case("reference_init", 29, cube=dict(SHIPPED), head_std=0.001),
    # C1 of BASELINE.json: 4 synthetic 512px images x 20 boxes, full-size head
    Case("c1_full", 1234, content=512, boxes_per_image=(20, 20, 20, 20), channels=256, fc_dim=1024,
         cube=dict(SHIPPED), pooled_stride=61),
]
CASE_BY_NAME = {c.name: c for c in CASES}


# --------------------------------------------------------------------------- edge-case mutators
def _mut_edge_boxes(inp: synth.LiftInputs, weights, case: Case) -> None:
    """Boxes partly / fully outside the map, zero-area, inverted, tiny, whole-image."""
    W = float(case.canvas)
    b = inp.boxes[0]
    b[0] = torch.tensor([-50.0, -30.0, 40.0, 60.0])            # hangs over the top-left corner
    b[1] = torch.tensor([W - 20.0, W - 35.0, W + 90.0, W + 10.0])  # hangs over the bottom-right corner
    b[2] = torch.tensor([W + 30.0, W + 30.0, W + 90.0, W + 80.0])  # fully outside
    b[3] = torch.tensor([-200.0, 100.0, -120.0, 180.0])        # fully outside (left)
    b[4] = torch.tensor([100.0, 100.0, 100.0, 100.0])          # zero area
    b[5] = torch.tensor([300.0, 200.0, 300.0, 400.0])          # zero width
    b[6] = torch.tensor([400.0, 400.0, 380.0, 390.0])          # inverted
    b[7] = torch.tensor([10.25, 10.5, 11.0, 11.75])            # sub-pixel on the finest level
    b[8] = torch.tensor([0.0, 0.0, W, W])                      # whole canvas
    b[9] = torch.tensor([0.0, 300.0, W, 330.0])                # very wide, thin
    b[10] = torch.tensor([500.0, 0.0, 520.0, W])               # very tall, thin
    b[11] = torch.tensor([W - 1.0, W - 1.0, W, W])             # last pixel
    # put one box centre exactly on the scaled principal point: the angle == 0 lane of R_from_allocentric
    K = inp.Ks[1] / inp.im_scales_ratio[1]
    cx, cy = float(K[0, 2]), float(K[1, 2])
    inp.boxes[1][0] = torch.tensor([cx - 32.0, cy - 48.0, cx + 32.0, cy + 48.0])
    for k in ("bbox_3D_center_deltas.weight", "bbox_3D_center_deltas.bias"):
        weights[k].zero_()


def _mut_edge_clips(inp: synth.LiftInputs, weights, case: Case) -> None:
    """Large biases so dims_raw > 5 (clip) and uncertainty < 0.01 (clip) both occur."""
    weights["bbox_3D_dims.bias"] += torch.tensor([6.0, -1.0, 4.9])
    weights["bbox_3D_uncertainty.bias"].fill_(-1.0)


def level_threshold_boxes(strides: Sequence[int], count: int, gen: torch.Generator) -> torch.Tensor:
    """Square boxes whose side sits within +-8 ulp of 224 * 2^(k-4) for every level edge k."""
    min_level = int(math.log2(strides[0]))
    max_level = int(math.log2(strides[-1]))
    sides = []
    for k in range(min_level, max_level + 2):
        base = np.float32(224.0 * 2.0 ** (k - 4))
        s = base
        lo = [base]
        for _ in range(8):
            s = np.nextafter(s, np.float32(0), dtype=np.float32)
            lo.append(s)
        s = base
        for _ in range(8):
            s = np.nextafter(s, np.float32(1e9), dtype=np.float32)
            lo.append(s)
        sides.extend(lo)
    sides = torch.tensor(np.array(sides, dtype=np.float32))
    idx = torch.randint(0, len(sides), (count,), generator=gen)
    side = sides[idx]
    x0 = torch.rand(count, generator=gen) * 64.0
    y0 = torch.rand(count, generator=gen) * 64.0
    # (x0 + side) - x0 must reproduce `side`'s neighbourhood; use x0 = y0 = 0 for half of them (exact)
    x0[::2] = 0.0
    y0[::2] = 0.0
    return torch.stack([x0, y0, x0 + side, y0 + side], dim=1)


def _mut_level_thresholds(inp: synth.LiftInputs, weights, case: Case) -> None:
    gen = torch.Generator().manual_seed(case.seed + 1000)
    for i in range(len(inp.boxes)):
        inp.boxes[i] = level_threshold_boxes(case.strides, len(inp.boxes[i]), gen)


MUTATORS = dict(edge_boxes=_mut_edge_boxes, edge_clips=_mut_edge_clips, level_thresholds=_mut_level_thresholds)


# --------------------------------------------------------------------------- builders
def build_case(case: Case):
    """-> (inputs, head weights, priors tuple or None).  Deterministic in `case.seed`."""
    cube = case.cube
    inp = synth.make_inputs(case.seed, len(case.boxes_per_image), list(case.boxes_per_image),
                            strides=case.strides, canvas=case.canvas, content=case.content,
                            channels=case.channels, num_classes=case.num_classes)
    weights = synth.make_head_weights(
        case.seed + 7, in_dim=case.channels * 49, fc_dim=case.fc_dim, num_classes=case.num_classes,
        dims_priors_enabled=cube.get("DIMS_PRIORS_ENABLED", True), cluster_bins=cube.get("CLUSTER_BINS", 1),
        pose_type=cube.get("POSE_TYPE", "6d"), use_confidence=bool(cube.get("USE_CONFIDENCE", 1.0)),
        shared_fc=cube.get("SHARED_FC", True), head_std=case.head_std, bias_std=case.bias_std)
    priors = synth.make_priors(case.seed + 13, case.num_classes, cube.get("CLUSTER_BINS", 1)) if case.use_priors else None
    if case.mutate:
        MUTATORS[case.mutate](inp, weights, case)
    return inp, weights, priors


def make_cfg(case: Case):
    from ovmono3d_b200.config import get_cfg

    cfg = get_cfg()
    cfg.MODEL.ROI_HEADS.IN_FEATURES = case.in_features
    cfg.MODEL.ROI_HEADS.NUM_CLASSES = case.num_classes
    cfg.MODEL.ROI_CUBE_HEAD.FC_DIM = case.fc_dim
    for k, v in case.cube.items():
        cfg.MODEL.ROI_CUBE_HEAD[k] = v
    return cfg


def input_checksum(inp: synth.LiftInputs, weights) -> np.ndarray:
    parts = [float(f.double().sum()) for f in inp.features]
    parts += [float(torch.cat(inp.boxes).double().sum()) if sum(map(len, inp.boxes)) else 0.0]
    parts += [float(sum(w.double().sum() for w in weights.values()))]
    return np.array(parts, dtype=np.float64)


def golden_path(case: Case) -> str:
    return os.path.join(GOLDEN_DIR, case.name + ".npz")
