"""Seeded synthetic inputs for the lift path (SURVEY.md §8d "Synthetic inputs").

Used by the tests, the golden-vector script and `bench.py`; everything is generated on the CPU
with a `torch.Generator` so that the oracle and the CUDA path see bit-identical inputs.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import torch


@dataclass
class LiftInputs:
    features: List[torch.Tensor]          # per level [N, C, H_l, W_l] fp32 NCHW
    strides: List[int]
    boxes: List[torch.Tensor]             # per image [n_i, 4] xyxy, network-resolution px
    classes: List[torch.Tensor]           # per image [n_i] int64
    scores: List[torch.Tensor]            # per image [n_i] fp32
    Ks: List[torch.Tensor]                # per image [3, 3] fp32 (original-resolution intrinsics)
    im_scales_ratio: List[float]
    im_dims: List[Tuple[int, int]]        # per image (H_cur, W_cur) of the padded network input


def make_boxes(gen: torch.Generator, n: int, content_w: float, content_h: float,
               min_side: float = 16.0) -> torch.Tensor:
    """Log-uniform box sides so every pyramid level is populated; clipped to the content area."""
    ctr = torch.rand(n, 2, generator=gen) * torch.tensor([content_w, content_h])
    lo, hi = math.log(min_side), math.log(0.9 * min(content_w, content_h))
    side = torch.exp(torch.rand(n, generator=gen) * (hi - lo) + lo)
    aspect = torch.exp(torch.rand(n, generator=gen) * 1.4 - 0.7)
    w = side * torch.sqrt(aspect)
    h = side / torch.sqrt(aspect)
    b = torch.stack([ctr[:, 0] - w / 2, ctr[:, 1] - h / 2, ctr[:, 0] + w / 2, ctr[:, 1] + h / 2], dim=1)
    b[:, 0::2] = b[:, 0::2].clamp(0, content_w)
    b[:, 1::2] = b[:, 1::2].clamp(0, content_h)
    return b.float()


def make_inputs(
    seed: int,
    num_images: int,
    boxes_per_image: int | Sequence[int],
    strides: Sequence[int] = (7, 14, 28),
    canvas: int = 896,
    content: Optional[int] = None,
    channels: int = 256,
    num_classes: int = 50,
) -> LiftInputs:
    gen = torch.Generator().manual_seed(seed)
    content = content or canvas
    feats = [torch.randn(num_images, channels, canvas // s, canvas // s, generator=gen) for s in strides]
    if isinstance(boxes_per_image, int):
        boxes_per_image = [boxes_per_image] * num_images
    boxes, classes, scores, Ks, ratios, dims = [], [], [], [], [], []
    for i in range(num_images):
        n_i = int(boxes_per_image[i])
        boxes.append(make_boxes(gen, n_i, float(content), float(content)))
        classes.append(torch.randint(0, num_classes, (n_i,), generator=gen))
        scores.append(torch.rand(n_i, generator=gen) * 0.95 + 0.05)
        ratio = float(torch.rand(1, generator=gen) * 1.5 + 0.5)
        h_orig = content * ratio
        f = float(torch.rand(1, generator=gen) * 1.2 + 0.8) * h_orig
        pp = torch.rand(2, generator=gen) * 40 - 20
        K = torch.tensor([[f, 0.0, h_orig / 2 + float(pp[0])],
                          [0.0, f, h_orig / 2 + float(pp[1])],
                          [0.0, 0.0, 1.0]])
        Ks.append(K.float())
        ratios.append(ratio)
        dims.append((canvas, canvas))
    return LiftInputs(feats, list(strides), boxes, classes, scores, Ks, ratios, dims)


def _xavier(gen: torch.Generator, out_f: int, in_f: int) -> torch.Tensor:
    # fvcore c2_xavier_fill == kaiming_uniform_(a=1): U(-sqrt(3/fan_in), sqrt(3/fan_in))  (cube_head.py:72)
    bound = math.sqrt(3.0 / in_f)
    return (torch.rand(out_f, in_f, generator=gen) * 2 - 1) * bound


def make_head_weights(
    seed: int,
    in_dim: int = 256 * 49,
    fc_dim: int = 1024,
    num_fc: int = 2,
    num_classes: int = 50,
    dims_priors_enabled: bool = False,
    cluster_bins: int = 1,
    pose_type: str = "6d",
    use_confidence: bool = True,
    shared_fc: bool = True,
    head_std: float = 0.05,
    bias_std: float = 0.0,
) -> Dict[str, torch.Tensor]:
    """Random-init CubeHead state dict with the reference's key names (cube_head.py:64-145).

    `head_std=0.001` is the reference's init (`cube_head.py:110-145`); the default 0.05 is the
    "trained-scale" variant of SURVEY.md §8d so that decoded outputs are not degenerate.
    """
    gen = torch.Generator().manual_seed(seed)
    w: Dict[str, torch.Tensor] = {}
    gens = ["feature_generator"] if shared_fc else (
        ["feature_generator_dims", "feature_generator_XY", "feature_generator_pose", "feature_generator_Z"]
        + (["feature_generator_conf"] if use_confidence else []))
    for g in gens:
        d = in_dim
        for k in range(num_fc):
            w[f"{g}.fc{k + 1}.weight"] = _xavier(gen, fc_dim, d)
            w[f"{g}.fc{k + 1}.bias"] = torch.randn(fc_dim, generator=gen) * bias_std
            d = fc_dim
    mult = num_classes if dims_priors_enabled else 1
    pose_dim = {"6d": 6, "quaternion": 4, "euler": 3}[pose_type]
    bins = cluster_bins if cluster_bins > 1 else 1
    heads = [("bbox_3D_dims", 3 * mult, 0.0), ("bbox_3D_center_deltas", 2 * mult, 0.0),
             ("bbox_3D_pose", pose_dim * mult, 0.0), ("bbox_3D_center_depth", bins * mult, 0.0)]
    if use_confidence:
        heads.append(("bbox_3D_uncertainty", mult, 5.0))
    for name, o, b in heads:
        w[f"{name}.weight"] = torch.randn(o, fc_dim, generator=gen) * head_std
        w[f"{name}.bias"] = torch.full((o,), b) + torch.randn(o, generator=gen) * bias_std
    if head_std > 0.01:
        # trained-scale variant: centre depth around a few metres, uncertainty around 1
        w["bbox_3D_center_depth.bias"] += 3.0
        if use_confidence:
            w["bbox_3D_uncertainty.bias"] -= 4.0
    return w


def make_priors(seed: int, num_classes: int, cluster_bins: int):
    gen = torch.Generator().manual_seed(seed)
    mean = torch.rand(num_classes, 3, generator=gen) * 2 + 0.2
    std = torch.rand(num_classes, 3, generator=gen) * 0.3 + 0.05
    priors_dims_per_cat = torch.stack([mean, std], dim=1).unsqueeze(0)            # [1, C, 2, 3]
    bins = max(cluster_bins, 1)
    z_scales = torch.sort(torch.rand(num_classes, bins, generator=gen) * 600 + 20, dim=1).values
    z_stats = torch.stack([torch.rand(num_classes, bins, generator=gen) * 20 + 2,
                           torch.rand(num_classes, bins, generator=gen) * 3 + 0.5], dim=2)  # [C, bins, 2]
    return priors_dims_per_cat.float(), z_scales.float(), z_stats.float()
