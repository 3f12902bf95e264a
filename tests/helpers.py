"""Shared test helpers: run the oracle on a golden case, tolerance metrics."""
from __future__ import annotations

import numpy as np
import torch

from oracle import lift_oracle as lo
from tests.golden import cases as gc


def lift_config(case: gc.Case) -> lo.LiftConfig:
    cube = case.cube
    return lo.LiftConfig(
        scales=tuple(1.0 / s for s in case.strides),
        pooler_resolution=7,
        sampling_ratio=cube.get("POOLER_SAMPLING_RATIO", 0),
        num_classes=case.num_classes,
        z_type=cube.get("Z_TYPE", "direct"),
        pose_type=cube.get("POSE_TYPE", "6d"),
        cluster_bins=cube.get("CLUSTER_BINS", 1),
        shared_fc=cube.get("SHARED_FC", True),
        dims_priors_enabled=cube.get("DIMS_PRIORS_ENABLED", True),
        dims_priors_func=cube.get("DIMS_PRIORS_FUNC", "exp"),
        use_confidence=cube.get("USE_CONFIDENCE", 1.0),
        virtual_depth=cube.get("VIRTUAL_DEPTH", True),
        virtual_focal=cube.get("VIRTUAL_FOCAL", 512.0),
        allocentric_pose=cube.get("ALLOCENTRIC_POSE", True),
        scale_roi_boxes=cube.get("SCALE_ROI_BOXES", 0.0),
    )


def round_tf32(x: torch.Tensor) -> torch.Tensor:
    """cvt.rna.tf32.f32 (nearest, ties away from zero, 10 mantissa bits) on an fp32 tensor."""
    i = x.contiguous().view(torch.int32)
    return ((i + 0x1000) & ~0x1FFF).view(torch.float32)


def round_f16(x: torch.Tensor) -> torch.Tensor:
    """fp32 -> fp16 -> fp32 as the kernels do it: round to nearest even, saturated to +-65504 (NaN stays NaN)."""
    return torch.where(x == x, x.clamp(-65504.0, 65504.0), x).half().float()


def tail_format(case: gc.Case) -> str:
    """Operand format of this case's last FC layer in the product: "f16" / "tf32" when it runs fused with the heads and
    the decode (`ovm_fc_tail`, csrc/fc_tail.cu: shared FC stack, class-agnostic heads, FC_DIM a power of two in
    128..1024; the format is OVM_TAIL_PRECISION, default f16), else "bf16"."""
    import os

    cube = case.cube
    fused = (cube.get("SHARED_FC", True) and not cube.get("DIMS_PRIORS_ENABLED", True) and cube.get("CLUSTER_BINS", 1) <= 1
             and case.fc_dim in (128, 256, 512, 1024) and os.environ.get("OVM_FUSE_TAIL", "1") != "0")
    return os.environ.get("OVM_TAIL_PRECISION", "f16").lower() if fused else "bf16"


def fc_stack_product_operands(tail: str):
    """`lift_oracle._fc_stack` with the operands of every Linear rounded the way the product rounds them -- bf16
    (round-to-nearest-even) everywhere, except fp16 (or TF32) for the last layer when it runs inside `ovm_fc_tail` -- and the
    products summed exactly (fp64), which is what the tcgen05 GEMMs compute up to fp32 summation order: the
    kernel-correctness oracle for the head, separate from the error of the number formats themselves."""
    import torch.nn.functional as F

    def stack(w, prefix, x):
        num = 0
        while f"{prefix}.fc{num + 1}.weight" in w:
            num += 1
        for k in range(1, num + 1):
            fmt = tail if (k == num and num >= 2) else "bf16"
            rnd = {"tf32": round_tf32, "f16": round_f16, "bf16": lambda t: t.bfloat16().float()}[fmt]
            a = rnd(x).double()
            wt = rnd(w[f"{prefix}.fc{k}.weight"]).double()
            x = F.relu(F.linear(a, wt, w[f"{prefix}.fc{k}.bias"].double())).float()
        return x

    return stack


def run_oracle(case: gc.Case, bf16_operands: bool = False):
    if bf16_operands:
        saved = lo._fc_stack
        lo._fc_stack = fc_stack_product_operands(tail_format(case))
        try:
            return run_oracle(case)
        finally:
            lo._fc_stack = saved
    inp, weights, priors = gc.build_case(case)
    cfg = lift_config(case)
    kw = {}
    if priors is not None:
        kw = dict(priors_dims_per_cat=priors[0], priors_z_scales=priors[1], priors_z_stats=priors[2])
    with torch.no_grad():
        out = lo.forward_cube(inp.features, inp.boxes, inp.classes, inp.scores if case.with_scores else None,
                              inp.Ks, inp.im_dims, inp.im_scales_ratio, weights, cfg, **kw)
    return inp, weights, priors, cfg, out


def rel_err(a, b, floor: float = 1e-6) -> float:
    """max |a-b| / (|b| + floor*max|b|): element-wise relative error with a scale-aware floor."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    if b.size == 0:
        return 0.0
    scale = max(float(np.abs(b).max()), 1e-30)
    return float((np.abs(a - b) / (np.abs(b) + floor * scale)).max())


def norm_err(a, b) -> float:
    """max |a-b| / max |b|."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    if b.size == 0:
        return 0.0
    return float(np.abs(a - b).max() / max(float(np.abs(b).max()), 1e-30))
