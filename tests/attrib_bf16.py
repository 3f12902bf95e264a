"""Where the bf16 head's pose / corner error comes from ("What's weak" 1).

Runs the oracle (fp32 CPU restatement of the reference's lift) on the full C2 batch of bench.py with round-to-nearest
bf16 applied to ONE operand set of the FC stack at a time -- pooled (A of fc1), W1, h1 (A of fc2), W2 -- and to the
combinations the kernels can run, and reports for every record field the error against the all-fp32 oracle:
max-norm relative to the field's scale, relative L2, and for the pose the per-box geodesic angle.

    python -m tests.attrib_bf16 [--seed 1235] [--out profiles/r02_bf16_attribution.json]

Lives under tests/ because it executes oracle/.  CPU only; a C2 batch takes about a second per variant.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import lift_oracle as lo  # noqa: E402

FIELDS = ("scores", "pred_bbox3D", "pred_center_cam", "pred_center_2D", "pred_dimensions", "pred_pose")


def _round(x: torch.Tensor, how: str) -> torch.Tensor:
    if how == "f32":
        return x
    if how == "bf16":
        return x.bfloat16().float()
    if how == "tf32":      # 10 explicit mantissa bits, round to nearest even (cvt.rna.tf32.f32 rounds half away; same size)
        i = x.contiguous().view(torch.int32)
        r = ((i + 0x0FFF + ((i >> 13) & 1)) & ~0x1FFF)
        return r.view(torch.float32)
    if how == "f16":       # 10 explicit mantissa bits like TF32, 5 exponent bits: saturating round-to-nearest-even
        return x.clamp(-65504.0, 65504.0).half().float()
    if how == "bf16x2":    # hi + lo bf16 pair: what a 3-term split product carries (16 mantissa bits)
        hi = x.bfloat16().float()
        return hi + (x - hi).bfloat16().float()
    raise ValueError(how)


def fc_stack_with(rounding):
    """rounding = {(layer k, 'a'|'w'): how}; products summed in fp64 (the tensor core's fp32 accumulation order is
    not what is being studied here)."""

    def stack(w, prefix, x):
        k = 1
        while f"{prefix}.fc{k}.weight" in w:
            a = _round(x, rounding.get((k, "a"), "f32")).double()
            wt = _round(w[f"{prefix}.fc{k}.weight"], rounding.get((k, "w"), "f32")).double()
            x = F.relu(F.linear(a, wt, w[f"{prefix}.fc{k}.bias"].double())).float()
            k += 1
        return x

    return stack


def geodesic_deg(Ra: torch.Tensor, Rb: torch.Tensor) -> torch.Tensor:
    M = Ra.double() @ Rb.double().transpose(1, 2)
    c = ((M[:, 0, 0] + M[:, 1, 1] + M[:, 2, 2]) - 1) / 2
    return torch.rad2deg(torch.acos(c.clamp(-1, 1)))


def run(inp, weights, cfg, rounding):
    saved = lo._fc_stack
    lo._fc_stack = fc_stack_with(rounding)
    try:
        with torch.no_grad():
            return lo.forward_cube(inp.features, inp.boxes, inp.classes, inp.scores, inp.Ks, inp.im_dims,
                                   inp.im_scales_ratio, weights, cfg)
    finally:
        lo._fc_stack = saved


def field_errors(out, ref):
    res = {}
    for f in FIELDS:
        a = torch.cat(getattr(out, f)).double()
        b = torch.cat(getattr(ref, f)).double()
        scale = b.abs().max().clamp_min(1e-30)
        # per image, like the tests: max over images of the per-image numbers
        mx, l2 = 0.0, 0.0
        for x, y in zip(getattr(out, f), getattr(ref, f)):
            x, y = x.double(), y.double()
            mx = max(mx, float((x - y).abs().max() / y.abs().max().clamp_min(1e-30)))
            l2 = max(l2, float((x - y).norm() / y.norm().clamp_min(1e-30)))
        res[f] = {"max_rel_per_image": mx, "l2_rel_per_image": l2,
                  "max_rel_batch": float((a - b).abs().max() / scale)}
    ang = geodesic_deg(torch.cat(out.pred_pose), torch.cat(ref.pred_pose))
    res["pose_geodesic_deg"] = {"max": float(ang.max()), "p99": float(ang.quantile(0.99)), "median": float(ang.median()),
                                "mean": float(ang.mean())}
    raw_a, raw_b = out.head_raw["pose"], ref.head_raw["pose"]   # rotation matrices before the allocentric step
    return res


VARIANTS = [
    ("pooled_bf16", {(1, "a"): "bf16"}),
    ("W1_bf16", {(1, "w"): "bf16"}),
    ("h1_bf16", {(2, "a"): "bf16"}),
    ("W2_bf16", {(2, "w"): "bf16"}),
    ("fc1_bf16 (pooled+W1)", {(1, "a"): "bf16", (1, "w"): "bf16"}),
    ("fc2_bf16 (h1+W2)", {(2, "a"): "bf16", (2, "w"): "bf16"}),
    ("all_bf16 (round 1 product)", {(1, "a"): "bf16", (1, "w"): "bf16", (2, "a"): "bf16", (2, "w"): "bf16"}),
    ("fc1_bf16 + fc2_tf32", {(1, "a"): "bf16", (1, "w"): "bf16", (2, "a"): "tf32", (2, "w"): "tf32"}),
    ("fc1_bf16 + fc2_f16 (this round's product)", {(1, "a"): "bf16", (1, "w"): "bf16", (2, "a"): "f16", (2, "w"): "f16"}),
    ("fc1_bf16 + fc2_bf16x2 split", {(1, "a"): "bf16", (1, "w"): "bf16", (2, "a"): "bf16x2", (2, "w"): "bf16x2"}),
    ("fc1 pooled_bf16x2 W1_bf16 + fc2 split", {(1, "a"): "bf16x2", (1, "w"): "bf16", (2, "a"): "bf16x2", (2, "w"): "bf16x2"}),
    ("fc1 split both + fc2 split", {(1, "a"): "bf16x2", (1, "w"): "bf16x2", (2, "a"): "bf16x2", (2, "w"): "bf16x2"}),
]


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--seed", type=int, default=1235)
    ap.add_argument("--out", default=None)
    args = ap.parse_args(argv)
    import bench

    torch.set_num_threads(os.cpu_count() or 1)
    inp = bench.make_batch(args.seed)
    weights = bench.make_weights()
    cfg = lo.LiftConfig(scales=tuple(1.0 / s for s in bench.STRIDES), dims_priors_enabled=False)
    ref = run(inp, weights, cfg, {})
    report = {"workload": bench.WORKLOAD, "seed": args.seed, "boxes": int(sum(len(b) for b in inp.boxes)),
              "reference": "oracle with every operand fp32 (products summed in fp64)", "variants": {}}
    raw6 = None
    for name, rounding in VARIANTS:
        out = run(inp, weights, cfg, rounding)
        report["variants"][name] = field_errors(out, ref)
        e = report["variants"][name]
        print("%-52s pose max %.2e l2 %.2e | corners max %.2e l2 %.2e | centre %.2e dims %.2e | geodesic max %.3f deg "
              "p99 %.3f" % (name, e["pred_pose"]["max_rel_per_image"], e["pred_pose"]["l2_rel_per_image"],
                            e["pred_bbox3D"]["max_rel_per_image"], e["pred_bbox3D"]["l2_rel_per_image"],
                            e["pred_center_cam"]["max_rel_per_image"], e["pred_dimensions"]["max_rel_per_image"],
                            e["pose_geodesic_deg"]["max"], e["pose_geodesic_deg"]["p99"]))
    if args.out:
        with open(args.out, "w") as f:
            json.dump(report, f, indent=1)
    return report


if __name__ == "__main__":
    main()
