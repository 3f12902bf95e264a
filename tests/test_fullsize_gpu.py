"""Full-size oracle parity on the BASELINE configs the bench times (VERDICT r1 "missing" 7): C2 (16 x 100 boxes,
DINOv2 SFP), C3 (1024-px 4-level SFP, 1 x 2000 and 8 x 2000 boxes) and the full-size C5 pyramids incl. C = 128 / 512.

Per workload, all through the C ABI and compared with `oracle.lift_oracle.forward_cube` on the same seeded inputs:
  * box-to-level assignment: bit-exact;
  * pooled features (fp32 `ovm_roi_align`): 1e-3 per box;
  * every record field of `ovm_lift_forward` against the oracle with the product's operand rounding (bf16 fc1,
    fp16 fc2 -- helpers.fc_stack_product_operands): 1e-3 (kernel correctness);
  * against the all-fp32 oracle (format error of the bf16 head): north_star's 1e-2 on 3-D centres, dimensions,
    2-D centres, scores and on the 13 raw head outputs; pose and corners -- a Gram-Schmidt of six of those raw numbers
    -- per box against the raw error times the conditioning of the box's 6-D vector, plus relative L2 and the
    geodesic-angle distribution (profiles/r02_bf16_attribution.json explains the outliers: short raw 6-D vectors).

`python -m tests.test_fullsize_gpu --format-error` prints the CPU emulation of the format error per workload
(no GPU needed): the bounds below come from that table.
"""
import sys

import numpy as np
import pytest
import torch

from oracle import lift_oracle as lo
from ovmono3d_b200 import workloads
from tests import helpers

FIELDS = ("scores", "pred_bbox3D", "pred_center_cam", "pred_center_2D", "pred_dimensions", "pred_pose")
DIRECT_FIELDS = ("scores", "pred_center_cam", "pred_center_2D", "pred_dimensions")   # no Gram-Schmidt on the way
KERNEL_TOL = 1e-3
FORMAT_TOL = 1e-2            # BASELINE.json north_star: "bf16 head within 1e-2"
# Pose and corners pass through rotation_6d_to_matrix: two normalisations that divide the error of the six raw head
# outputs by the lengths |a1| and |a2 - (b1.a2) b1|.  With random-init heads some of 16 000 boxes have a short leg, and
# the emulated format error (--format-error) reaches 0.25 in single pose entries (18 degrees) on such boxes while the
# raw outputs are within 4e-3 everywhere.  So: the raw outputs are held to 1e-2, the pose of every box to the raw
# error times the conditioning of ITS 6-D vector, and the distribution over boxes to the bounds below.
POSE_L2_TOL = 3e-2           # relative L2 of the pose matrices per image (emulation: 5e-3 .. 2.6e-2)
CORNER_L2_TOL = 1e-2         # relative L2 of the corners per image (emulation: 6e-4 .. 5.7e-3)
POSE_P99_DEG = 2.5           # 99th percentile of the per-box geodesic angle to the fp32 pose (emulation: 1.1 .. 1.8)
POSE_MEDIAN_DEG = 0.5        # (emulation: 0.15 .. 0.23)
COND_SAFETY = 6.0
NAMES = ["C2", "C3-1x2000", "C3-8x2000", "C5-sfp4", "C5-dla", "C5-c128", "C5-c512"]


def _cfg(w):
    return lo.LiftConfig(scales=tuple(1.0 / s for s in w.strides), dims_priors_enabled=False)


def _oracle(w, inp, weights, product_operands):
    """-> (LiftOutput, raw 6-D pose vectors [n, 6] as the oracle's heads emit them)."""
    saved, saved_r6 = lo._fc_stack, lo.tp.rotation_6d_to_matrix
    raw6 = []

    def capture(d6):
        raw6.append(d6.detach().clone())
        return saved_r6(d6)

    lo.tp.rotation_6d_to_matrix = capture
    if product_operands:
        lo._fc_stack = helpers.fc_stack_product_operands("f16")
    try:
        with torch.no_grad():
            out = lo.forward_cube(inp.features, inp.boxes, inp.classes, inp.scores, inp.Ks, inp.im_dims,
                                  inp.im_scales_ratio, weights, _cfg(w))
    finally:
        lo._fc_stack, lo.tp.rotation_6d_to_matrix = saved, saved_r6
    return out, raw6[0]


def _conditioning(raw6):
    """1 / (shorter leg of the Gram-Schmidt): how much rotation_6d_to_matrix amplifies an absolute error of its input."""
    a1, a2 = raw6[:, :3].double(), raw6[:, 3:].double()
    b1 = a1 / a1.norm(dim=1, keepdim=True).clamp_min(1e-12)
    b2 = a2 - (b1 * a2).sum(1, keepdim=True) * b1
    return 1.0 / torch.minimum(a1.norm(dim=1), b2.norm(dim=1)).clamp_min(1e-12)


def _per_image_err(got_list, want_list):
    mx, l2 = 0.0, 0.0
    for g, t in zip(got_list, want_list):
        g, t = g.double().cpu(), t.double()
        if t.numel() == 0:
            continue
        assert torch.isfinite(g).all()
        mx = max(mx, float((g - t).abs().max() / t.abs().max().clamp_min(1e-30)))
        l2 = max(l2, float((g - t).norm() / t.norm().clamp_min(1e-30)))
    return mx, l2


def _geodesic_deg(Ra, Rb):
    M = Ra.double() @ Rb.double().transpose(1, 2)
    c = ((M[:, 0, 0] + M[:, 1, 1] + M[:, 2, 2]) - 1) / 2
    return torch.rad2deg(torch.acos(c.clamp(-1, 1)))


def _raw_heads(out):
    """[n, 13] raw head outputs of an oracle run in the record's raw order: deltas 2, z 1, dims 3, uncertainty 1, and the
    rotation (3 x 3, before the allocentric step) flattened -- compared only through what the kernel also emits."""
    r = out.head_raw
    return torch.cat([r["deltas"], r["z"].reshape(-1, 1), r["dims"], r["uncert"].reshape(-1, 1)], dim=1)


def _split_records(rec, counts):
    out = {f: [] for f in FIELDS}
    start = 0
    for k in counts:
        r = rec[start:start + k]
        start += k
        out["scores"].append(r[:, 45])
        out["pred_bbox3D"].append(r[:, 0:24].reshape(k, 8, 3))
        out["pred_center_cam"].append(r[:, 24:27])
        out["pred_center_2D"].append(r[:, 27:29])
        out["pred_dimensions"].append(r[:, 29:32])
        out["pred_pose"].append(r[:, 32:41].reshape(k, 3, 3))
    return out


def format_error_table(names=NAMES):
    """CPU only: product-operand oracle against the fp32 oracle, per workload."""
    rows = {}
    for name in names:
        w = workloads.get(name)
        inp, weights = w.make_inputs(), w.make_weights()
        ref, _ = _oracle(w, inp, weights, False)
        emu, _ = _oracle(w, inp, weights, True)
        row = {}
        for f in FIELDS:
            row[f] = _per_image_err(getattr(emu, f), getattr(ref, f))
        ang = _geodesic_deg(torch.cat(emu.pred_pose), torch.cat(ref.pred_pose))
        row["geodesic_deg"] = dict(median=float(ang.median()), p99=float(ang.quantile(0.99)), max=float(ang.max()))
        ra, rb = _raw_heads(emu), _raw_heads(ref)
        row["raw_heads"] = float(((ra - rb).abs().amax(0) / rb.abs().amax(0).clamp_min(1e-30)).max())
        rows[name] = row
        print(name, {k: (tuple(round(x, 5) for x in v) if isinstance(v, tuple) else v) for k, v in row.items()}, flush=True)
    return rows


@pytest.mark.gpu
@pytest.mark.parametrize("layout", ["nchw", "channels_last"])
@pytest.mark.parametrize("name", NAMES)
def test_fullsize_parity(name, layout):
    from ovmono3d_b200 import ops
    from ovmono3d_b200.compat import ShapeSpec
    from ovmono3d_b200.config import get_cfg
    from ovmono3d_b200.roi_heads import build_roi_heads

    w = workloads.get(name)
    dev = torch.device("cuda:0")
    inp, weights = w.make_inputs(), w.make_weights()
    counts = [len(b) for b in inp.boxes]
    n = sum(counts)
    c = get_cfg()
    c.MODEL.ROI_HEADS.NAME = "ROIHeads3DB200"
    c.MODEL.ROI_HEADS.IN_FEATURES = list(w.in_features)
    c.MODEL.ROI_CUBE_HEAD.DIMS_PRIORS_ENABLED = False
    head = build_roi_heads(c, {f: ShapeSpec(channels=w.channels, stride=s) for f, s in zip(w.in_features, w.strides)})
    head.cube_head.load_state_dict(weights)
    head = head.to(dev).eval()
    eng = head._get_engine()
    eng.prepare(dev, head.priors_dims_per_cat, head.priors_z_scales, getattr(head, "priors_z_stats", None))
    if eng.tail_precision != "f16":
        pytest.skip("OVM_TAIL_PRECISION=%s: this test emulates the default fp16 tail operands" % eng.tail_precision)
    assert eng.tail_fused, "the full-size workloads run the fused tail"

    feats = [f.to(dev) for f in inp.features]
    if layout == "channels_last":     # the channels-last ROIAlign kernel: same bounds, nothing copied back to NCHW
        feats = [f.contiguous(memory_format=torch.channels_last) for f in feats]
    boxes = torch.cat(inp.boxes).to(dev)
    classes = torch.cat(inp.classes).to(dev)
    scores = torch.cat(inp.scores).to(dev)
    bidx = torch.from_numpy(np.repeat(np.arange(len(counts), dtype=np.int32), counts)).to(dev)
    cams = torch.zeros(len(counts), 12)
    for i in range(len(counts)):
        cams[i, :9] = inp.Ks[i].reshape(9)
        cams[i, 9] = inp.im_scales_ratio[i]
        cams[i, 10] = inp.im_dims[i][0]
    cams = cams.to(dev)
    rec, raw = eng.forward(feats, boxes, classes, scores, bidx, cams, want_raw=True)
    rec2, _ = eng.forward(feats, boxes, classes, scores, bidx, cams)      # second call: workspaces handed back clean
    pooled, levels = ops.roi_align(feats, _cfg(w).scales, boxes, bidx, 7, 0, torch.float32)
    torch.cuda.synchronize()
    rec, raw, pooled, levels = rec.cpu(), raw.cpu(), pooled.cpu(), levels.cpu()
    import os
    if os.environ.get("OVM_GEMM_SLOTS", "0") != "0":
        # fc1's split-K partials are summed from per-cluster slots in a fixed order: no atomics anywhere, bit-reproducible
        assert torch.equal(rec, rec2.cpu()), "a repeated call must reproduce the records bit for bit"
    else:
        # default: fc1's split-K partial sums meet in L2 atomics (red.global.add), whose order -- hence the last bits of
        # h1 -- varies from run to run
        assert float((rec - rec2.cpu()).abs().max()) <= 1e-4 * float(rec.abs().max()), "a repeated call must reproduce the records"

    ref, ref_raw6 = _oracle(w, inp, weights, False)
    # ---- levels: bit-exact, both from the stand-alone call and from inside the lift
    assert torch.equal(levels.long(), ref.levels)
    assert torch.equal(eng.last_levels[:n].cpu().long(), ref.levels)
    # ---- pooled features: 1e-3 per box
    want = ref.pooled.flatten(1)
    got = pooled.flatten(1)
    scale = want.abs().amax(dim=1).clamp_min(1e-6)
    assert float(((got - want).abs().amax(dim=1) / scale).max()) <= 1e-3
    del pooled, got, want

    # ---- records against the oracle with the product's operand rounding: kernel correctness
    # (fp32 accumulation over K = 49 C against the emulation's exact sums: ~1e-5 of the raw outputs; pose and corners
    # again through the conditioning of each box's 6-D vector)
    emu, emu_raw6 = _oracle(w, inp, weights, True)
    mine = _split_records(rec, counts)
    for f in DIRECT_FIELDS:
        mx, _ = _per_image_err(mine[f], getattr(emu, f))
        assert mx <= KERNEL_TOL, "%s vs product-operand oracle: %.3g" % (f, mx)
    emu_raw = torch.cat([_raw_heads(emu), emu_raw6], dim=1)
    kraw_err = (raw[:, :13] - emu_raw).abs()
    kraw_rel = float((kraw_err.amax(0) / emu_raw.abs().amax(0).clamp_min(1e-30)).max())
    assert kraw_rel <= KERNEL_TOL, "raw head outputs vs emulation: %.3g" % kraw_rel
    k_allowed = COND_SAFETY * kraw_err[:, 7:13].double().norm(dim=1) * _conditioning(emu_raw6) + KERNEL_TOL
    k_pose = (torch.cat(mine["pred_pose"]).double() - torch.cat(emu.pred_pose).double()).abs().amax(dim=(1, 2))
    assert bool((k_pose <= k_allowed).all()), "pose vs product-operand oracle: %.3g" % float((k_pose / k_allowed).max())
    k_dims = torch.cat(emu.pred_dimensions).double()
    k_scale = torch.cat([t.abs().max().expand(len(t)) for t in emu.pred_bbox3D]).double()
    k_corner = (torch.cat(mine["pred_bbox3D"]).double() - torch.cat(emu.pred_bbox3D).double()).abs().amax(dim=(1, 2))
    assert bool((k_corner <= KERNEL_TOL * k_scale + 0.5 * k_dims.norm(dim=1) * k_allowed * 3 ** 0.5).all()), "corners vs emulation"
    _, l2p = _per_image_err(mine["pred_pose"], emu.pred_pose)
    _, l2c = _per_image_err(mine["pred_bbox3D"], emu.pred_bbox3D)
    assert l2p <= KERNEL_TOL and l2c <= KERNEL_TOL, "pose / corner L2 vs emulation: %.3g %.3g" % (l2p, l2c)
    # ---- against the fp32 oracle: the format error of the bf16 head
    for f in DIRECT_FIELDS:
        mx, _ = _per_image_err(mine[f], getattr(ref, f))
        assert mx <= FORMAT_TOL, "%s vs fp32 oracle: %.3g" % (f, mx)
    want_raw = torch.cat([_raw_heads(ref), ref_raw6], dim=1)          # deltas 2, z, dims 3, uncertainty, pose 6
    got_raw = raw[:, :13]
    raw_err = (got_raw - want_raw).abs()
    assert float((raw_err.amax(0) / want_raw.abs().amax(0).clamp_min(1e-30)).max()) <= FORMAT_TOL, "raw head outputs"
    # pose of every box: explained by the error of its six raw numbers times the conditioning of the Gram-Schmidt
    eps6 = raw_err[:, 7:13].double().norm(dim=1)                       # per box
    allowed = COND_SAFETY * eps6 * _conditioning(ref_raw6) + 1e-3      # (first-order bound; the emulation's worst box needs 2.0, the kernel's 4.5)
    pose_err = (torch.cat(mine["pred_pose"]).double() - torch.cat(ref.pred_pose).double()).abs().amax(dim=(1, 2))
    worst = int((pose_err / allowed).argmax())
    assert bool((pose_err <= allowed).all()), "pose of box %d off by %.3g, allowed %.3g" % (
        worst, float(pose_err[worst]), float(allowed[worst]))
    # corners = R . (+-dims / 2) + centre
    dims = torch.cat(ref.pred_dimensions).double()
    scale = torch.cat([t.abs().max().expand(len(t)) for t in ref.pred_bbox3D]).double()
    corner_err = (torch.cat(mine["pred_bbox3D"]).double() - torch.cat(ref.pred_bbox3D).double()).abs().amax(dim=(1, 2))
    corner_allowed = FORMAT_TOL * scale + 0.5 * dims.norm(dim=1) * allowed * 3 ** 0.5
    assert bool((corner_err <= corner_allowed).all()), "corners"
    _, l2c = _per_image_err(mine["pred_bbox3D"], ref.pred_bbox3D)
    _, l2p = _per_image_err(mine["pred_pose"], ref.pred_pose)
    ang = _geodesic_deg(torch.cat(mine["pred_pose"]), torch.cat(ref.pred_pose))
    assert l2c <= CORNER_L2_TOL and l2p <= POSE_L2_TOL, "corner L2 %.3g, pose L2 %.3g" % (l2c, l2p)
    assert float(ang.median()) <= POSE_MEDIAN_DEG and float(ang.quantile(0.99)) <= POSE_P99_DEG, \
        "pose geodesic median %.3f p99 %.3f deg" % (float(ang.median()), float(ang.quantile(0.99)))


@pytest.mark.gpu
@pytest.mark.parametrize("fc_dim", [128, 256, 512])
def test_fused_tail_finishes_fc1_at_every_fc_dim(fc_dim):
    """FC_DIM 128 / 256 / 512 (clusters of 1 / 2 / 4 CTAs, 128 / 64 / 32 rows converted per CTA, the ungrouped and the
    grouped conversion): fc1 runs split-K and only accumulates, the fused tail applies its bias / ReLU / fp16 rounding on
    the way in and hands the accumulator back zeroed.  Kernel correctness against the oracle with the product's operand
    rounding, three calls in a row (a dirty accumulator would show in the second)."""
    from ovmono3d_b200.compat import ShapeSpec
    from ovmono3d_b200.config import get_cfg
    from ovmono3d_b200.roi_heads import build_roi_heads

    w = workloads.Workload("X-fc%d" % fc_dim, "fused tail at FC_DIM %d" % fc_dim, canvas=448, channels=64, num_images=2,
                           boxes_per_image=160, seed=77 + fc_dim, fc_dim=fc_dim)
    dev = torch.device("cuda:0")
    inp, weights = w.make_inputs(), w.make_weights()
    counts = [len(b) for b in inp.boxes]
    n = sum(counts)
    assert n >= 256, "the deferred finish is used from 256 boxes"
    c = get_cfg()
    c.MODEL.ROI_HEADS.NAME = "ROIHeads3DB200"
    c.MODEL.ROI_HEADS.IN_FEATURES = list(w.in_features)
    c.MODEL.ROI_CUBE_HEAD.DIMS_PRIORS_ENABLED = False
    c.MODEL.ROI_CUBE_HEAD.FC_DIM = fc_dim
    head = build_roi_heads(c, {f: ShapeSpec(channels=w.channels, stride=s) for f, s in zip(w.in_features, w.strides)})
    head.cube_head.load_state_dict(weights)
    head = head.to(dev).eval()
    eng = head._get_engine()
    eng.prepare(dev, head.priors_dims_per_cat, head.priors_z_scales, getattr(head, "priors_z_stats", None))
    if eng.tail_precision != "f16":
        pytest.skip("emulates the default fp16 tail operands")
    assert eng.tail_fused
    feats = [f.to(dev) for f in inp.features]
    boxes = torch.cat(inp.boxes).to(dev)
    classes = torch.cat(inp.classes).to(dev)
    scores = torch.cat(inp.scores).to(dev)
    bidx = torch.from_numpy(np.repeat(np.arange(len(counts), dtype=np.int32), counts)).to(dev)
    cams = torch.zeros(len(counts), 12)
    for i in range(len(counts)):
        cams[i, :9] = inp.Ks[i].reshape(9)
        cams[i, 9] = inp.im_scales_ratio[i]
        cams[i, 10] = inp.im_dims[i][0]
    cams = cams.to(dev)
    recs = [eng.forward(feats, boxes, classes, scores, bidx, cams, want_raw=True) for _ in range(3)]
    torch.cuda.synchronize()
    rec, raw = recs[0][0].cpu(), recs[0][1].cpu()
    for r2, _ in recs[1:]:
        assert float((rec - r2.cpu()).abs().max()) <= 1e-4 * float(rec.abs().max()), "a repeated call must reproduce the records"
    emu, emu_raw6 = _oracle(w, inp, weights, True)
    mine = _split_records(rec, counts)
    for f in DIRECT_FIELDS:
        mx, _ = _per_image_err(mine[f], getattr(emu, f))
        assert mx <= KERNEL_TOL, "%s vs product-operand oracle: %.3g" % (f, mx)
    emu_raw = torch.cat([_raw_heads(emu), emu_raw6], dim=1)
    kraw_err = (raw[:, :13] - emu_raw).abs()
    assert float((kraw_err.amax(0) / emu_raw.abs().amax(0).clamp_min(1e-30)).max()) <= KERNEL_TOL, "raw head outputs"
    k_allowed = COND_SAFETY * kraw_err[:, 7:13].double().norm(dim=1) * _conditioning(emu_raw6) + KERNEL_TOL
    k_pose = (torch.cat(mine["pred_pose"]).double() - torch.cat(emu.pred_pose).double()).abs().amax(dim=(1, 2))
    assert bool((k_pose <= k_allowed).all()), "pose vs product-operand oracle"


if __name__ == "__main__":
    if "--format-error" in sys.argv:
        torch.set_num_threads(torch.get_num_threads())
        format_error_table([a for a in sys.argv[1:] if not a.startswith("--")] or NAMES)
