"""`ovm_fc_tail` (last FC layer + heads + decode in one launch, csrc/fc_tail.cu) against the unfused sequence it
replaces -- an fp64 matmul of the SAME rounded operands + bias + ReLU, then `ovm_head_decode` (itself held to the
oracle in test_lift_gpu.py) -- for every cluster size (FC_DIM 128 ... 1024), both operand formats, ragged n and K.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

FIELD_SLICES = dict(pred_bbox3D=slice(0, 24), pred_center_cam=slice(24, 27), pred_center_2D=slice(27, 29),
                    pred_dimensions=slice(29, 32), pred_pose=slice(32, 41), scores=slice(45, 46))


def _tf32(x: torch.Tensor) -> torch.Tensor:
    """cvt.rna.tf32.f32: nearest, ties away from zero, 10 mantissa bits."""
    i = x.contiguous().view(torch.int32)
    return ((i + 0x1000) & ~0x1FFF).view(torch.float32)


def _setup(fc_dim, pose_type="6d", z_type="direct", use_confidence=1.0, seed=0):
    from ovmono3d_b200 import ops, synth
    from ovmono3d_b200.engine import DecodeSettings, LiftEngine
    from ovmono3d_b200.roi_heads import CubeHead
    from ovmono3d_b200.compat import ShapeSpec
    from ovmono3d_b200.config import get_cfg

    dev = torch.device("cuda:0")
    c = get_cfg()
    c.MODEL.ROI_CUBE_HEAD.DIMS_PRIORS_ENABLED = False
    c.MODEL.ROI_CUBE_HEAD.FC_DIM = fc_dim
    c.MODEL.ROI_CUBE_HEAD.POSE_TYPE = pose_type
    c.MODEL.ROI_CUBE_HEAD.Z_TYPE = z_type
    c.MODEL.ROI_CUBE_HEAD.USE_CONFIDENCE = use_confidence
    head = CubeHead(c, ShapeSpec(channels=4, height=7, width=7))
    w = synth.make_head_weights(seed + 3, in_dim=4 * 49, fc_dim=fc_dim, pose_type=pose_type,
                                use_confidence=bool(use_confidence))
    head.load_state_dict(w, strict=True)
    head = head.to(dev)
    s = DecodeSettings(z_type=z_type, pose_type=pose_type, dims_priors_enabled=False, use_confidence=use_confidence)
    eng = LiftEngine(head, (1 / 7, 1 / 14, 1 / 28), 7, 0, s)
    eng.prepare(dev, None, None, None)
    return ops, eng, dev


def _inputs(n, num_images, dev, seed):
    from ovmono3d_b200 import synth

    per = [n // num_images + (i < n % num_images) for i in range(num_images)]
    inp = synth.make_inputs(seed, num_images, per, channels=1)
    boxes = torch.cat(inp.boxes).to(dev)
    classes = torch.cat(inp.classes).to(dev)
    scores = torch.cat(inp.scores).to(dev)
    batch_idx = torch.repeat_interleave(torch.arange(num_images, dtype=torch.int32), torch.tensor(per)).to(dev)
    cams = torch.zeros(num_images, 12)
    for i in range(num_images):
        cams[i, :9] = inp.Ks[i].reshape(9)
        cams[i, 9] = inp.im_scales_ratio[i]
        cams[i, 10] = inp.im_dims[i][0]
    return boxes, classes, scores, batch_idx, cams.to(dev)


@pytest.mark.parametrize("fc_dim", [128, 256, 512, 1024])
@pytest.mark.parametrize("fmt", ["f16", "tf32", "bf16"])
@pytest.mark.parametrize("n,K", [(1, 1024), (128, 64), (300, 200), (1600, 1024)])
def test_fc_tail_matches_unfused_sequence(fc_dim, fmt, n, K):
    ops, eng, dev = _setup(fc_dim)
    g = torch.Generator().manual_seed(fc_dim + n + K)
    h1 = torch.relu(torch.randn(n, K, generator=g)).to(dev)                 # the layer before ends in a ReLU
    W2 = ((torch.rand(fc_dim, K, generator=g) * 2 - 1) * (3.0 / K) ** 0.5).to(dev)
    b2 = (torch.randn(fc_dim, generator=g) * 0.1).to(dev)
    boxes, classes, scores, batch_idx, cams = _inputs(n, 3, dev, 7)
    if fmt == "tf32":
        A, W = _tf32(h1), ops.round_tf32(W2.contiguous())
        assert torch.equal(W, _tf32(W2)), "ovm_round_tf32 differs from cvt.rna emulation"
    elif fmt == "f16":
        ld = (K + 7) // 8 * 8
        A = torch.zeros(n, ld, device=dev, dtype=torch.float16)[:, :K]
        A.copy_(h1)
        W = ops.cast_f16(W2.contiguous())
        assert torch.equal(W, W2.half()), "ovm_cast_f16 differs from torch's fp16 rounding"
    else:
        ld = (K + 7) // 8 * 8
        A = torch.zeros(n, ld, device=dev, dtype=torch.bfloat16)[:, :K]
        A.copy_(h1)
        W = ops.cast_bf16(W2.contiguous())
    rec, raw = ops.fc_tail(eng._cfg, eng._heads, A, W, b2, boxes, classes, scores, batch_idx, cams, want_raw=True)
    h2 = torch.relu(A.double() @ W.double().t() + b2.double()).float().contiguous()
    want, want_raw = ops.head_decode(eng._cfg, eng._heads, h2, boxes, classes, scores, batch_idx, cams, want_raw=True)
    torch.cuda.synchronize()
    assert torch.isfinite(rec).all()
    # the 13 head outputs: same operands, fp32 accumulation in a different order
    scale = want_raw.abs().max().clamp_min(1e-6)
    assert float((raw - want_raw).abs().max() / scale) <= 2e-5
    for f, sl in FIELD_SLICES.items():
        a, b = rec[:, sl].double(), want[:, sl].double()
        e = float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))
        assert e <= 1e-3, "%s: %.3g" % (f, e)
    assert torch.equal(rec[:, 41:45], want[:, 41:45]) and torch.equal(rec[:, 46:], want[:, 46:])   # box, class pass through


@pytest.mark.parametrize("pose_type,z_type,conf", [("quaternion", "log", 1.0), ("euler", "sigmoid", 0.0)])
def test_fc_tail_head_variants(pose_type, z_type, conf):
    ops, eng, dev = _setup(256, pose_type, z_type, conf, seed=5)
    n, K = 200, 256
    g = torch.Generator().manual_seed(99)
    h1 = _tf32(torch.relu(torch.randn(n, K, generator=g)).to(dev))
    W = ops.round_tf32(((torch.rand(256, K, generator=g) * 2 - 1) * (3.0 / K) ** 0.5).to(dev))
    b2 = (torch.randn(256, generator=g) * 0.1).to(dev)
    boxes, classes, scores, batch_idx, cams = _inputs(n, 2, dev, 8)
    rec, _ = ops.fc_tail(eng._cfg, eng._heads, h1, W, b2, boxes, classes, scores, batch_idx, cams)
    h2 = torch.relu(h1.double() @ W.double().t() + b2.double()).float().contiguous()
    want, _ = ops.head_decode(eng._cfg, eng._heads, h2, boxes, classes, scores, batch_idx, cams)
    for f, sl in FIELD_SLICES.items():
        a, b = rec[:, sl].double(), want[:, sl].double()
        ok = torch.isfinite(b)      # without the confidence head the score is sqrt(score * v): NaN for v < 0 in both
        assert torch.equal(torch.isfinite(a), ok), f
        e = float((a - b)[ok].abs().max() / b[ok].abs().max().clamp_min(1e-30))
        assert e <= 1e-3, "%s: %.3g (max |want| %.3g)" % (f, e, float(b[ok].abs().max()))


def test_fc_tail_is_deterministic_and_rejects_unsupported():
    ops, eng, dev = _setup(1024)
    n, K = 777, 1024
    g = torch.Generator().manual_seed(1)
    h1 = _tf32(torch.relu(torch.randn(n, K, generator=g)).to(dev))
    W = ops.round_tf32(((torch.rand(1024, K, generator=g) * 2 - 1) * (3.0 / K) ** 0.5).to(dev))
    b2 = torch.zeros(1024, device=dev)
    boxes, classes, scores, batch_idx, cams = _inputs(n, 4, dev, 9)
    r0, _ = ops.fc_tail(eng._cfg, eng._heads, h1, W, b2, boxes, classes, scores, batch_idx, cams)
    for _ in range(3):
        r1, _ = ops.fc_tail(eng._cfg, eng._heads, h1, W, b2, boxes, classes, scores, batch_idx, cams)
        assert torch.equal(r0, r1), "partial sums are combined in a fixed order: results must be bit-identical"
    # FC_DIM that does not split into 128-column CTAs of a power-of-two cluster
    ops2, eng2, _ = _setup(384)
    W3 = ops.round_tf32(torch.zeros(384, K, device=dev))
    with pytest.raises(ops.LiftError):
        ops.fc_tail(eng2._cfg, eng2._heads, h1, W3, torch.zeros(384, device=dev), boxes, classes, scores, batch_idx, cams)
