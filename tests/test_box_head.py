"""2-D box branch at inference (`ovm_box_inference`, ovmono3d_b200/box_head.py).

CPU (`-m "not gpu"`): the oracle's `fast_rcnn_inference` against vectors written by the reference's own function
(tests/golden/box_inference_*.npz); module names / shapes of the branch.  GPU: `ovm_box_inference` through the C ABI on
the same logits / deltas -- kept proposal rows and classes bit-exact, scores to 2e-6 relative and boxes to 1e-5
relative + 1e-4 px (device `expf` vs torch's vectorised CPU `exp`) -- and the whole branch (ROIAlign -> tcgen05 FC
stack -> predictor -> inference) against the oracle with bf16 GEMM operands."""
import numpy as np
import pytest
import torch

from oracle import box_head_oracle as bo
from tests.golden import box_cases as bc

SCORE_RTOL = 2e-6
BOX_RTOL, BOX_ATOL = 1e-5, 1e-4


@pytest.fixture(autouse=True)
def _cpu_oracle_nms_rule(monkeypatch):
    """The oracle is torchvision's CPU batched_nms, which leaves the coordinate trick above 1000 candidates (CUDA: 5000,
    the plugin's default): run the plugin with the CPU rule so that both sides take the same arithmetic."""
    from ovmono3d_b200.box_head import BoxBranchB200

    monkeypatch.setattr(BoxBranchB200, "COORD_TRICK_MAX", 1000)



def _oracle(case):
    logits, deltas, props = bc.build_case(case)
    boxes, probs = bo.predict(logits, deltas, props, bc.REG_WEIGHTS)
    out = bo.fast_rcnn_inference(boxes, probs, list(case.image_sizes), case.score_thresh, case.nms_thresh, case.topk)
    cands = [int((p[torch.isfinite(b).all(1) & torch.isfinite(p).all(1)][:, :-1] > case.score_thresh).sum())
             for b, p in zip(boxes, probs)]
    return logits, deltas, props, out, cands


@pytest.mark.parametrize("case", bc.CASES, ids=lambda c: c.name)
def test_oracle_matches_reference_vectors(case):
    logits, deltas, props, out, _ = _oracle(case)
    g = np.load(bc.golden_path(case))
    assert str(g["checksum"]) == bc.input_checksum(logits, deltas, props)
    step = max(1, case.num_classes // 5)
    for i, (b, s, sf, c, idx) in enumerate(out):
        assert np.array_equal(b.numpy(), g["im%d_boxes" % i])
        assert np.array_equal(s.numpy(), g["im%d_scores" % i])
        assert np.array_equal(sf.numpy()[:, ::step], g["im%d_scores_full" % i])
        assert np.array_equal(c.numpy(), g["im%d_classes" % i])
        assert np.array_equal(idx.numpy(), g["im%d_index" % i])


def test_cases_cover_both_torchvision_paths():
    cands = {c.name: _oracle(c)[4] for c in bc.CASES}
    assert 0 < max(cands["box_inference_small"]) <= 1000          # coordinate trick
    assert min(cands["box_inference_1000x50"]) > 1000             # class-by-class loop
    assert cands["box_inference_small"][1] == 0                   # an image without proposals


def test_branch_modules_have_the_reference_names():
    from ovmono3d_b200.compat import ShapeSpec
    from ovmono3d_b200.config import get_cfg
    from ovmono3d_b200.roi_heads import build_roi_heads
    from ovmono3d_b200 import ops

    c = get_cfg()
    c.MODEL.ROI_HEADS.NAME = "ROIHeads3DB200"
    c.MODEL.ROI_HEADS.IN_FEATURES = ["p2", "p3", "p4"]
    c.MODEL.ROI_HEADS.NUM_CLASSES = 50
    c.MODEL.ROI_BOX_HEAD.FC_DIM = 64
    shape = {f: ShapeSpec(channels=8, stride=s) for f, s in zip(["p2", "p3", "p4"], (7, 14, 28))}
    plain = build_roi_heads(c, shape).eval()
    assert not any(k.startswith("box_") for k in plain.state_dict())
    with pytest.raises(ops.LiftError):
        plain._forward_box({}, [])
    c.MODEL.ROI_BOX_HEAD.BUILD = True
    head = build_roi_heads(c, shape)
    sd = head.state_dict()
    # nohup.out:645-661
    assert sd["box_head.fc1.weight"].shape == (64, 8 * 49) and sd["box_head.fc2.weight"].shape == (64, 64)
    assert sd["box_predictor.cls_score.weight"].shape == (51, 64)
    assert sd["box_predictor.bbox_pred.weight"].shape == (200, 64)
    assert head.box_predictor.test_topk_per_image == 100 and head.box_predictor.box2box_weights == (10.0, 10.0, 5.0, 5.0)


# ------------------------------------------------------------------------------------------------------------ GPU

def _dev():
    return torch.device("cuda:0")


def _run_kernel(case, logits, deltas, props):
    from ovmono3d_b200 import ops

    dev = _dev()
    counts = [len(p) for p in props]
    off = torch.tensor([0] + list(np.cumsum(counts)), dtype=torch.int32, device=dev)
    sizes = torch.tensor([list(s) for s in case.image_sizes], dtype=torch.int32, device=dev)
    before = ops.launch_count()
    out = ops.box_inference(logits.to(dev), deltas.to(dev), torch.cat(props).to(dev), off, sizes, max(counts),
                            case.num_classes, bc.REG_WEIGHTS, case.score_thresh, case.nms_thresh, case.topk,
                            case.cls_agnostic, 1000)
    assert ops.launch_count() - before == 4        # stats memset + three kernels
    return out


def _compare(case, got, want, index_exact=True):
    ob, osc, oc, oi, ofull, cnt = got
    for i, (b, s, sf, c, idx) in enumerate(want):
        k = int(cnt[i])
        assert k == len(s), (i, k, len(s))
        assert oc[i, :k].cpu().tolist() == c.tolist()
        if index_exact:
            assert oi[i, :k].cpu().long().tolist() == idx.tolist()
        gs = osc[i, :k].cpu()
        assert torch.all((gs - s).abs() <= SCORE_RTOL * s.abs())
        assert torch.all((ob[i, :k].cpu() - b).abs() <= BOX_RTOL * b.abs() + BOX_ATOL)
        assert torch.all((ofull[i, :k].cpu() - sf).abs() <= SCORE_RTOL * sf.abs() + 1e-12)


@pytest.mark.gpu
@pytest.mark.parametrize("case", bc.CASES, ids=lambda c: c.name)
def test_box_inference_kernel_matches_oracle(case):
    logits, deltas, props, want, _ = _oracle(case)
    got = _run_kernel(case, logits, deltas, props)
    # with non-finite rows the reference's kept indices count only the finite rows (fast_rcnn.py:75-78); the kernel
    # reports the proposal row itself -- map it before comparing
    if case.nonfinite_rows:
        boxes, probs = bo.predict(logits, deltas, props, bc.REG_WEIGHTS)
        ob, osc, oc, oi, ofull, cnt = got
        oi = oi.clone()
        for i, (b, p) in enumerate(zip(boxes, probs)):
            valid = torch.isfinite(b).all(1) & torch.isfinite(p).all(1)
            assert i > 0 or not valid.all()
            rank = (torch.cumsum(valid.long(), 0) - 1).to(oi.device)
            k = int(cnt[i])
            oi[i, :k] = rank[oi[i, :k].long()].int()
        got = (ob, osc, oc, oi, ofull, cnt)
    _compare(case, got, want)


@pytest.mark.gpu
def test_box_inference_errors_and_empty():
    from ovmono3d_b200 import ops

    dev = _dev()
    z = lambda *s: torch.zeros(*s, device=dev)
    off = torch.tensor([0, 0], dtype=torch.int32, device=dev)
    sizes = torch.tensor([[8, 8]], dtype=torch.int32, device=dev)
    out = ops.box_inference(z(0, 6), z(0, 20), z(0, 4), off, sizes, 0, 5, bc.REG_WEIGHTS, 0.05, 0.5, 10)
    assert out[5].tolist() == [0]
    with pytest.raises(ops.LiftError):            # merge capacity: 1024 classes x 100 survivors
        off2 = torch.tensor([0, 4], dtype=torch.int32, device=dev)
        ops.box_inference(z(4, 1025), z(4, 4096), z(4, 4), off2, sizes, 200, 1024, bc.REG_WEIGHTS, 0.05, 0.5, 100)
    with pytest.raises(ops.LiftError):
        ops.box_inference(torch.zeros(4, 6), z(4, 20), z(4, 4), off, sizes, 4, 5, bc.REG_WEIGHTS, 0.05, 0.5, 10)


@pytest.mark.gpu
def test_forward_box_branch_end_to_end():
    """ROIHeads3DB200._forward_box on a 3-level DINOv2-style pyramid: predictor outputs against the bf16-operand
    oracle (1e-2 of the row scale), then the detections against the oracle's inference run on the branch's own
    predictor outputs (rows / classes exact), then forward() -> _forward_box -> _forward_cube."""
    from ovmono3d_b200.compat import Boxes, Instances, ShapeSpec
    from ovmono3d_b200.config import get_cfg
    from ovmono3d_b200.roi_heads import build_roi_heads
    from ovmono3d_b200 import synth

    dev = _dev()
    inp = synth.make_inputs(4321, 2, 150)
    feats_cpu = [f[:, :32].contiguous() for f in inp.features]        # 32 channels keep the CPU oracle quick
    c = get_cfg()
    c.MODEL.ROI_HEADS.NAME = "ROIHeads3DB200"
    c.MODEL.ROI_HEADS.IN_FEATURES = ["p2", "p3", "p4"]
    c.MODEL.ROI_HEADS.NUM_CLASSES = 50
    c.MODEL.ROI_HEADS.SCORE_THRESH_TEST = 0.01
    c.MODEL.ROI_CUBE_HEAD.DIMS_PRIORS_ENABLED = False
    c.MODEL.ROI_BOX_HEAD.BUILD = True
    shape = {f: ShapeSpec(channels=32, stride=s) for f, s in zip(["p2", "p3", "p4"], (7, 14, 28))}
    torch.manual_seed(3)
    head = build_roi_heads(c, shape)
    with torch.no_grad():                 # trained-scale predictor weights: peaky scores, real box deltas
        head.box_predictor.cls_score.weight.normal_(0, 0.08)
        head.box_predictor.bbox_pred.weight.normal_(0, 0.02)
    weights = {k: v.detach().clone() for k, v in head.state_dict().items() if k.startswith("box_")}
    head = head.to(dev).eval()
    feats = {f: t.to(dev) for f, t in zip(["p2", "p3", "p4"], feats_cpu)}
    sizes = [tuple(d) for d in inp.im_dims]
    props = []
    for i in range(2):
        p = Instances(sizes[i])
        p.proposal_boxes = Boxes(inp.boxes[i].clone().to(dev))
        props.append(p)
    res = head._forward_box(feats, props)
    br = head._box_branch
    # predictor outputs vs the oracle with bf16 operands
    x = bo.box_head_features(feats_cpu, inp.boxes, (1 / 7, 1 / 14, 1 / 28), weights, bf16_operands=True)
    want_logits, want_deltas = bo.box_predictor(x, weights, bf16_operands=True)
    for got, want in ((br.last_logits, want_logits), (br.last_deltas, want_deltas)):
        err = (got.cpu() - want).abs().max() / want.abs().max()
        assert err < 1e-2, float(err)
    # detections vs the oracle's inference on the branch's own predictor outputs
    boxes, probs = bo.predict(br.last_logits.cpu(), br.last_deltas.cpu(), inp.boxes, (10.0, 10.0, 5.0, 5.0))
    want = bo.fast_rcnn_inference(boxes, probs, sizes, 0.01, 0.5, 100)
    for i, (b, s, sf, cl, idx) in enumerate(want):
        assert len(res[i]) == len(s) > 0
        assert res[i].pred_classes.cpu().tolist() == cl.tolist()
        assert br.last_index[i].cpu().tolist() == idx.tolist()
        assert torch.all((res[i].scores.cpu() - s).abs() <= SCORE_RTOL * s.abs())
        assert torch.all((res[i].pred_boxes.tensor.cpu() - b).abs() <= BOX_RTOL * b.abs() + BOX_ATOL)
        assert res[i].scores_full.shape == (len(s), 50)
    # forward(): proposals -> 2-D detections -> lift
    out, _ = head([torch.zeros(3, *sizes[0], device=dev)] * 2, feats, props, [k.clone() for k in inp.Ks],
                  list(inp.im_scales_ratio))
    assert all(o.has("pred_bbox3D") and o.pred_bbox3D.shape == (len(o), 8, 3) for o in out)
    assert all(torch.isfinite(o.pred_bbox3D).all() for o in out)
