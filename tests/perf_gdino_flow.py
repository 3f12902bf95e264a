"""The open-vocabulary flow of `ROIHeads3DGDINO.forward` after the detector itself (roi_heads_gdino.py:137-171) for ONE
image: raw GroundingDINO outputs (900 queries x 256 tokens, stand-in tensors) -> post-processing -> prompt boxes ->
lift, through `ROIHeads3DGDINOB200` with `GroundingDinoPostB200` as its detector; next to the reference's own sequence
on the same GPU (oracle post-processing with its `.cpu()` copies and host NMS, then the oracle lift on CUDA tensors).
Wall clock per call with a device synchronisation at the end (the caller consumes the Instances).  One JSON line.

    python tests/perf_gdino_flow.py [--iters 100]
"""
import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402
from oracle import gdino_post_oracle as go, lift_oracle as lo   # noqa: E402  (baseline leg only)
from ovmono3d_b200 import gdino_post as gp, ops                 # noqa: E402
from tests.golden import gdino_cases as gd                       # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=100)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    ops.load_library()
    head, names = bench.build_head(dev)
    inp = bench.make_batch(999)
    feats = {f: t[:1].to(dev).contiguous() for f, t in zip(names, inp.features)}
    H, W = [int(x) for x in inp.im_dims[0]]
    case = gd.GdinoCase("flow", 31, 900, 256, gd.OMNI, (H, W), box_threshold=0.05, clusters=40)
    logits, boxes, tok, caption, prompt = gd.build_case(case)
    cats = [c[0] for c in prompt]
    lg, bx = logits.to(dev), boxes.to(dev)

    class Model:
        tokenizer = tok

        def __call__(self, image, captions):
            return {"pred_logits": lg, "pred_boxes": bx}

    det = gp.GroundingDinoPostB200(Model(), box_threshold=0.05)
    from ovmono3d_b200.roi_heads import ROIHeads3DGDINOB200
    head.__class__ = ROIHeads3DGDINOB200          # same modules / parameters; the GDINO forward on top
    head.groundingdino_model = det
    images = [torch.zeros(3, H, W, device=dev)]
    Ks, ratios = [inp.Ks[0]], [inp.im_scales_ratio[0]]

    def ours():
        res, _ = head(images, feats, None, Ks, ratios, category_list=cats)
        torch.cuda.synchronize()
        return res

    for _ in range(5):
        res = ours()
    t0 = time.perf_counter()
    for _ in range(a.iters):
        res = ours()
    ours_ms = (time.perf_counter() - t0) * 1e3 / a.iters
    n = len(res[0])
    # where the call's time goes: detector wrapper (post-processing + count read-back) and the lift, separately
    img = images[0].flip(0)
    prompt_l = [[c] for c in cats]
    t0 = time.perf_counter()
    for _ in range(a.iters):
        ov = det(img, prompt_l)
    torch.cuda.synchronize()
    det_ms = (time.perf_counter() - t0) * 1e3 / a.iters
    from ovmono3d_b200.compat import Boxes, Instances
    t0 = time.perf_counter()
    for _ in range(a.iters):
        tgt = Instances((H, W))
        tgt.pred_boxes, tgt.pred_classes, tgt.scores = Boxes(ov["bboxes"]), ov["pred_classes"], ov["scores"]
        head._forward_cube(feats, [tgt], Ks, [(H, W)], ratios)
        torch.cuda.synchronize()
    cube_ms = (time.perf_counter() - t0) * 1e3 / a.iters

    ids = go.phrase_token_ids(tok(caption)["input_ids"], tok(cats, add_special_tokens=False)["input_ids"])
    weights = {k: v.detach() for k, v in head.cube_head.state_dict().items()}     # on the GPU: the reference's GPU path
    cfg = lo.LiftConfig(scales=tuple(1.0 / s for s in bench.STRIDES), dims_priors_enabled=False)
    feat_list = [feats[f] for f in names]

    def ref():
        out = go.gdino_post(lg[0], bx[0], ids, (H, W), 0.05, 0.5)
        with torch.no_grad():     # target.to(device) + _forward_cube, roi_heads_gdino.py:168-170
            r = lo.forward_cube(feat_list, [out["boxes"].to(dev)], [out["classes"].to(dev)],
                                [out["scores"].to(dev)], Ks, [inp.im_dims[0]], ratios, weights, cfg)
        torch.cuda.synchronize()
        return out, r

    for _ in range(3):
        out, r = ref()
    t0 = time.perf_counter()
    k = max(5, a.iters // 10)
    for _ in range(k):
        out, r = ref()
    ref_ms = (time.perf_counter() - t0) * 1e3 / k
    assert len(out["scores"]) == n
    got, want = res[0].pred_center_cam.cpu(), r.pred_center_cam[0].cpu()
    assert float((got - want).abs().max() / want.abs().max()) <= 1e-2       # same answer (bf16-operand head)
    res = {"workload": "one image, 900 queries x 256 tokens, %d phrases -> %d prompt boxes -> lift "
                                  "(DINOv2 SFP pyramid, head 12544-1024-1024-13)" % (len(cats), n),
                      "ours_ms_per_call": ours_ms, "ours_detector_wrapper_ms": det_ms, "ours_forward_cube_ms": cube_ms, "reference_gpu_sequence_ms_per_call": ref_ms,
                      "speedup": ref_ms / ours_ms}
    print(json.dumps(res))
    return res


if __name__ == "__main__":
    main()
