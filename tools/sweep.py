"""BASELINE.json configs[2] (C3: box-count sweep at 1024-px SFP features) and configs[4] (C5: backbone-variant
pyramids, channel / level shape changes) through the same lift path, on one B200.

    python tools/sweep.py > gpurun_out/sweep.jsonl

Per case: device-resident `LiftEngine.forward` (ROIAlign + FC stack + heads/decode) and `ovm_roi_align` alone, CUDA
events, 256 MiB L2 flush before every call, median of 15.  Reports lifted boxes/s, the ROIAlign time and its
algorithmic HBM GB/s (clipped footprint at the assigned level + bf16 output per box).
Parity of these shapes is covered by tests/ (clip_4level, dla_5level, ... golden cases); this is measurement only.
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ovmono3d_b200 import ops, synth  # noqa: E402
from ovmono3d_b200.compat import ShapeSpec  # noqa: E402
from ovmono3d_b200.config import get_cfg  # noqa: E402
from ovmono3d_b200.roi_heads import build_roi_heads  # noqa: E402


def algorithmic_bytes(boxes, levels, strides, canvas, channels, out_elt=2):
    total = 0.0
    for l, s in enumerate(strides):
        m = levels == l
        if not m.any():
            continue
        b = boxes[m].astype(np.float64)
        size = canvas // s
        w = np.clip(np.floor(b[:, 2] / s - 0.5) - np.floor(b[:, 0] / s - 0.5) + 2, 1, size)
        h = np.clip(np.floor(b[:, 3] / s - 0.5) - np.floor(b[:, 1] / s - 0.5) + 2, 1, size)
        total += float((w * h).sum()) * channels * 4
    return total + boxes.shape[0] * channels * 49 * out_elt


def median_ms(fn, flush, iters=15):
    ts = []
    for i in range(iters + 3):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        if i >= 3:
            ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


def run_case(name, strides, canvas, channels, num_images, boxes_per_image, dev, flush):
    names = ["p%d" % (2 + i) for i in range(len(strides))]
    c = get_cfg()
    c.MODEL.ROI_HEADS.NAME = "ROIHeads3DB200"
    c.MODEL.ROI_HEADS.IN_FEATURES = names
    c.MODEL.ROI_CUBE_HEAD.DIMS_PRIORS_ENABLED = False
    head = build_roi_heads(c, {f: ShapeSpec(channels=channels, stride=s) for f, s in zip(names, strides)})
    head.cube_head.load_state_dict(synth.make_head_weights(77, in_dim=channels * 49, fc_dim=1024))
    head = head.to(dev).eval()
    inp = synth.make_inputs(4321, num_images, boxes_per_image, strides=strides, canvas=canvas, channels=channels)
    feats = [f.to(dev) for f in inp.features]
    counts = [len(b) for b in inp.boxes]
    n = sum(counts)
    boxes = torch.cat(inp.boxes).to(dev)
    classes = torch.cat(inp.classes).to(dev)
    scores = torch.cat(inp.scores).to(dev)
    bidx = torch.from_numpy(np.repeat(np.arange(num_images, dtype=np.int32), counts)).to(dev)
    cams = torch.zeros(num_images, 12)
    for i in range(num_images):
        cams[i, :9] = inp.Ks[i].reshape(9)
        cams[i, 9] = inp.im_scales_ratio[i]
        cams[i, 10] = inp.im_dims[i][0]
    cams = cams.to(dev)
    eng = head._get_engine()
    eng.prepare(dev, head.priors_dims_per_cat, head.priors_z_scales, getattr(head, "priors_z_stats", None))
    scales = [1.0 / s for s in strides]

    lift_ms = median_ms(lambda: eng.forward(feats, boxes, classes, scores, bidx, cams), flush)
    roi_ms = median_ms(lambda: ops.roi_align(feats, scales, boxes, bidx, 7, 0, torch.bfloat16), flush)
    rec = eng.forward(feats, boxes, classes, scores, bidx, cams)[0]
    assert torch.isfinite(rec).all()
    levels = eng.last_levels[:n].cpu().numpy()
    alg = algorithmic_bytes(boxes.cpu().numpy(), levels, strides, canvas, channels)
    print(json.dumps({"case": name, "strides": list(strides), "canvas": canvas, "channels": channels,
                      "images": num_images, "boxes_per_image": boxes_per_image, "boxes": n,
                      "pyramid_MB": sum(f.numel() * 4 for f in feats) / 1e6,
                      "lift_ms": lift_ms, "boxes_per_s": n / lift_ms * 1e3, "roi_align_ms": roi_ms,
                      "roi_align_algorithmic_GBps": alg / roi_ms / 1e6,
                      "boxes_per_level": np.bincount(levels[levels >= 0], minlength=len(strides)).tolist()}), flush=True)
    del feats, head, eng
    torch.cuda.empty_cache()


def main():
    dev = torch.device("cuda:0")
    ops.load_library()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    # C3: 1024-px SFP pyramid of the patch-16 variants (strides 4/8/16/32), N = 1 and 8, 10 .. 2000 boxes per image
    for N in (1, 8):
        for k in (10, 30, 100, 300, 1000, 2000):
            run_case("C3", (4, 8, 16, 32), 1024, 256, N, k, dev, flush)
    # C5: pyramids of the other shipped backbones and the channel-count change in fc1's K = 49 C
    run_case("C5 dinov2 SFP", (7, 14, 28), 896, 256, 16, 100, dev, flush)
    run_case("C5 clip/mae/sam/midas SFP", (4, 8, 16, 32), 1024, 256, 16, 100, dev, flush)
    run_case("C5 DLA-FPN p2-p6", (4, 8, 16, 32, 64), 1024, 256, 16, 100, dev, flush)
    run_case("C5 channels 128", (7, 14, 28), 896, 128, 16, 100, dev, flush)
    run_case("C5 channels 512", (7, 14, 28), 896, 512, 16, 100, dev, flush)


if __name__ == "__main__":
    main()
