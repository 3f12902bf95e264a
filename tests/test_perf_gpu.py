"""The lift against the reference's GPU PyTorch path on the C2 workload (baseline ii).

The reference's GPU path is restated by `oracle/` moved to `cuda`: torchvision's CUDA `roi_align` per level with the
`nonzero` syncs of d2's ROIPooler, fp32 cuBLAS `F.linear`, the eager decode, and the per-box camera tensors built on
the CPU and copied over exactly as roi_heads.py:378-400 does.  BASELINE.json's target is >= 50x that path in lifted
boxes per second on one B200; the test asserts that target and writes the measured ratio to
`gpurun_out/ref_gpu_path.json` (copied into profiles/ by hand).
"""
import json
import os

import numpy as np
import pytest
import torch

from oracle import lift_oracle as lo

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TARGET_SPEEDUP = 50.0       # BASELINE.json north_star: >= 50x the reference GPU PyTorch path on one B200
POST_TARGET = 10.0          # the rows either side of the lift against the reference's own call sequence on the same GPU


def _time_ms(fn, iters, flush):
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts)), float(min(ts))


@pytest.mark.gpu
def test_c2_speed_against_reference_gpu_pytorch_path():
    import bench
    from ovmono3d_b200 import ops

    dev = torch.device("cuda:0")
    ops.load_library()
    head, names = bench.build_head(dev)
    inp = bench.make_batch(1235)
    db = bench.DeviceBatch(inp, dev)
    eng = head._get_engine()
    eng.prepare(dev, head.priors_dims_per_cat, head.priors_z_scales, getattr(head, "priors_z_stats", None))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def ours():
        return eng.forward(db.features, db.boxes, db.classes, db.scores, db.batch_idx, db.cams)[0]

    cfg = lo.LiftConfig(scales=tuple(1.0 / s for s in bench.STRIDES), dims_priors_enabled=False)
    weights = {k: v.to(dev) for k, v in bench.make_weights().items()}
    feats = [f.to(dev) for f in inp.features]
    boxes = [b.to(dev) for b in inp.boxes]
    classes = [c.to(dev) for c in inp.classes]
    scores = [s.to(dev) for s in inp.scores]

    def ref():
        with torch.no_grad():
            return lo.forward_cube(feats, boxes, classes, scores, inp.Ks, inp.im_dims, inp.im_scales_ratio, weights, cfg)

    for _ in range(3):
        rec = ours()
        out = ref()
    torch.cuda.synchronize()
    # same answer first (fp32 reference vs bf16-operand head: centre within 1e-2 of the scene scale)
    got = rec[:, ops.REC_CENTER_CAM].cpu()
    want = torch.cat(out.pred_center_cam).cpu()
    assert float((got - want).abs().max() / want.abs().max()) <= 1e-2
    ours_ms, ours_min = _time_ms(ours, 20, flush)
    ref_ms, ref_min = _time_ms(ref, 10, flush)
    if ref_ms / ours_ms < TARGET_SPEEDUP:      # one re-measurement: the reference arm is host-launch-bound and noisy
        ours_ms, ours_min = _time_ms(ours, 30, flush)
        ref_ms, ref_min = _time_ms(ref, 15, flush)
    n = db.n
    res = {"workload": bench.WORKLOAD, "boxes": n, "ours_ms_median": ours_ms, "ours_ms_min": ours_min,
           "reference_gpu_pytorch_ms_median": ref_ms, "reference_gpu_pytorch_ms_min": ref_min,
           "ours_boxes_per_s": n / ours_ms * 1e3, "reference_gpu_pytorch_boxes_per_s": n / ref_ms * 1e3,
           "speedup_median": ref_ms / ours_ms,
           "what": "device-resident inputs, 256 MiB L2 flush before each call, CUDA events around the call; the "
                   "reference arm is oracle/ on cuda (torchvision CUDA roi_align, fp32 cuBLAS, eager decode)"}
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "ref_gpu_path.json"), "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res))
    assert res["speedup_median"] >= TARGET_SPEEDUP, res


def _run_perf(module_name, argv, monkeypatch):
    """tests/perf_*.py are runnable scripts (they live under tests/ because their reference leg executes oracle/);
    here each runs briefly, the JSON goes to gpurun_out/ and a conservative floor is asserted."""
    import importlib
    import sys

    mod = importlib.import_module("tests." + module_name)
    monkeypatch.setattr(sys, "argv", [module_name + ".py"] + argv)
    res = mod.main()
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", module_name + ".json"), "w") as f:
        json.dump(res, f, indent=1)
    return res


@pytest.mark.gpu
def test_post_processing_speed_against_reference_sequence(monkeypatch):
    res = _run_perf("perf_gdino_post", ["--iters", "30"], monkeypatch)
    assert res["gdino_post"]["speedup_vs_reference_sequence"] >= POST_TARGET, res
    assert res["nms_16x1000"]["speedup"] >= POST_TARGET, res


@pytest.mark.gpu
def test_box_inference_speed_against_reference_sequence(monkeypatch):
    res = _run_perf("perf_box_head", ["--iters", "30"], monkeypatch)
    assert res["box_inference"]["same_classes_as_reference_gpu_sequence"], res
    assert res["box_inference"]["speedup"] >= POST_TARGET, res


@pytest.mark.gpu
def test_open_vocabulary_flow_speed_against_reference_sequence(monkeypatch):
    res = _run_perf("perf_gdino_flow", ["--iters", "30"], monkeypatch)
    assert res["speedup"] >= POST_TARGET, res
