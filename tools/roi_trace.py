"""Per-CTA timeline of the ROIAlign kernel on the C2 workload (ovm_roi_set_trace).

    python tools/roi_trace.py            # on a B200
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ovmono3d_b200 import ops, synth  # noqa: E402


def main():
    lib = ops.load_library()
    dev = torch.device("cuda:0")
    # optional: strides (comma list), canvas, images, boxes per image   e.g.  4,8,16,32 1024 16 100
    argv = [a for a in sys.argv if not a.startswith("--")]
    strides = tuple(int(v) for v in argv[1].split(",")) if len(argv) > 1 else (7, 14, 28)
    canvas = int(argv[2]) if len(argv) > 2 else 896
    N = int(argv[3]) if len(argv) > 3 else 16
    k = int(argv[4]) if len(argv) > 4 else 100
    inp = synth.make_inputs(1235 if len(argv) == 1 else 4321, N, k, strides=strides, canvas=canvas)
    feats = [f.to(dev) for f in inp.features]
    if os.environ.get("OVM_TRACE_LAYOUT", "nchw") == "channels_last":
        feats = [f.contiguous(memory_format=torch.channels_last) for f in feats]
        print("pyramid: torch.channels_last")
    counts = [len(b) for b in inp.boxes]
    boxes = torch.cat(inp.boxes).to(dev)
    bidx = torch.from_numpy(np.repeat(np.arange(N, dtype=np.int32), counts)).to(dev)
    scales = [1.0 / v for v in strides]
    trace = torch.zeros(65536 * 16, dtype=torch.int64, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for it in range(3):
        flush.zero_()
        if it == 2:
            lib.ovm_roi_set_trace(trace.data_ptr())
        ops.roi_align(feats, scales, boxes, bidx, 7, 0, torch.bfloat16)
    lib.ovm_roi_set_trace(None)
    torch.cuda.synchronize()
    t = trace.cpu().numpy().reshape(-1, 16)
    if "--plan" in sys.argv:
        pl = t[65000:65000 + 512].astype(np.float64)
        pl = pl[pl[:, 0] > 0]
        p0 = pl[:, 0].min()
        names = ["entry", "after the wait", "boxes planned", "last CTA elected", "mean known", "histogram", "scan", "items written"]
        print("roi_plan_kernel: %d CTAs; first ROIAlign CTA enters %.2f us after the planner's first" % (
            len(pl), (t[:60000][t[:60000, 0] > 0][:, 0].min() - p0) / 1e3))
        for j, nm in enumerate(names):
            v = pl[:, j][pl[:, j] > 0]
            print("  %-18s min %6.2f  median %6.2f  max %6.2f us  (%d CTAs)" % (nm, (v.min() - p0) / 1e3, (np.median(v) - p0) / 1e3,
                                                                            (v.max() - p0) / 1e3, len(v)))
    t = t[:60000]
    t = t[t[:, 0] > 0]
    t0 = t[:, 0].min()
    tma = t[:, 12] > 0
    print("CTAs %d, on the TMA path %d; kernel span %.1f us" % (len(t), tma.sum(), (t[:, :13].max() - t0) / 1e3))
    if (~tma).any():
        dd = t[~tma]
        print("  direct-path / early-exit CTAs: %d (they stamp only slot 0)" % len(dd))
    tt = t[tma].astype(np.float64)
    d = lambda a, b: (tt[:, b] - tt[:, a]) / 1e3
    for name, v in [("entry -> bin ranges", d(0, 1)), ("bin ranges -> tables", d(1, 2)),
                    ("tables -> first stage landed", d(2, 3)), ("first stage -> chunk 0 written", d(3, 4)),
                    ("chunk 0 -> chunk 7", d(4, 11)), ("whole CTA", d(0, 12))]:
        print("  %-32s mean %6.2f  median %6.2f  p90 %6.2f  max %6.2f us" % (name, v.mean(), np.median(v),
                                                                        np.percentile(v, 90), v.max()))
    print("  consumer warp 0 waiting for data: mean %.2f us/CTA (%.0f%% of the chunk loop); producer waiting for a free "
          "stage: mean %.2f us/CTA" % (tt[:, 13].mean() / 1965.0, 100 * tt[:, 13].sum() / 1965.0 / max(d(3, 12).sum(), 1e-9),
                                        tt[:, 14].mean() / 1965.0))
    packed = t[tma][:, 15]
    smid = packed & 0xFFFFFF
    rows, fw, wsel = packed >> 48, (packed >> 32) & 0xFFFF, (packed >> 24) & 0xFF
    print("  tile width index histogram", np.bincount(wsel.astype(int), minlength=3), " rows mean %.1f  fw mean %.1f" % (
        rows.mean(), fw.mean()))
    busy = np.zeros(148)
    for s_, a_, b_ in zip(smid, tt[:, 0], tt[:, 12]):
        busy[int(s_) % 148] += (b_ - a_) / 1e3
    print("  per-SM sum of CTA lifetimes: mean %.1f  max %.1f us (2 CTAs run side by side)" % (busy.mean(), busy.max()))
    start_rel = (tt[:, 0] - t0) / 1e3
    print("  last CTA starts at %.1f us; CTAs starting after 80%% of the span: %d" % (
        start_rel.max(), (start_rel > 0.8 * (t[:, :13].max() - t0) / 1e3).sum()))


if __name__ == "__main__":
    main()
