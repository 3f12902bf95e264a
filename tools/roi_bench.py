"""Time `ovm_roi_align` alone on the C2 workload under the development knobs of csrc/roi_align.cu.

    python tools/roi_bench.py                    # sweep (one subprocess per knob setting, knobs are read once)
    python tools/roi_bench.py --one              # time the current environment's setting

Conditions match bench.py: three rotating 352 MB pyramids, 256 MiB L2 flush before every launch (outside the event
pair), bf16 output, CUDA events on the launching stream.
"""
import itertools
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
KNOBS = ("OVM_ROI_SPLIT", "OVM_ROI_PROMO", "OVM_ROI_ALIGN8", "OVM_ROI_STAGES")


def one(iters=30):
    import numpy as np
    import torch

    import bench
    from ovmono3d_b200 import ops

    ops.load_library()
    dev = torch.device("cuda:0")
    sets = []
    for i in range(3):
        inp = bench.make_batch(1235 + i)
        counts = [len(b) for b in inp.boxes]
        sets.append(([f.to(dev) for f in inp.features], torch.cat(inp.boxes).to(dev),
                     torch.from_numpy(np.repeat(np.arange(len(counts), dtype=np.int32), counts)).to(dev)))
    scales = [1.0 / s for s in bench.STRIDES]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ts = []
    for it in range(iters + 5):
        f, b, bi = sets[it % 3]
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out, lv = ops.roi_align(f, scales, b, bi, 7, 0, torch.bfloat16)
        e1.record()
        torch.cuda.synchronize()
        if it >= 5:
            ts.append(e0.elapsed_time(e1) * 1e3)
    f, b, bi = sets[0]
    alg = bench.roi_align_algorithmic_bytes(b.cpu().numpy(), ops.roi_align(f, scales, b, bi, 7, 0, torch.bfloat16)[1]
                                            .cpu().numpy(), 2)
    ts = np.array(ts)
    print(json.dumps({"knobs": {k: os.environ.get(k) for k in KNOBS}, "us_mean": float(ts.mean()),
                      "us_min": float(ts.min()), "us_median": float(np.median(ts)),
                      "GBps_at_mean": alg / ts.mean() / 1e3}))


def sweep():
    grid = {"OVM_ROI_SPLIT": ["400", "100000"], "OVM_ROI_PROMO": ["1", "2"], "OVM_ROI_ALIGN8": ["0", "1"],
            "OVM_ROI_STAGES": ["2"]}
    if len(sys.argv) > 1 and sys.argv[1] == "--full":
        grid["OVM_ROI_PROMO"] = ["0", "1", "2", "3"]
    for combo in itertools.product(*grid.values()):
        env = dict(os.environ)
        env.update(dict(zip(grid.keys(), combo)))
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--one"], env=env, capture_output=True, text=True)
        line = [l for l in r.stdout.splitlines() if l.startswith("{")]
        print(line[-1] if line else "FAILED %s: %s" % (combo, r.stderr[-400:]), flush=True)


if __name__ == "__main__":
    if "--one" in sys.argv:
        one()
    else:
        sweep()
