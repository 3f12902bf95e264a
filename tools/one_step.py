"""Two C2 steps of the lift on cuda:0 (device-resident inputs), finite-checked: the target of compute-sanitizer runs
(tools/sanitize.sh) and of single-kernel ncu captures.

    python tools/one_step.py [--boxes 100] [--images 16] [--steps 2]
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--boxes", type=int, default=100)
    ap.add_argument("--images", type=int, default=16)
    ap.add_argument("--steps", type=int, default=2)
    args = ap.parse_args()
    import bench
    from ovmono3d_b200 import ops, synth

    dev = torch.device("cuda:0")
    ops.load_library()
    head, names = bench.build_head(dev)
    inp = synth.make_inputs(1235, args.images, args.boxes, strides=bench.STRIDES, canvas=bench.CANVAS,
                            channels=bench.CHANNELS)
    db = bench.DeviceBatch(inp, dev)
    eng = head._get_engine()
    eng.prepare(dev, head.priors_dims_per_cat, head.priors_z_scales, getattr(head, "priors_z_stats", None))
    before = ops.launch_count()
    for _ in range(args.steps):
        rec = eng.forward(db.features, db.boxes, db.classes, db.scores, db.batch_idx, db.cams)[0]
    torch.cuda.synchronize()
    assert torch.isfinite(rec).all()
    print("one_step ok: %d boxes, %d launches, checksum %.6f" % (db.n, ops.launch_count() - before,
                                                                float(rec.double().abs().sum())))


if __name__ == "__main__":
    main()
