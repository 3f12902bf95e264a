"""Per-CTA timeline of the heads + decode kernel inside the C2 lift (ovm_head_set_trace).  Run on a B200."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from ovmono3d_b200 import ops  # noqa: E402


def main():
    lib = ops.load_library()
    dev = torch.device("cuda:0")
    head, names = bench.build_head(dev)
    db = bench.DeviceBatch(bench.make_batch(1235), dev)
    eng = head._get_engine()
    eng.prepare(dev, head.priors_dims_per_cat, head.priors_z_scales, getattr(head, "priors_z_stats", None))
    trace = torch.zeros(4096 * 8, dtype=torch.int64, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for it in range(4):
        flush.zero_()
        if it == 3:
            lib.ovm_head_set_trace(trace.data_ptr())
        eng.forward(db.features, db.boxes, db.classes, db.scores, db.batch_idx, db.cams)
    lib.ovm_head_set_trace(None)
    torch.cuda.synchronize()
    t = trace.cpu().numpy().reshape(-1, 8).astype(np.float64)
    t = t[t[:, 0] > 0]
    t0 = t[:, 0].min()
    if eng.tail_fused:
        names = ["entry", "set-up + staging done", "previous FC layer complete", "first operand stage landed",
                 "accumulator complete", "epilogue + head partials", "partials exchanged", "records written"]
    else:
        names = ["entry", "rows staged", "FC features ready", "box/camera loads", "dot products", "record written"]
    last = len(names) - 1
    print("%s: CTAs %d; span %.1f us (first entry -> last record)" % (
        "fc_tail_kernel (%s)" % eng.tail_precision if eng.tail_fused else "head_decode_kernel", len(t),
        (t[:, last].max() - t0) / 1e3))
    for j, nm in enumerate(names):
        v = (t[:, j] - t0) / 1e3
        print("  %-28s min %6.2f  median %6.2f  max %6.2f us" % (nm, v.min(), np.median(v), v.max()))
    d = np.diff(t[:, :len(names)], axis=1) / 1e3
    print("  per-CTA phase durations (median): " + ", ".join("%s %.2f" % (names[j + 1], np.median(d[:, j])) for j in range(d.shape[1])))


if __name__ == "__main__":
    main()
