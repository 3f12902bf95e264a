"""Phase timeline of the tcgen05 GEMM (ovm_gemm_set_trace): where a CTA's time goes, for the lift's two shapes.

    python tools/gemm_trace.py            # on a B200
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ovmono3d_b200 import ops  # noqa: E402

NAMES = ["entry", "setup", "last_tma", "first_tile", "last_mma", "acc_ready", "epi_done", "pre_release"] + \
    ["chunk%d" % i for i in range(8)]


def main():
    lib = ops.load_library()
    dev = torch.device("cuda:0")
    for (M, N, K, mode) in [(1600, 1024, 12544, ops.GEMM_STREAMK), (1600, 1024, 1024, ops.GEMM_DIRECT),
                            (1600, 1024, 12544, ops.GEMM_DIRECT)]:
        A = torch.randn(M, K, device=dev).bfloat16()
        B = torch.randn(N, K, device=dev).bfloat16()
        bias = torch.randn(N, device=dev)
        trace = torch.zeros(4096 * 16, dtype=torch.int64, device=dev)
        for it in range(3):
            if it == 2:
                lib.ovm_gemm_set_trace(trace.data_ptr())
            if mode == ops.GEMM_STREAMK:
                ops.gemm_bf16(A, B, None, False, torch.float32, mode)
            else:
                ops.gemm_bf16(A, B, bias, True, torch.float32, mode)
        lib.ovm_gemm_set_trace(None)
        torch.cuda.synchronize()
        t = trace.cpu().numpy().reshape(-1, 16)
        t = t[t[:, 0] > 0].astype(np.float64)
        t0 = t[:, 0].min()
        rel = (t - t0) / 1e3
        print("M=%d N=%d K=%d mode=%d ctas=%d  (us since first CTA entry: min / median / max over CTAs)" % (
            M, N, K, mode, len(t)))
        for k, nm in enumerate(NAMES):
            col = rel[:, k][t[:, k] > 0]
            if len(col):
                print("  %-12s %8.2f %8.2f %8.2f" % (nm, col.min(), np.median(col), col.max()))


def pipeline():
    """fc1 / fc2 as they run inside the lift at C2 (after ROIAlign, 256 MiB L2 flush before the step): the pair kernel's
    stamps are entry, set-up done, last TMA, wait on the previous kernel over, last MMA issued, (last) accumulator ready,
    epilogue + finish done, (last) partial tile added."""
    import bench
    lib = ops.load_library()
    dev = torch.device("cuda:0")
    head, _ = bench.build_head(dev)
    db = bench.DeviceBatch(bench.make_batch(1235), dev)
    eng = head._get_engine()
    eng.prepare(dev, head.priors_dims_per_cat, head.priors_z_scales, getattr(head, "priors_z_stats", None))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    trace = torch.zeros(4096 * 16, dtype=torch.int64, device=dev)
    for it in range(4):
        flush.zero_()
        if it == 3:
            lib.ovm_gemm_set_trace(trace.data_ptr())
        eng.forward(db.features, db.boxes, db.classes, db.scores, db.batch_idx, db.cams)
    lib.ovm_gemm_set_trace(None)
    torch.cuda.synchronize()
    t = trace.cpu().numpy().reshape(-1, 16)
    if not eng.tail_fused:
        t = t[104:148]    # fc2 (104 CTAs) overwrote rows 0..103: fc1's CTAs 104..147
    t = t[t[:, 0] > 0].astype(np.float64)
    names = ["entry", "setup", "last_tma", "prev_done", "last_mma", "acc_ready", "epi_done", "partial_added"]
    t0 = t[:, 0].min()
    print("fc1 in the lift (cold L2), CTAs 104..147, us since their first entry: min / median / max")
    for k, nm in enumerate(names):
        col = t[:, k][t[:, k] > 0]
        if len(col):
            col = (col - t0) / 1e3
            print("  %-14s %8.2f %8.2f %8.2f" % (nm, col.min(), np.median(col), col.max()))
    d = lambda a, b: (t[:, a] - t[:, b]) / 1e3
    for nm, v in (("main loop (prev_done -> last_mma)", d(4, 3)[t[:, 4] > 0]), ("drain (last_mma -> acc_ready)", d(5, 4)[t[:, 4] > 0]),
                  ("partial add (acc_ready -> added)", d(7, 5)), ("finish (added -> epi_done)", d(6, 7))):
        print("  %-36s min %6.2f  median %6.2f  p90 %6.2f  max %6.2f   sorted: %s" % (
            nm, v.min(), np.median(v), np.percentile(v, 90), v.max(), " ".join("%.1f" % x for x in np.sort(v)[::4])))
    own = t[:, 8] > 0       # slot exchange (gemm_pair_kernel<true>): CTAs whose last segment owns a tile
    if own.any():
        o = t[own]
        print("  owners of a tile (last segment): %d CTAs; partial sources per tile %s" % (
            len(o), np.bincount(o[:, 11].astype(int)).tolist()))
        for nm, v in (("acc_ready -> all partials published", (o[:, 8] - o[:, 5]) / 1e3),
                      ("published -> first chunk merged", (o[:, 9] - o[:, 8]) / 1e3),
                      ("first chunk -> last chunk merged", (o[:, 10] - o[:, 9]) / 1e3)):
            v = v[np.isfinite(v)]
            print("    %-38s min %6.2f  median %6.2f  p90 %6.2f  max %6.2f" % (nm, v.min(), np.median(v), np.percentile(v, 90), v.max()))


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "pipeline":
        pipeline()
    else:
        main()
