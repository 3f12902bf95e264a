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


if __name__ == "__main__":
    main()
