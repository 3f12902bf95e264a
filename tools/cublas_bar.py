"""The Blackwell library bar for the cube-head GEMMs (SURVEY.md section 2a row 2, section 8d baseline iii; VERDICT r1
"missing" 2): torch.matmul (cuBLAS / cuBLASLt, bf16 operands, fp32 accumulate) on the three FC shapes of CubeHead
(reference cubercnn/modeling/roi_heads/cube_head.py:157-164), next to `ovm_gemm_bf16` on the same operands.

    python tools/cublas_bar.py [--out gpurun_out/cublas_bar.json] [--iters 30]

Every call is bracketed by its own CUDA-event pair; a 256 MiB memset flushes the L2 before each call (outside the
pair), so operands come from HBM as they do for the first layer of a step.  `--warm` additionally times back-to-back
calls without the flush (operands L2-resident where they fit), which is the state fc2 sees inside the lift.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def time_call(fn, iters, flush):
    ts = []
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return float(np.median(ts)), float(min(ts))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "cublas_bar.json"))
    ap.add_argument("--iters", type=int, default=30)
    args = ap.parse_args()
    from ovmono3d_b200 import ops

    ops.load_library()
    dev = torch.device("cuda:0")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    gen = torch.Generator(device="cpu").manual_seed(5)
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(
        os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    rows = []
    for n in (80, 1600, 16000):
        for name, K, N in (("fc1", 12544, 1024), ("fc2", 1024, 1024), ("heads", 1024, 16)):
            A = (torch.randn(n, K, generator=gen) * 0.5).to(dev).bfloat16()
            B = (torch.randn(N, K, generator=gen) * 0.02).to(dev).bfloat16()
            Bt = B.t()
            out_cublas = torch.empty(n, N, device=dev, dtype=torch.bfloat16)
            out_f32 = torch.empty(n, N, device=dev, dtype=torch.float32)

            def cublas_bf16_out():
                torch.matmul(A, Bt, out=out_cublas)

            ours_out = torch.zeros(n, (N + 7) // 8 * 8, device=dev, dtype=torch.float32)

            def ours_direct():
                ops.gemm_bf16(A, B, out=ours_out[:, :N], mode=ops.GEMM_DIRECT)

            def ours_auto():
                ours_out.zero_()
                ops.gemm_bf16(A, B, out=ours_out[:, :N], mode=ops.GEMM_AUTO)

            flops = 2.0 * n * K * N
            r = {"layer": name, "M": n, "N": N, "K": K, "gflop": flops / 1e9}
            for label, fn in (("cublas_bf16", cublas_bf16_out), ("ovm_gemm_direct", ours_direct),
                              ("ovm_gemm_auto_incl_memset", ours_auto)):
                med, mn = time_call(fn, args.iters, flush)
                r[label + "_us_flushed"] = med
                r[label + "_tflops_flushed"] = flops / med / 1e6
                medw, mnw = time_call(fn, args.iters, None)
                r[label + "_us_warm"] = medw
                r[label + "_tflops_warm"] = flops / medw / 1e6
            # same answer (fp32 out of ours vs bf16 out of cuBLAS: one bf16 rounding apart)
            ours_direct()
            cublas_bf16_out()
            torch.cuda.synchronize()
            ref = out_cublas.float()
            r["max_abs_diff_vs_cublas"] = float((ours_out[:, :N] - ref).abs().max())
            r["ref_scale"] = float(ref.abs().max())
            rows.append(r)
            print(json.dumps(r))
    res = {"what": "torch.matmul bf16 (cuBLAS) vs ovm_gemm_bf16 on CubeHead's FC shapes; per-call CUDA events; "
                   "'flushed' = 256 MiB memset before every call, 'warm' = back to back",
           "torch": torch.__version__, "gpu": torch.cuda.get_device_name(0), "peaks": peaks, "rows": rows}
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    with open(args.out, "w") as f:
        json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
