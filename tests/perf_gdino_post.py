"""Time `ovm_gdino_post` / `ovm_nms` on the GPU next to the reference's own sequence of calls
(roi_heads_gdino.py:186-202, 269-275: CUDA sigmoid / gathers / max, three `.cpu()` copies, torchvision nms on the
host) run through oracle/gdino_post_oracle.py on CUDA tensors.  Prints one JSON line.

    python tests/perf_gdino_post.py [--queries 900] [--iters 200]
"""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from oracle import gdino_post_oracle as go          # noqa: E402  (baseline leg only)
from ovmono3d_b200 import gdino_post as gp, ops     # noqa: E402
from tests.golden import gdino_cases as gd          # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--queries", type=int, default=900)
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--iters", type=int, default=200)
    ap.add_argument("--trace", action="store_true", help="per-phase %%globaltimer stamps of one call (ovm_post_set_trace)")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    case = gd.GdinoCase("bench", 5, a.queries, 256, gd.OMNI, (896, 1184))
    logits, boxes, tok, caption, prompt = gd.build_case(case)
    cap_list = [c[0] for c in prompt]
    ids = go.phrase_token_ids(tok(caption)["input_ids"], tok(cap_list, add_special_tokens=False)["input_ids"])
    ranges = gp.phrase_token_ranges(tok, cap_list, caption).to(dev)
    lg = logits.repeat(a.batch, 1, 1).to(dev)
    bx = boxes.repeat(a.batch, 1, 1).to(dev)
    for _ in range(5):
        out = ops.gdino_post(lg, bx, ranges, case.image_hw, 0.001, 0.5)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.iters):
        out = ops.gdino_post(lg, bx, ranges, case.image_hw, 0.001, 0.5)
    e1.record()
    torch.cuda.synchronize()
    ours_us = e0.elapsed_time(e1) * 1e3 / a.iters
    kept = int(out[4][0])
    # with the one host read the plugin does
    t0 = time.perf_counter()
    for _ in range(a.iters):
        out = ops.gdino_post(lg, bx, ranges, case.image_hw, 0.001, 0.5)
        out[4].tolist()
    ours_sync_us = (time.perf_counter() - t0) * 1e6 / a.iters
    for _ in range(3):
        ref = go.gdino_post(lg[0], bx[0], ids, case.image_hw, 0.001, 0.5)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n_ref = max(10, a.iters // 10)
    for _ in range(n_ref):
        for b in range(a.batch):
            ref = go.gdino_post(lg[b], bx[b], ids, case.image_hw, 0.001, 0.5)
    torch.cuda.synchronize()
    ref_us = (time.perf_counter() - t0) * 1e6 / n_ref
    assert kept == len(ref["scores"])
    phases = None
    if a.trace:
        buf = torch.zeros(a.batch * 8, dtype=torch.int64, device=dev)
        lib = ops.load_library()
        lib.ovm_post_set_trace(buf.data_ptr())
        ops.gdino_post(lg, bx, ranges, case.image_hw, 0.001, 0.5)
        torch.cuda.synchronize()
        lib.ovm_post_set_trace(None)
        st = buf.view(a.batch, 8)[0].tolist()
        names = ["load query scores", "sort", "box staging", "suppression", "gather"]
        phases = {nm: (st[i + 1] - st[i]) / 1e3 for i, nm in enumerate(names)}
        phases["suppression_cycles_step1_pair_tests"] = st[6] >> 32
        phases["suppression_cycles_step2_serial_resolve"] = st[6] & 0xFFFFFFFF
        phases["suppression_cycles_step3_later_positions"] = st[7] & 0xFFFFFFFF
        phases["suppression_cycles_step3_of_which_cluster_barrier"] = st[7] >> 32
    # NMS alone, 16 images x 1000 boxes per launch
    g = torch.Generator().manual_seed(1)
    n, S = 1000, 16
    c = torch.rand(S * n, 2, generator=g) * 800
    s = torch.exp(torch.rand(S * n, 2, generator=g) * 3 + 2)
    nb = torch.cat([c - s / 2, c + s / 2], 1).to(dev)
    ns = torch.rand(S * n, generator=g).to(dev)
    off = torch.arange(0, S * n + 1, n, dtype=torch.int32, device=dev)
    for _ in range(3):
        ops.nms(nb, ns, 0.5, seg_offsets=off, max_segment_len=n)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(a.iters):
        k, cnt = ops.nms(nb, ns, 0.5, seg_offsets=off, max_segment_len=n)
    e1.record()
    torch.cuda.synchronize()
    nms_us = e0.elapsed_time(e1) * 1e3 / a.iters
    from torchvision.ops import nms as tv_nms
    for _ in range(3):
        tv_nms(nb[:n], ns[:n], 0.5)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n_ref):
        for i in range(S):
            tv_nms(nb[i * n:(i + 1) * n], ns[i * n:(i + 1) * n], 0.5)
    torch.cuda.synchronize()
    tv_us = (time.perf_counter() - t0) * 1e6 / n_ref
    res = {
        "gdino_post": {"queries": a.queries, "batch": a.batch, "phrases": len(cap_list), "kept": kept,
                       "ours_us_device": ours_us, "ours_us_with_count_readback": ours_sync_us,
                       "reference_sequence_us": ref_us, "speedup_vs_reference_sequence": ref_us / ours_sync_us, "phase_us": phases},
        "nms_16x1000": {"ours_us": nms_us, "kept_mean": float(cnt.float().mean()),
                        "torchvision_cuda_nms_16_calls_us": tv_us, "speedup": tv_us / nms_us}}
    print(json.dumps(res))
    return res


if __name__ == "__main__":
    main()
