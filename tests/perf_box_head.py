"""Time the 2-D box branch's inference tail (`ovm_box_inference`: softmax, apply_deltas, clip, threshold, per-class
NMS, top-k) and the whole branch next to the reference's own call sequence run on the GPU through
oracle/box_head_oracle.py (torch CUDA ops + torchvision CUDA batched_nms, fast_rcnn.py:57-116).  One JSON line.

    python tests/perf_box_head.py [--images 2] [--proposals 1000] [--iters 100]
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from oracle import box_head_oracle as bo            # noqa: E402  (baseline leg only)
from ovmono3d_b200 import ops                       # noqa: E402
from tests.golden import box_cases as bc            # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--images", type=int, default=2)
    ap.add_argument("--proposals", type=int, default=1000)
    ap.add_argument("--iters", type=int, default=100)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    case = bc.BoxCase("bench", 9, (a.proposals,) * a.images, 50, ((896, 1184),) * a.images, logit_scale=4.0)
    logits, deltas, props = bc.build_case(case)
    counts = [len(p) for p in props]
    off = torch.tensor([0] + list(np.cumsum(counts)), dtype=torch.int32, device=dev)
    sizes = torch.tensor([list(s) for s in case.image_sizes], dtype=torch.int32, device=dev)
    lg, dl, pr = logits.to(dev), deltas.to(dev), torch.cat(props).to(dev)

    def ours():
        return ops.box_inference(lg, dl, pr, off, sizes, max(counts), 50, bc.REG_WEIGHTS, case.score_thresh,
                                 case.nms_thresh, case.topk, False, 5000)

    for _ in range(5):
        out = ours()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.iters):
        out = ours()
    e1.record()
    torch.cuda.synchronize()
    ours_us = e0.elapsed_time(e1) * 1e3 / a.iters
    props_d = [p.to(dev) for p in props]

    def ref():
        boxes, probs = bo.predict(lg, dl, props_d, bc.REG_WEIGHTS)
        return bo.fast_rcnn_inference(boxes, probs, list(case.image_sizes), case.score_thresh, case.nms_thresh, case.topk)

    for _ in range(3):
        want = ref()
    torch.cuda.synchronize()
    n_ref = max(5, a.iters // 10)
    t0 = time.perf_counter()
    for _ in range(n_ref):
        want = ref()
    torch.cuda.synchronize()
    ref_us = (time.perf_counter() - t0) * 1e6 / n_ref
    same = all(int(out[5][i]) == len(w[1]) and out[2][i, :len(w[1])].tolist() == w[3].tolist()
               for i, w in enumerate(want))
    cands = [int((torch.softmax(lg[o:o + c], -1)[:, :-1] > case.score_thresh).sum()) for o, c in zip(off.tolist(), counts)]
    res = {"box_inference": {"images": a.images, "proposals_per_image": a.proposals, "classes": 50,
                                        "candidates_per_image": cands, "ours_us": ours_us,
                                        "reference_gpu_sequence_us": ref_us, "speedup": ref_us / ours_us,
                                        "same_classes_as_reference_gpu_sequence": same}}
    print(json.dumps(res))
    return res


if __name__ == "__main__":
    main()
