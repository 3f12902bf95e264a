"""cProfile of the plugin call `_forward_cube` on the C2 workload with host-resident Instances (run on a B200)."""
import cProfile
import os
import pstats
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from ovmono3d_b200.compat import Boxes, Instances  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    head, names = bench.build_head(dev)
    hb = bench.HostBatch(bench.make_batch(1235))
    feats = dict(zip(names, [f.to(dev) for f in hb.features]))

    def step():
        inst = []
        for i in range(len(hb.boxes)):
            x = Instances(hb.im_dims[i])
            x.pred_boxes = Boxes(hb.boxes[i])
            x.pred_classes = hb.classes[i]
            x.scores = hb.scores[i]
            inst.append(x)
        head._forward_cube(feats, inst, hb.Ks, hb.im_dims, hb.ratios)
        torch.cuda.current_stream().synchronize()

    for _ in range(20):
        step()
    if "--timeline" in sys.argv:
        # wall-clock marks inside one call (no profiler): where the 0.30 ms of e2e_resident_features goes
        import time
        eng = head._get_engine()
        marks = {}

        def wrap(name):
            real = getattr(eng, name)

            def f(*a, **k):
                marks[name + "_in"] = time.perf_counter()
                r = real(*a, **k)
                marks[name + "_out"] = time.perf_counter()
                return r

            setattr(eng, name, f)

        wrap("begin")
        wrap("finish")
        rec_host = torch.empty((hb.n, 48), dtype=torch.float32).pin_memory()
        rows = []
        for _ in range(300):
            inst = []
            for i in range(len(hb.boxes)):
                x = Instances(hb.im_dims[i])
                x.pred_boxes = Boxes(hb.boxes[i])
                x.pred_classes = hb.classes[i]
                x.scores = hb.scores[i]
                inst.append(x)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter()
            e0.record()
            head._forward_cube(feats, inst, hb.Ks, hb.im_dims, hb.ratios)
            t1 = time.perf_counter()
            e1.record()
            rec_host.copy_(head.last_records, non_blocking=True)
            torch.cuda.current_stream().synchronize()
            t2 = time.perf_counter()
            rows.append(((marks["begin_in"] - t0) * 1e6, (marks["begin_out"] - marks["begin_in"]) * 1e6,
                         (marks["finish_in"] - marks["begin_out"]) * 1e6, (marks["finish_out"] - marks["finish_in"]) * 1e6,
                         (t1 - marks["finish_out"]) * 1e6, (t2 - t1) * 1e6, (t2 - t0) * 1e6, e0.elapsed_time(e1) * 1e3))
        import numpy as np
        m = np.median(np.array(rows[50:]), axis=0)
        print("median us: box staging + H2D %.1f | engine.begin (plan + pool launches) %.1f | camera / score / class staging + "
              "H2D %.1f | engine.finish (FC + decode launches) %.1f | write-back %.1f | D2H + sync wait %.1f | total %.1f | "
              "device events around the call %.1f" % tuple(m))
        return
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(200):
        step()
    pr.disable()
    pstats.Stats(pr).sort_stats("cumulative").print_stats(28)


if __name__ == "__main__":
    main()
