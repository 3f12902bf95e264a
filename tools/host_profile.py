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
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(200):
        step()
    pr.disable()
    pstats.Stats(pr).sort_stats("cumulative").print_stats(28)


if __name__ == "__main__":
    main()
