"""Debug helper: one box per call on a single-level map, footprint widths chosen to hit each TMA tile width."""
import os
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def one(width, height, C):
    import torch
    from ovmono3d_b200 import ops
    dev = torch.device("cuda:0")
    f = torch.randn(1, C, 64, 64, device=dev)
    boxes = torch.tensor([[8.0, 8.0, 8.0 + width, 8.0 + height]], device=dev)
    bi = torch.zeros(1, dtype=torch.int32, device=dev)
    out, lv = ops.roi_align([f], [1.0], boxes, bi)
    torch.cuda.synchronize()
    import torchvision
    ref = torchvision.ops.roi_align(f, torch.cat([torch.zeros(1, 1, device=dev), boxes], 1), 7, 1.0, 0, True)
    print("w=%g h=%g C=%d ok, max err %.3g" % (width, height, C, float((out - ref).abs().max())))


if __name__ == "__main__":
    if len(sys.argv) > 1:
        one(float(sys.argv[1]), float(sys.argv[2]), int(sys.argv[3]))
    else:
        for w, h, C in [(4, 4, 32), (12, 4, 32), (28, 4, 32), (4, 20, 32), (12, 12, 64), (4, 4, 16), (40, 30, 256)]:
            r = subprocess.run([sys.executable, __file__, str(w), str(h), str(C)], capture_output=True, text=True,
                               timeout=120)
            print((r.stdout.strip() or r.stderr.strip().splitlines()[-1])[:200])
