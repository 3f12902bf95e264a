"""The synthetic workloads of BASELINE.json `configs`, by name.

One definition shared by `bench.py --config`, `tools/sweep.py`, `tools/stream_c4.py` and the full-size parity tests, so
that what is timed is what is checked.  Pyramid shapes follow the reference's configs: DINOv2 simple feature pyramid
p2-p4 at strides 7/14/28 on an 896-px canvas (configs/OVMono3D_dinov2_SFP.yaml:37-44, cubercnn/modeling/backbone/
dino.py:143); the patch-16 variants (clip / mae / sam / midas) p2-p5 at strides 4/8/16/32 on 1024 px
(configs/OVMono3D_clip_SFP.yaml:34-44, clip.py:155); DLA-FPN p2-p6 at strides 4..64 (configs/Base.yaml:44,63,
dla.py:460-461).  256 channels everywhere in the shipped configs; 128 / 512 exercise the change of fc1's K = 49 C.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Sequence, Tuple


@dataclass(frozen=True)
class Workload:
    name: str
    what: str
    strides: Tuple[int, ...] = (7, 14, 28)
    canvas: int = 896
    channels: int = 256
    num_images: int = 16
    boxes_per_image: int = 100
    content: int = 0            # side of the image content inside the canvas (0 = the whole canvas)
    seed: int = 1235
    fc_dim: int = 1024

    @property
    def boxes(self) -> int:
        return self.num_images * self.boxes_per_image

    @property
    def in_features(self) -> Sequence[str]:
        return ["p%d" % (2 + i) for i in range(len(self.strides))]

    def describe(self) -> str:
        return ("%s: %s, %d images x %d boxes per rank-step, %d-level pyramid %dch @%s (%dpx, strides %s) fp32 NCHW, "
                "head %d-%d-%d-13 bf16 operands / fp32 accumulate" % (
                    self.name, self.what, self.num_images, self.boxes_per_image, len(self.strides), self.channels,
                    "/".join(str(self.canvas // s) for s in self.strides), self.canvas,
                    "/".join(str(s) for s in self.strides), self.channels * 49, self.fc_dim, self.fc_dim))

    def make_inputs(self, seed_offset: int = 0):
        from . import synth

        return synth.make_inputs(self.seed + seed_offset, self.num_images, self.boxes_per_image, strides=self.strides,
                                 canvas=self.canvas, content=self.content or None, channels=self.channels)

    def make_weights(self):
        from . import synth

        return synth.make_head_weights(77, in_dim=self.channels * 49, fc_dim=self.fc_dim)


_SFP4 = dict(strides=(4, 8, 16, 32), canvas=1024)
WORKLOADS: Dict[str, Workload] = {w.name: w for w in [
    Workload("C1", "OVMono3D_dinov2_SFP lift, 512-px content (BASELINE configs[0])", num_images=4, boxes_per_image=20,
             content=512, seed=1234),
    Workload("C2", "OVMono3D_dinov2_SFP lift (BASELINE configs[1])"),
    Workload("C3-1x2000", "box-count sweep, 1024-px SFP (BASELINE configs[2])", num_images=1, boxes_per_image=2000,
             seed=4321, **_SFP4),
    Workload("C3-8x2000", "box-count sweep, 1024-px SFP (BASELINE configs[2])", num_images=8, boxes_per_image=2000,
             seed=4321, **_SFP4),
    Workload("C3-8x300", "box-count sweep, 1024-px SFP (BASELINE configs[2])", num_images=8, boxes_per_image=300,
             seed=4321, **_SFP4),
    Workload("C5-sfp4", "clip / mae / sam / midas SFP pyramid (BASELINE configs[4])", seed=4321, **_SFP4),
    Workload("C5-dla", "DLA-FPN p2-p6 pyramid (BASELINE configs[4])", strides=(4, 8, 16, 32, 64), canvas=1024, seed=4321),
    Workload("C5-c128", "DINOv2 SFP with 128 channels (BASELINE configs[4])", channels=128, seed=4321),
    Workload("C5-c512", "DINOv2 SFP with 512 channels (BASELINE configs[4])", channels=512, seed=4321),
]}
# C4 (BASELINE configs[3]) is the C2 batch streamed 20 000 images deep over 1/2/4/8 ranks: bench.py --config C4


def get(name: str) -> Workload:
    if name not in WORKLOADS:
        raise KeyError("unknown workload %r (known: %s)" % (name, ", ".join(WORKLOADS)))
    return WORKLOADS[name]
