"""Golden vectors for SURVEY.md section 8f rank 1 from the reference's OWN `Omni3DEvaluator.instances_to_coco_json`
(cubercnn/evaluation/omni3d_evaluation.py:1200-1252), imported unmodified from /root/reference through the stubs of
`ref_harness.py` (detectron2 / pytorch3d / pycocotools are not installed; nothing of them is executed by this method
except `BoxMode.convert(XYXY_ABS -> XYWH_ABS)`, restated below from d2's published source).

    python -m tests.golden.make_golden_postprocess        # writes tests/golden/coco_json_cases.json

The d2 half of the row -- `detector_postprocess` (`Boxes.scale` / `clip` / `nonempty`, called from
`GeneralizedRCNN._postprocess`, rcnn3d.py:113-115) -- lives in the un-vendored fork and stays our restatement
(`oracle/postprocess_oracle.detector_postprocess`); the vectors below take its output as the method's input, so they
pin the reference-owned half exactly and the d2 half only through `oracle/`.
"""
from __future__ import annotations

import importlib
import json
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

CASES = [  # (seed, boxes per image, network (H, W), original (height, width), image ids)
    dict(seed=7, counts=[20, 0, 7, 1, 33], sizes=[(512, 683), (512, 512), (384, 512), (600, 800), (512, 640)],
         outs=[(480, 640), (1024, 1024), (768, 1024), (333, 444), (1920, 2400)], ids=[3, 5, 8, 13, 21]),
    dict(seed=8, counts=[100, 100], sizes=[(896, 896), (896, 896)], outs=[(1080, 1920), (375, 1242)], ids=[101, 102]),
]


def random_records(seed, counts, size):
    """Seeded [n, 48] records with boxes that partly leave the image and some empty ones (also used by the tests)."""
    from ovmono3d_b200 import ops

    g = torch.Generator().manual_seed(seed)
    n = sum(counts)
    rec = torch.randn(n, ops.RECORD_WIDTH, generator=g)
    xy = torch.rand(n, 2, generator=g) * torch.tensor([size[1], size[0]]) * 1.2 - 30
    wh = torch.rand(n, 2, generator=g) * 200
    wh[::7] = 0.0
    rec[:, ops.REC_BOX2D] = torch.cat([xy, xy + wh], 1)
    rec[:, ops.REC_SCORE] = torch.rand(n, generator=g)
    rec[:, ops.REC_CLASS] = torch.randint(0, 50, (n,), generator=g).float()
    rec[:, 47] = 0
    return rec


def fields_of(r):
    from ovmono3d_b200 import ops

    k = r.shape[0]
    return dict(pred_boxes=r[:, ops.REC_BOX2D].clone(), scores=r[:, ops.REC_SCORE].clone(),
                pred_classes=r[:, ops.REC_CLASS].long(), pred_bbox3D=r[:, ops.REC_BBOX3D].reshape(k, 8, 3),
                pred_center_cam=r[:, ops.REC_CENTER_CAM], pred_center_2D=r[:, ops.REC_CENTER_2D],
                pred_dimensions=r[:, ops.REC_DIMENSIONS], pred_pose=r[:, ops.REC_POSE].reshape(k, 3, 3))


def reference_method():
    from tests.golden import ref_harness as rh

    rh.install_stubs()
    d2s = importlib.import_module("detectron2.structures")

    class BoxMode:      # d2 structures/boxes.py, the one conversion the method uses
        XYXY_ABS, XYWH_ABS = 0, 1

        @staticmethod
        def convert(box, from_mode, to_mode):
            assert from_mode == BoxMode.XYXY_ABS and to_mode == BoxMode.XYWH_ABS
            arr = torch.from_numpy(np.asarray(box)).clone()
            arr[:, 2] -= arr[:, 0]
            arr[:, 3] -= arr[:, 1]
            return arr.numpy()

    d2s.BoxMode = BoxMode
    for pkg in ("cubercnn.evaluation",):
        m = types.ModuleType(pkg)
        m.__path__ = [os.path.join(rh.REFERENCE_ROOT, *pkg.split("."))]
        sys.modules[pkg] = m
    for pkg in ("cubercnn.data", "cubercnn.vis", "cubercnn.vis.logperf", "cubercnn.util"):   # imported, never executed here
        m = rh._StubModule(pkg)
        m.__path__ = []
        sys.modules[pkg] = m
    sys.modules["cubercnn"].util = sys.modules["cubercnn.util"]
    mod = importlib.import_module("cubercnn.evaluation.omni3d_evaluation")
    return mod.Omni3DEvaluator.instances_to_coco_json


def main():
    from oracle import postprocess_oracle as po
    from ovmono3d_b200.compat import Boxes, Instances

    method = reference_method()
    out = []
    for case in CASES:
        rec = random_records(case["seed"], case["counts"], case["sizes"][0])
        per_image, start = [], 0
        for i, k in enumerate(case["counts"]):
            r = rec[start:start + k]
            start += k
            f = po.detector_postprocess(fields_of(r), case["sizes"][i], case["outs"][i][0], case["outs"][i][1])
            inst = Instances(case["outs"][i])
            if len(f["pred_boxes"]):
                inst.pred_boxes = Boxes(f["pred_boxes"])
                for name in ("scores", "pred_classes", "pred_bbox3D", "pred_center_cam", "pred_center_2D",
                             "pred_dimensions", "pred_pose"):
                    inst.set(name, f[name])
                per_image.append(method(types.SimpleNamespace(), inst, case["ids"][i]))
            else:
                per_image.append([])      # len(instances) == 0 -> [] (:1204-1206)
        out.append(dict(case=case, results=per_image))
    path = os.path.join(HERE, "coco_json_cases.json")
    with open(path, "w") as f:
        json.dump(out, f)
    print("wrote %s (%d cases, %d detections)" % (path, len(out), sum(len(x) for c in out for x in c["results"])))


if __name__ == "__main__":
    main()
