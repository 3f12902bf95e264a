"""Records -> reference result schema (ovmono3d_b200/records.py) against the oracle restatement of d2
`detector_postprocess` + `Omni3DEvaluator.instances_to_coco_json` (oracle/postprocess_oracle.py).  Host logic: runs
without a GPU; the numbers must be bit-identical (same fp32 operations in the same order)."""
import json

import numpy as np
import torch

from oracle import postprocess_oracle as po
from ovmono3d_b200 import ops, records


def _random_records(seed, counts, size):
    g = torch.Generator().manual_seed(seed)
    n = sum(counts)
    rec = torch.randn(n, ops.RECORD_WIDTH, generator=g)
    xy = torch.rand(n, 2, generator=g) * torch.tensor([size[1], size[0]]) * 1.2 - 30   # some boxes leave the image
    wh = torch.rand(n, 2, generator=g) * 200
    wh[::7] = 0.0                                                                       # empty boxes are dropped
    rec[:, ops.REC_BOX2D] = torch.cat([xy, xy + wh], 1)
    rec[:, ops.REC_SCORE] = torch.rand(n, generator=g)
    rec[:, ops.REC_CLASS] = torch.randint(0, 50, (n,), generator=g).float()
    rec[:, 47] = 0
    return rec


def _oracle(rec, counts, ids, sizes, outs):
    res, start = [], 0
    for i, k in enumerate(counts):
        r = rec[start:start + k]
        start += k
        fields = dict(pred_boxes=r[:, ops.REC_BOX2D].clone(), scores=r[:, ops.REC_SCORE].clone(),
                      pred_classes=r[:, ops.REC_CLASS].long(), pred_bbox3D=r[:, ops.REC_BBOX3D].reshape(k, 8, 3),
                      pred_center_cam=r[:, ops.REC_CENTER_CAM], pred_center_2D=r[:, ops.REC_CENTER_2D],
                      pred_dimensions=r[:, ops.REC_DIMENSIONS], pred_pose=r[:, ops.REC_POSE].reshape(k, 3, 3))
        res.append(po.instances_to_coco_json(po.detector_postprocess(fields, sizes[i], outs[i][0], outs[i][1]), ids[i]))
    return res


def test_records_to_coco_json_bit_identical():
    counts = [20, 0, 7, 1, 33]
    sizes = [(512, 683), (512, 512), (384, 512), (600, 800), (512, 640)]
    outs = [(480, 640), (1024, 1024), (768, 1024), (333, 444), (1920, 2400)]
    ids = [3, 5, 8, 13, 21]
    rec = _random_records(7, counts, (512, 683))
    got = records.records_to_coco_json(rec, counts, ids, sizes, outs)
    want = _oracle(rec, counts, ids, sizes, outs)
    assert [len(g) for g in got] == [len(w) for w in want]
    assert any(len(g) < c for g, c in zip(got, counts)), "the case must exercise the empty-box filter"
    assert got == want                       # exact: same floats, same ints, same nesting
    json.dumps(got)                          # serialisable as the reference writes it (omni_instances_results.json)


def test_records_to_coco_json_empty_and_errors():
    assert records.records_to_coco_json(torch.zeros(0, ops.RECORD_WIDTH), [0, 0], [1, 2], [(8, 8)] * 2, [(8, 8)] * 2) == [[], []]
    try:
        records.records_to_coco_json(np.zeros((3, 5), np.float32), [3], [1], [(8, 8)], [(8, 8)])
    except ops.LiftError:
        pass
    else:
        raise AssertionError("a malformed record array must be refused")
    try:
        records.records_to_coco_json(np.zeros((3, ops.RECORD_WIDTH), np.float32), [2], [1], [(8, 8)], [(8, 8)])
    except ops.LiftError:
        pass
    else:
        raise AssertionError("counts that do not cover the records must be refused")
