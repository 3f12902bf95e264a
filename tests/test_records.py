"""Records -> reference result schema.

  * the oracle (oracle/postprocess_oracle.py) against vectors written by the reference's OWN
    `Omni3DEvaluator.instances_to_coco_json` (tests/golden/make_golden_postprocess.py -> coco_json_cases.json): exact;
  * the host-side formatting of `ovmono3d_b200/records.py` against the same vectors (CPU);
  * `ovm_records_postprocess` + formatting (`records.records_to_coco_json` on CUDA records) against both: exact --
    same fp32 operations as `Boxes.scale` / `clip` / `nonempty` and `BoxMode.convert`.
"""
import json
import os

import numpy as np
import pytest
import torch

from oracle import postprocess_oracle as po
from ovmono3d_b200 import ops, records
from tests.golden import make_golden_postprocess as mg

FIXTURE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "coco_json_cases.json")


def _cases():
    with open(FIXTURE) as f:
        return json.load(f)


def _oracle(rec, case):
    res, start = [], 0
    for i, k in enumerate(case["counts"]):
        r = rec[start:start + k]
        start += k
        f = po.detector_postprocess(mg.fields_of(r), tuple(case["sizes"][i]), case["outs"][i][0], case["outs"][i][1])
        res.append(po.instances_to_coco_json(f, case["ids"][i]))
    return res


def test_oracle_matches_the_reference_methods_own_output():
    cases = _cases()
    assert [c["case"] for c in cases] == json.loads(json.dumps(mg.CASES)), "fixture is stale: re-run make_golden_postprocess"
    for c in cases:
        rec = mg.random_records(c["case"]["seed"], c["case"]["counts"], tuple(c["case"]["sizes"][0]))
        got = _oracle(rec, c["case"])
        assert any(len(g) < k for g, k in zip(got, c["case"]["counts"])), "the case must exercise the empty-box filter"
        assert got == c["results"]           # exact: same floats, same ints, same nesting


def test_rows_to_coco_json_formatting():
    """Rows as the kernel leaves them (scaled / clipped / filtered, XYWH) -> the reference's dicts."""
    for c in _cases():
        case = c["case"]
        rec = mg.random_records(case["seed"], case["counts"], tuple(case["sizes"][0]))
        rows = np.zeros((rec.shape[0], ops.RECORD_WIDTH), np.float32)
        counts, offsets, start = [], [], 0
        for i, k in enumerate(case["counts"]):
            r = rec[start:start + k]
            f = po.detector_postprocess(mg.fields_of(r), tuple(case["sizes"][i]), case["outs"][i][0], case["outs"][i][1])
            keep = len(f["pred_boxes"])
            out = torch.zeros(keep, ops.RECORD_WIDTH)
            b = f["pred_boxes"].clone()
            b[:, 2] -= b[:, 0]
            b[:, 3] -= b[:, 1]
            out[:, ops.REC_BOX2D] = b
            out[:, ops.REC_SCORE] = f["scores"]
            out[:, ops.REC_CLASS] = f["pred_classes"].float()
            out[:, ops.REC_BBOX3D] = f["pred_bbox3D"].reshape(keep, 24)
            out[:, ops.REC_CENTER_CAM], out[:, ops.REC_CENTER_2D] = f["pred_center_cam"], f["pred_center_2D"]
            out[:, ops.REC_DIMENSIONS], out[:, ops.REC_POSE] = f["pred_dimensions"], f["pred_pose"].reshape(keep, 9)
            rows[start:start + keep] = out.numpy()
            counts.append(keep)
            offsets.append(start)
            start += k
        got = records.rows_to_coco_json(rows, counts, offsets, case["ids"])
        assert got == c["results"]
        json.dumps(got)                      # serialisable as the reference writes it (omni_instances_results.json)


def test_records_on_the_host_are_refused():
    with pytest.raises(ops.LiftError):
        records.records_to_coco_json(torch.zeros(3, ops.RECORD_WIDTH), [3], [1], [(8, 8)], [(8, 8)])
    with pytest.raises(ops.LiftError):
        records.records_to_coco_json(np.zeros((3, ops.RECORD_WIDTH), np.float32), [3], [1], [(8, 8)], [(8, 8)])
    g = records.postprocess_geometry([(512, 683)], [(480, 640)])
    assert g.dtype == np.float32 and g[0, 0] == np.float32(640 / 683) and g[0, 3] == 480


@pytest.mark.gpu
def test_records_to_coco_json_on_the_device_is_bit_identical():
    dev = torch.device("cuda:0")
    for c in _cases():
        case = c["case"]
        rec = mg.random_records(case["seed"], case["counts"], tuple(case["sizes"][0]))
        got = records.records_to_coco_json(rec.to(dev), case["counts"], case["ids"], [tuple(s) for s in case["sizes"]],
                                           [tuple(o) for o in case["outs"]])
        assert got == c["results"] and got == _oracle(rec, case)
    # empty batch / empty images / malformed input
    assert records.records_to_coco_json(torch.zeros(0, ops.RECORD_WIDTH, device=dev), [0, 0], [1, 2], [(8, 8)] * 2,
                                        [(8, 8)] * 2) == [[], []]
    with pytest.raises(ops.LiftError):
        records.records_to_coco_json(torch.zeros(3, 5, device=dev), [3], [1], [(8, 8)], [(8, 8)])
    with pytest.raises(ops.LiftError):
        records.records_to_coco_json(torch.zeros(3, ops.RECORD_WIDTH, device=dev), [2], [1], [(8, 8)], [(8, 8)])


@pytest.mark.gpu
def test_records_postprocess_large_image_and_nan_boxes():
    """More records than one pass of the CTA holds, NaN coordinates (dropped, like `nonempty()` does)."""
    dev = torch.device("cuda:0")
    counts = [1500, 3]
    rec = mg.random_records(11, counts, (896, 896))
    rec[5, 41] = float("nan")
    got = records.records_to_coco_json(rec.to(dev), counts, [1, 2], [(896, 896)] * 2, [(1080, 1920), (600, 800)])
    want = _oracle(rec, dict(counts=counts, sizes=[(896, 896)] * 2, outs=[(1080, 1920), (600, 800)], ids=[1, 2]))
    assert got == want
