"""The oracle (oracle/lift_oracle.py) against vectors produced by the reference's own code.

`tests/golden/*.npz` were written by `tests/golden/make_golden.py`, which runs the unmodified
`ROIHeads3D._forward_cube` / `CubeHead` from /root/reference with stubbed third-party imports.
The oracle restates the same op sequence, so agreement is required to be (near) bit-exact.
"""
import json
import os

import numpy as np
import pytest
import torch

from oracle import lift_oracle as lo
from tests import helpers
from tests.golden import cases as gc

FIELDS = ("scores", "pred_bbox3D", "pred_center_cam", "pred_center_2D", "pred_dimensions", "pred_pose")


@pytest.mark.parametrize("case", gc.CASES, ids=lambda c: c.name)
def test_oracle_matches_reference_vectors(case):
    g = np.load(gc.golden_path(case))
    inp, weights, priors, cfg, out = helpers.run_oracle(case)
    np.testing.assert_allclose(gc.input_checksum(inp, weights), g["checksum"], rtol=1e-12,
                               err_msg="synthetic inputs drifted: regenerate goldens")
    n = int(g["n"])
    assert out.pooled.shape == tuple(g["pooled_shape"])
    np.testing.assert_array_equal(out.pooled.flatten()[:: case.pooled_stride].numpy(), g["pooled"])
    if n == 0:
        return
    np.testing.assert_array_equal(out.levels.numpy(), g["levels"])
    for k, name in (("deltas", "raw_deltas"), ("z", "raw_z"), ("dims", "raw_dims"), ("pose", "raw_pose"),
                    ("uncert", "raw_uncert")):
        if name in g:
            np.testing.assert_allclose(out.head_raw[k].numpy(), g[name], rtol=1e-6, atol=1e-7)
    j = 0
    for i in range(len(inp.boxes)):
        if "im%d_scores" % i not in g:
            assert len(inp.boxes[i]) == 0
            j += 1 if len(out.scores) > j and len(out.scores[j]) == 0 else 0
            continue
        for f in FIELDS:
            got = getattr(out, f)[i].numpy()
            np.testing.assert_allclose(got, g["im%d_%s" % (i, f)], rtol=1e-6, atol=1e-7, err_msg=f)


def test_cuboid_corners_against_logged_detections():
    """Known-answer: detections the reference logged (err.txt:5): bbox3D == R.local + centre."""
    with open(os.path.join(gc.GOLDEN_DIR, "logged_detections.json")) as f:
        dets = json.load(f)["detections"]
    assert len(dets) >= 1
    for d in dets:
        box = torch.tensor([d["center_cam"] + d["dimensions"]], dtype=torch.float32)
        R = torch.tensor([d["pose"]], dtype=torch.float32)
        verts = lo.get_cuboid_verts(box, R)[0].numpy()
        np.testing.assert_allclose(verts, np.array(d["bbox3D"], dtype=np.float32), atol=2e-6)
        assert abs(d["depth"] - d["center_cam"][2]) < 1e-12
