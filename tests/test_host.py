"""CPU-side checks: the C-ABI library loads and exports what include/ovmono3d_lift.h declares, the plugin registry,
config loading and the compat structures behave like the reference's, and nothing routes around the CUDA path."""
import os
import re

import pytest
import torch

from tests.golden import cases as gc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from ovmono3d_b200 import build, ops

    build.build_library()
    lib = ops.load_library()
    with open(os.path.join(ROOT, "include", "ovmono3d_lift.h")) as f:
        header = f.read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(ovm_[a-z0-9_]+)\s*\(", header))
    assert declared == set(ops.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.ovm_abi_version() == ops.OVM_ABI_VERSION
    assert ops.launch_count() == 0          # no compute without a GPU


def test_no_cpu_fallback():
    from ovmono3d_b200 import ops

    with pytest.raises(ops.LiftError):
        ops.roi_align([torch.zeros(1, 8, 4, 4)], [0.25], torch.zeros(1, 4), torch.zeros(1, dtype=torch.int32))
    head = _make_head(gc.CASE_BY_NAME["dinov2_agnostic"]).eval()
    from ovmono3d_b200.compat import Boxes, Instances
    inst = Instances((896, 896))
    inst.pred_boxes = Boxes(torch.tensor([[0.0, 0.0, 10.0, 10.0]]))
    inst.pred_classes = torch.zeros(1, dtype=torch.int64)
    feats = {f: torch.zeros(1, 16, 896 // s, 896 // s) for f, s in zip(["p2", "p3", "p4"], (7, 14, 28))}
    with pytest.raises(ops.LiftError):
        head._forward_cube(feats, [inst], [torch.eye(3)], [(896, 896)], [1.0])
    with pytest.raises(ops.LiftError):
        head.cube_head(torch.zeros(1, 16 * 49))


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "ovmono3d_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h")):
                with open(os.path.join(dirpath, fn)) as f:
                    src = f.read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), fn
                assert "/root/reference" not in src, fn


def _make_head(case, name="ROIHeads3DB200"):
    from ovmono3d_b200.compat import ShapeSpec
    from ovmono3d_b200.roi_heads import build_roi_heads

    c = gc.make_cfg(case)
    c.MODEL.ROI_HEADS.NAME = name
    shape = {f: ShapeSpec(channels=case.channels, stride=s) for f, s in zip(case.in_features, case.strides)}
    return build_roi_heads(c, shape, priors=None)


@pytest.mark.parametrize("case", gc.CASES, ids=lambda c: c.name)
def test_state_dict_keys_are_the_references(case):
    """The plugin's cube head accepts a reference CubeHead state dict unchanged (cube_head.py:64-145)."""
    head = _make_head(case)
    _, weights, _ = gc.build_case(case)
    res = head.cube_head.load_state_dict(weights, strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    keys = set(head.state_dict().keys())
    assert {"priors_dims_per_cat", "priors_z_scales"} <= keys
    assert ("priors_z_stats" in keys) == (case.cube.get("Z_TYPE") == "clusters")
    assert head.priors_dims_per_cat.shape == (1, case.num_classes, 2, 3)


def test_registry_and_cfg_selection():
    from ovmono3d_b200.compat import ROI_HEADS_REGISTRY
    from ovmono3d_b200 import roi_heads

    assert ROI_HEADS_REGISTRY.get("ROIHeads3DB200") is roi_heads.ROIHeads3DB200
    assert ROI_HEADS_REGISTRY.get("ROIHeads3DGDINOB200") is roi_heads.ROIHeads3DGDINOB200
    head = _make_head(gc.CASE_BY_NAME["clip_4level"], "ROIHeads3DGDINOB200")
    assert isinstance(head, roi_heads.ROIHeads3DGDINOB200)
    assert head.cube_pooler.scales == (1 / 4, 1 / 8, 1 / 16, 1 / 32)
    assert (head.cube_pooler.min_level, head.cube_pooler.max_level) == (2, 5)
    d = _make_head(gc.CASE_BY_NAME["dinov2_agnostic"])
    assert (d.cube_pooler.min_level, d.cube_pooler.max_level) == (2, 4)     # strides 7/14/28 (nohup.out:676-682)
    with pytest.raises(KeyError):
        ROI_HEADS_REGISTRY.get("NoSuchHead")
    bad = gc.make_cfg(gc.CASE_BY_NAME["dinov2_agnostic"])
    bad.MODEL.ROI_CUBE_HEAD.POSE_TYPE = "matrix"
    from ovmono3d_b200.compat import ShapeSpec
    with pytest.raises(ValueError):
        roi_heads.build_roi_heads(bad, {f: ShapeSpec(channels=16, stride=s) for f, s in
                                        zip(["p2", "p3", "p4"], (7, 14, 28))})


def test_config_defaults_match_reference_values():
    from ovmono3d_b200.config import get_cfg

    cube = get_cfg().MODEL.ROI_CUBE_HEAD
    assert (cube.POOLER_RESOLUTION, cube.POOLER_SAMPLING_RATIO, cube.POOLER_TYPE) == (7, 0, "ROIAlignV2")
    assert (cube.NUM_FC, cube.FC_DIM, cube.Z_TYPE, cube.POSE_TYPE) == (2, 1024, "direct", "6d")
    assert (cube.VIRTUAL_DEPTH, cube.VIRTUAL_FOCAL, cube.CLUSTER_BINS) == (True, 512.0, 1)
    assert (cube.SHARED_FC, cube.DIMS_PRIORS_ENABLED, cube.DIMS_PRIORS_FUNC, cube.USE_CONFIDENCE) == \
        (True, True, "exp", 1.0)
    cfg = get_cfg()
    cfg.merge_from_list(["MODEL.ROI_CUBE_HEAD.Z_TYPE", "log", "MODEL.ROI_HEADS.NAME", "ROIHeads3DGDINOB200",
                         "MODEL.ROI_CUBE_HEAD.SCALE_ROI_BOXES", "1.5"])
    assert cfg.MODEL.ROI_CUBE_HEAD.Z_TYPE == "log" and cfg.MODEL.ROI_CUBE_HEAD.SCALE_ROI_BOXES == 1.5


def test_yaml_with_base_inheritance(tmp_path):
    from ovmono3d_b200.config import load_reference_yaml

    (tmp_path / "Base.yaml").write_text(
        "MODEL:\n  ROI_CUBE_HEAD:\n    DIMS_PRIORS_ENABLED: False\n    Z_TYPE: 'direct'\n"
        "  ROI_HEADS:\n    IN_FEATURES: ['p2', 'p3', 'p4', 'p5', 'p6']\n")
    (tmp_path / "child.yaml").write_text(
        "_BASE_: \"Base.yaml\"\nMODEL:\n  ROI_HEADS:\n    IN_FEATURES: ['p2', 'p3', 'p4']\n    NAME: 'ROIHeads3D'\n")
    cfg = load_reference_yaml(str(tmp_path / "child.yaml"), ["MODEL.ROI_HEADS.NAME", "ROIHeads3DGDINOB200"])
    assert cfg.MODEL.ROI_HEADS.IN_FEATURES == ["p2", "p3", "p4"]
    assert cfg.MODEL.ROI_CUBE_HEAD.DIMS_PRIORS_ENABLED is False
    assert cfg.MODEL.ROI_HEADS.NAME == "ROIHeads3DGDINOB200"
    assert cfg.MODEL.ROI_CUBE_HEAD.FC_DIM == 1024      # default kept


def test_instances_and_boxes_behave_like_detectron2():
    from ovmono3d_b200.compat import Boxes, Instances

    b = Boxes(torch.tensor([[0.0, 0.0, 4.0, 2.0], [1.0, 1.0, 3.0, 5.0]]))
    assert b.area().tolist() == [8.0, 8.0] and b.get_centers().tolist() == [[2.0, 1.0], [2.0, 3.0]]
    inst = Instances((10, 20))
    assert not inst.has("scores")
    inst.pred_boxes = b
    inst.scores = torch.tensor([0.5, 0.25])
    assert len(inst) == 2 and inst.image_size == (10, 20) and inst.has("scores")
    with pytest.raises(AssertionError):
        inst.pred_classes = torch.zeros(3)
    sub = inst[1]
    assert len(sub) == 1 and sub.scores.tolist() == [0.25]
    with pytest.raises(AttributeError):
        _ = inst.nope
    assert len(Boxes(torch.zeros(0))) == 0
