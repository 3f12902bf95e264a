"""Config node + the `MODEL.ROI_CUBE_HEAD.*` keys the lift path honours.

The reference uses yacs `CfgNode` through detectron2 (`cubercnn/config/config.py:4-241`) with yaml
`_BASE_` inheritance and `KEY VALUE` command-line overrides (`tools/train_net.py:363-374`).  Only
the keys that change the lift's arithmetic are declared here (SURVEY.md §5 "Config / flags"); the
defaults are the reference's (`config.py:118-214,239`) and `load_reference_yaml` reads the shipped
`configs/*.yaml` files unchanged, ignoring keys that belong to other subsystems.
"""
from __future__ import annotations

import ast
import copy
import os
from typing import Any, Dict, Iterable, List

import yaml

BASE_KEY = "_BASE_"


class CfgNode(dict):
    """Attribute-access nested dict, API-compatible with the yacs subset the reference uses."""

    def __init__(self, init_dict: Dict[str, Any] | None = None):
        super().__init__()
        for k, v in (init_dict or {}).items():
            self[k] = CfgNode(v) if isinstance(v, dict) and not isinstance(v, CfgNode) else v

    def __getattr__(self, name: str) -> Any:
        if name in self:
            return self[name]
        raise AttributeError(name)

    def __setattr__(self, name: str, value: Any) -> None:
        self[name] = value

    def clone(self) -> "CfgNode":
        return copy.deepcopy(self)

    def merge_from_other_cfg(self, other: Dict[str, Any], allow_new: bool = True) -> None:
        for k, v in other.items():
            if isinstance(v, dict) and isinstance(self.get(k), dict):
                self[k].merge_from_other_cfg(v, allow_new)
            elif k in self or allow_new:
                self[k] = CfgNode(v) if isinstance(v, dict) else v

    def merge_from_file(self, path: str, allow_new: bool = True) -> None:
        self.merge_from_other_cfg(_load_yaml_with_base(path), allow_new)

    def merge_from_list(self, cfg_list: Iterable[Any]) -> None:
        cfg_list = list(cfg_list)
        assert len(cfg_list) % 2 == 0, "Override list has odd length: {}".format(cfg_list)
        for full_key, v in zip(cfg_list[0::2], cfg_list[1::2]):
            node = self
            keys = full_key.split(".")
            for sub in keys[:-1]:
                if sub not in node:
                    node[sub] = CfgNode()
                node = node[sub]
            if isinstance(v, str):
                try:
                    v = ast.literal_eval(v)
                except (ValueError, SyntaxError):
                    pass
            node[keys[-1]] = v

    def freeze(self) -> None:  # yacs API; the lift path never mutates cfg
        pass

    def defrost(self) -> None:
        pass


CN = CfgNode


def _load_yaml_with_base(path: str) -> Dict[str, Any]:
    with open(path, "r") as f:
        cfg = yaml.unsafe_load(f) or {}
    if BASE_KEY in cfg:
        base = cfg.pop(BASE_KEY)
        if not os.path.isabs(base):
            base = os.path.join(os.path.dirname(path), base)
        merged = CfgNode(_load_yaml_with_base(base))
        merged.merge_from_other_cfg(cfg)
        return merged
    return cfg


def add_lift_defaults(cfg: CfgNode) -> CfgNode:
    """Declare every key the lift path reads, with the reference's default values."""
    cfg.setdefault("MODEL", CN())
    m = cfg.MODEL
    m.setdefault("DEVICE", "cuda")
    m.setdefault("ROI_HEADS", CN())
    m.ROI_HEADS.setdefault("NAME", "ROIHeads3DGDINOB200")
    m.ROI_HEADS.setdefault("IN_FEATURES", ["p2", "p3", "p4", "p5", "p6"])
    m.ROI_HEADS.setdefault("NUM_CLASSES", 43)
    m.setdefault("FPN", CN())
    m.FPN.setdefault("OUT_CHANNELS", 256)
    m.setdefault("RPN", CN())
    m.RPN.setdefault("IGNORE_THRESHOLD", 0.5)

    m.setdefault("ROI_CUBE_HEAD", CN())
    h = m.ROI_CUBE_HEAD
    defaults = dict(
        NAME="CubeHead",
        POOLER_RESOLUTION=7,          # config.py:120
        POOLER_SAMPLING_RATIO=0,      # config.py:121
        POOLER_TYPE="ROIAlignV2",     # config.py:122
        NUM_CONV=0,
        CONV_DIM=256,
        NUM_FC=2,                     # config.py:127
        FC_DIM=1024,                  # config.py:128
        Z_TYPE="direct",              # config.py:133   direct | sigmoid | log | clusters
        POSE_TYPE="6d",               # config.py:137   6d | quaternion | euler
        INVERSE_Z_WEIGHT=False,
        VIRTUAL_DEPTH=True,           # config.py:144
        VIRTUAL_FOCAL=512.0,          # config.py:145
        DISENTANGLED_LOSS=True,
        CLUSTER_BINS=1,               # config.py:156
        ALLOCENTRIC_POSE=True,        # config.py:165
        CHAMFER_POSE=True,
        SHARED_FC=True,               # config.py:174
        DIMS_PRIORS_ENABLED=True,     # config.py:183 (shipped yaml turns it off, Base.yaml:81)
        DIMS_PRIORS_FUNC="exp",       # config.py:189   exp | sigmoid
        USE_CONFIDENCE=1.0,           # config.py:192
        LOSS_W_3D=1.0,
        LOSS_W_XY=1.0,
        LOSS_W_Z=1.0,
        LOSS_W_DIMS=1.0,
        LOSS_W_POSE=1.0,
        LOSS_W_JOINT=1.0,
        SCALE_ROI_BOXES=0.0,          # config.py:239
    )
    for k, v in defaults.items():
        h.setdefault(k, v)
    # 2-D box branch (d2 defaults; Base.yaml:60-70 overrides SCORE_THRESH_TEST to 0.01)
    m.ROI_HEADS.setdefault("SCORE_THRESH_TEST", 0.05)
    m.ROI_HEADS.setdefault("NMS_THRESH_TEST", 0.5)
    m.setdefault("ROI_BOX_HEAD", CN())
    for k, v in dict(NAME="FastRCNNConvFCHead", NUM_FC=2, FC_DIM=1024, NUM_CONV=0, POOLER_RESOLUTION=7,
                     POOLER_SAMPLING_RATIO=0, POOLER_TYPE="ROIAlignV2", CLS_AGNOSTIC_BBOX_REG=False,
                     BBOX_REG_WEIGHTS=(10.0, 10.0, 5.0, 5.0),
                     BUILD=False).items():   # BUILD is this package's key: the GDINO head never uses the branch
        m.ROI_BOX_HEAD.setdefault(k, v)
    cfg.setdefault("INPUT", CN())
    cfg.INPUT.setdefault("MIN_SIZE_TEST", 512)
    cfg.setdefault("TEST", CN())
    cfg.TEST.setdefault("DETECTIONS_PER_IMAGE", 100)  # config.py:220
    return cfg


def get_cfg() -> CfgNode:
    return add_lift_defaults(CfgNode())


def load_reference_yaml(path: str, overrides: List[Any] | None = None) -> CfgNode:
    """`get_cfg()` + a shipped reference yaml (with `_BASE_`) + `KEY VALUE` overrides."""
    cfg = get_cfg()
    cfg.merge_from_file(path)
    if overrides:
        cfg.merge_from_list(overrides)
    return cfg


# Pyramids of the shipped configs: name -> (in_features, strides, canvas).
PYRAMIDS = {
    "dinov2": (["p2", "p3", "p4"], [7, 14, 28], 896),                       # OVMono3D_dinov2_SFP.yaml:37-44, dino.py:143
    "clip":   (["p2", "p3", "p4", "p5"], [4, 8, 16, 32], 1024),             # OVMono3D_clip_SFP.yaml:34-44, clip.py:155
    "mae":    (["p2", "p3", "p4", "p5"], [4, 8, 16, 32], 1024),             # mae.py:140
    "sam":    (["p2", "p3", "p4", "p5"], [4, 8, 16, 32], 1024),             # sam.py:130
    "midas":  (["p2", "p3", "p4", "p5"], [4, 8, 16, 32], 1024),             # midas_final.py:108
    "dla_fpn": (["p2", "p3", "p4", "p5", "p6"], [4, 8, 16, 32, 64], 1024),  # Base.yaml:44,63, dla.py:460-461
}
