"""Run the reference's OWN `ROIHeads3D._forward_cube` / `CubeHead` code in this container.

`/root/reference` imports detectron2, pytorch3d, fvcore, groundingdino and matplotlib at module
top; none is installed and the submodule directories are empty.  This harness installs stub
modules for those names -- permissive dummies for everything the lift path never executes, and
`oracle/thirdparty.py`'s restatements for the handful of functions it does execute
(`ROIPooler`, `rotation_6d_to_matrix`, `axis_angle_to_matrix`, ...) -- then imports the reference
files *unmodified from where they lie* and drives them.  It exists only to generate
`tests/golden/*.npz` (see `make_golden.py`); `/root/reference` does not exist on the GPU box, so
nothing at test/bench time imports this file's reference side.
"""
from __future__ import annotations

import importlib
import importlib.abc
import importlib.machinery
import math
import os
import sys
import types
from typing import Any, Dict, List

import torch
from torch import nn

REFERENCE_ROOT = os.environ.get("OVMONO3D_REFERENCE", "/root/reference")

_STUB_ROOTS = ("detectron2", "pytorch3d", "fvcore", "groundingdino", "matplotlib", "pycocotools",
               "iopath", "termcolor")


class _PermissiveMeta(type):
    def __getattr__(cls, name: str) -> Any:
        if name.startswith("__"):
            raise AttributeError(name)
        return lambda *a, **k: None


def _dummy_class(name: str) -> type:
    return _PermissiveMeta(name, (), {"__init__": lambda self, *a, **k: None})


class _StubModule(types.ModuleType):
    def __getattr__(self, name: str) -> Any:
        if name.startswith("__"):
            raise AttributeError(name)
        obj = _dummy_class(name)
        setattr(self, name, obj)
        return obj


class _StubFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, fullname, path=None, target=None):
        if fullname.split(".")[0] in _STUB_ROOTS:
            return importlib.machinery.ModuleSpec(fullname, self, is_package=True)
        return None

    def create_module(self, spec):
        m = _StubModule(spec.name)
        m.__path__ = []
        return m

    def exec_module(self, module):
        pass


_installed = False


def install_stubs() -> None:
    """Idempotently install the stub finder + the functional overrides, and put the reference on sys.path."""
    global _installed
    if _installed:
        return
    _installed = True
    repo_root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    if repo_root not in sys.path:
        sys.path.insert(0, repo_root)
    from oracle import thirdparty as tp
    from ovmono3d_b200 import compat

    assert not compat.HAVE_DETECTRON2, "real detectron2 present: run the reference directly instead"
    sys.meta_path.insert(0, _StubFinder())

    # ---- pytorch3d: the functions the path executes (cube_head.py:10-15, math_util.py:34)
    p3d_t = importlib.import_module("pytorch3d.transforms")
    p3d_rc = importlib.import_module("pytorch3d.transforms.rotation_conversions")
    for mod in (p3d_t, p3d_rc):
        mod.rotation_6d_to_matrix = tp.rotation_6d_to_matrix
        mod.euler_angles_to_matrix = tp.euler_angles_to_matrix
        mod.quaternion_to_matrix = tp.quaternion_to_matrix
        mod.axis_angle_to_matrix = tp.axis_angle_to_matrix
        mod._copysign = tp._copysign

    # ---- fvcore: c2_xavier_fill (cube_head.py:8,72)
    wi = importlib.import_module("fvcore.nn.weight_init")

    def c2_xavier_fill(module: nn.Module) -> None:
        nn.init.kaiming_uniform_(module.weight, a=1)
        if module.bias is not None:
            nn.init.constant_(module.bias, 0)

    wi.c2_xavier_fill = c2_xavier_fill
    importlib.import_module("fvcore.nn").weight_init = wi

    # ---- detectron2: registry / structures / configurable / poolers / StandardROIHeads
    importlib.import_module("detectron2.utils.registry").Registry = compat.Registry
    d2_cfg = importlib.import_module("detectron2.config")
    d2_cfg.configurable = compat.configurable
    d2_struct = importlib.import_module("detectron2.structures")
    d2_struct.Instances = compat.Instances
    d2_struct.Boxes = compat.Boxes
    d2_layers = importlib.import_module("detectron2.layers")
    d2_layers.ShapeSpec = compat.ShapeSpec
    d2_layers.nonzero_tuple = lambda x: torch.nonzero(x, as_tuple=True)

    class ROIPooler(nn.Module):
        """d2 `ROIPooler` facade over the restated algorithm (takes `List[Boxes]` like the real one)."""

        def __init__(self, output_size, scales, sampling_ratio, pooler_type, canonical_box_size=224,
                     canonical_level=4):
            super().__init__()
            self.impl = tp.ROIPoolerOracle(output_size, scales, sampling_ratio, pooler_type,
                                           canonical_box_size, canonical_level)

        def forward(self, x: List[torch.Tensor], box_lists):
            return self.impl(x, [b.tensor for b in box_lists])

    importlib.import_module("detectron2.modeling.poolers").ROIPooler = ROIPooler

    class StandardROIHeads(nn.Module):
        """Just enough of d2 `StandardROIHeads` for `ROIHeads3D.__init__` / `_forward_cube`."""

        def __init__(self, *, num_classes=None, box_in_features=None, **kwargs):
            super().__init__()
            self.num_classes = num_classes
            self.in_features = self.box_in_features = box_in_features

    d2_rh = importlib.import_module("detectron2.modeling.roi_heads")
    d2_rh.StandardROIHeads = StandardROIHeads
    d2_rh.ROI_HEADS_REGISTRY = compat.Registry("ROI_HEADS_REFERENCE")

    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    # keep `cubercnn/modeling/roi_heads/__init__.py` from dragging in every other subsystem
    for pkg in ("cubercnn", "cubercnn.modeling", "cubercnn.modeling.roi_heads",
                "cubercnn.modeling.proposal_generator"):
        m = types.ModuleType(pkg)
        m.__path__ = [os.path.join(REFERENCE_ROOT, *pkg.split("."))]
        sys.modules[pkg] = m
    for name in ("cubercnn.modeling.proposal_generator.rpn", "cubercnn.modeling.roi_heads.fast_rcnn"):
        sys.modules[name] = _StubModule(name)
    sys.modules["cubercnn"].util = importlib.import_module("cubercnn.util")


def reference_modules():
    install_stubs()
    roi_heads = importlib.import_module("cubercnn.modeling.roi_heads.roi_heads")
    cube_head = importlib.import_module("cubercnn.modeling.roi_heads.cube_head")
    math_util = importlib.import_module("cubercnn.util.math_util")
    return roi_heads, cube_head, math_util


def build_reference_head(cfg, in_features: List[str], strides: List[int], channels: int, priors=None):
    """Construct the reference `ROIHeads3D` through its own `_init_cube_head` + `__init__`."""
    roi_heads, _, _ = reference_modules()
    from ovmono3d_b200.compat import ShapeSpec

    input_shape = {f: ShapeSpec(channels=channels, stride=s) for f, s in zip(in_features, strides)}
    kwargs: Dict[str, Any] = roi_heads.ROIHeads3D._init_cube_head(cfg, input_shape)
    kwargs["priors"] = priors
    kwargs["num_classes"] = cfg.MODEL.ROI_HEADS.NUM_CLASSES
    kwargs["box_in_features"] = in_features
    head = roi_heads.ROIHeads3D(**kwargs)
    head.eval()
    return head


def run_reference_forward_cube(head, inputs, with_scores: bool = True):
    """Drive `ROIHeads3D._forward_cube` (roi_heads.py:329) on `synth.LiftInputs`."""
    from ovmono3d_b200.compat import Boxes, Instances

    feats = {f: t for f, t in zip(head.in_features, inputs.features)}
    instances = []
    for i in range(len(inputs.boxes)):
        inst = Instances(tuple(inputs.im_dims[i]))
        inst.pred_boxes = Boxes(inputs.boxes[i].clone())
        inst.pred_classes = inputs.classes[i].clone()
        if with_scores:
            inst.scores = inputs.scores[i].clone()
        instances.append(inst)
    with torch.no_grad():
        out = head._forward_cube(feats, instances, [k.clone() for k in inputs.Ks],
                                 [tuple(d) for d in inputs.im_dims], list(inputs.im_scales_ratio))
    return out


__all__ = ["install_stubs", "reference_modules", "build_reference_head", "run_reference_forward_cube",
           "REFERENCE_ROOT"]
_ = math  # keep linters quiet about the stdlib import used by callers
