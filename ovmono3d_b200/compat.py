"""Thin stand-ins for the detectron2 names the lift path touches.

The reference imports these from detectron2 (`cubercnn/modeling/roi_heads/roi_heads.py:12-20`,
`cube_head.py:2-4`).  detectron2 is not vendored in the reference checkout and is not installed in
this image, so the plugin ships the minimal surface it needs.  When the real detectron2 *is*
importable the real objects are used instead, so the head registers into the real
`ROI_HEADS_REGISTRY` and consumes real `Instances`/`Boxes`.

Only behaviour that the lift path relies on is implemented:
  Registry.register/get, Instances (field bag with image_size/__len__/has/to/fields),
  Boxes (.tensor, area, get_centers, to, __len__), ShapeSpec, configurable/from_config.
"""
from __future__ import annotations

import functools
import inspect
from dataclasses import dataclass
from typing import Any, Dict, Iterator, List, Optional, Tuple

import torch

try:  # pragma: no cover - exercised only where detectron2 exists
    from detectron2.config import configurable  # type: ignore
    from detectron2.layers import ShapeSpec  # type: ignore
    from detectron2.modeling.roi_heads import ROI_HEADS_REGISTRY  # type: ignore
    from detectron2.structures import Boxes, Instances  # type: ignore
    from detectron2.utils.registry import Registry  # type: ignore

    HAVE_DETECTRON2 = True
except Exception:  # ImportError, or a broken install
    HAVE_DETECTRON2 = False

    class Registry:
        """name -> object map with decorator registration (d2 `fvcore.common.registry.Registry`)."""

        def __init__(self, name: str) -> None:
            self._name = name
            self._obj_map: Dict[str, Any] = {}

        def _do_register(self, name: str, obj: Any) -> None:
            assert name not in self._obj_map, (
                "An object named '{}' was already registered in '{}' registry!".format(name, self._name)
            )
            self._obj_map[name] = obj

        def register(self, obj: Any = None) -> Any:
            if obj is None:
                def deco(func_or_class: Any) -> Any:
                    self._do_register(func_or_class.__name__, func_or_class)
                    return func_or_class
                return deco
            self._do_register(obj.__name__, obj)
            return obj

        def get(self, name: str) -> Any:
            ret = self._obj_map.get(name)
            if ret is None:
                raise KeyError("No object named '{}' found in '{}' registry!".format(name, self._name))
            return ret

        def __contains__(self, name: str) -> bool:
            return name in self._obj_map

        def __iter__(self) -> Iterator[Tuple[str, Any]]:
            return iter(self._obj_map.items())

    ROI_HEADS_REGISTRY = Registry("ROI_HEADS")

    @dataclass
    class ShapeSpec:
        channels: Optional[int] = None
        height: Optional[int] = None
        width: Optional[int] = None
        stride: Optional[int] = None

    class Boxes:
        """[k,4] xyxy float boxes (d2 `structures/boxes.py`)."""

        def __init__(self, tensor: torch.Tensor):
            if not isinstance(tensor, torch.Tensor):
                tensor = torch.as_tensor(tensor, dtype=torch.float32, device=torch.device("cpu"))
            else:
                tensor = tensor.to(torch.float32)
            if tensor.numel() == 0:
                tensor = tensor.reshape((-1, 4)).to(dtype=torch.float32)
            assert tensor.dim() == 2 and tensor.size(-1) == 4, tensor.size()
            self.tensor = tensor

        def clone(self) -> "Boxes":
            return Boxes(self.tensor.clone())

        def to(self, *args: Any, **kwargs: Any) -> "Boxes":
            return Boxes(self.tensor.to(*args, **kwargs))

        def area(self) -> torch.Tensor:
            box = self.tensor
            return (box[:, 2] - box[:, 0]) * (box[:, 3] - box[:, 1])

        def get_centers(self) -> torch.Tensor:
            return (self.tensor[:, :2] + self.tensor[:, 2:]) / 2

        def scale(self, scale_x: float, scale_y: float) -> None:
            self.tensor[:, 0::2] *= scale_x
            self.tensor[:, 1::2] *= scale_y

        def clip(self, box_size: Tuple[int, int]) -> None:
            assert torch.isfinite(self.tensor).all(), "Box tensor contains infinite or NaN!"
            h, w = box_size
            x1 = self.tensor[:, 0].clamp(min=0, max=w)
            y1 = self.tensor[:, 1].clamp(min=0, max=h)
            x2 = self.tensor[:, 2].clamp(min=0, max=w)
            y2 = self.tensor[:, 3].clamp(min=0, max=h)
            self.tensor = torch.stack((x1, y1, x2, y2), dim=-1)

        def __getitem__(self, item: Any) -> "Boxes":
            if isinstance(item, int):
                return Boxes(self.tensor[item].view(1, -1))
            b = self.tensor[item]
            assert b.dim() == 2, "Indexing on Boxes with {} failed to return a matrix!".format(item)
            return Boxes(b)

        def __len__(self) -> int:
            return self.tensor.shape[0]

        def __repr__(self) -> str:
            return "Boxes(" + str(self.tensor) + ")"

        @property
        def device(self) -> torch.device:
            return self.tensor.device

        @classmethod
        def cat(cls, boxes_list: List["Boxes"]) -> "Boxes":
            if len(boxes_list) == 0:
                return cls(torch.empty(0))
            return cls(torch.cat([b.tensor for b in boxes_list], dim=0))

    class Instances:
        """Per-image field bag (d2 `structures/instances.py`)."""

        def __init__(self, image_size: Tuple[int, int], **kwargs: Any):
            object.__setattr__(self, "_image_size", image_size)
            object.__setattr__(self, "_fields", {})
            for k, v in kwargs.items():
                self.set(k, v)

        @property
        def image_size(self) -> Tuple[int, int]:
            return self._image_size

        def __setattr__(self, name: str, val: Any) -> None:
            if name.startswith("_"):
                object.__setattr__(self, name, val)
            else:
                self.set(name, val)

        def __getattr__(self, name: str) -> Any:
            if name == "_fields" or name not in self._fields:
                raise AttributeError("Cannot find field '{}' in the given Instances!".format(name))
            return self._fields[name]

        def set(self, name: str, value: Any) -> None:
            data_len = len(value)
            if len(self._fields):
                assert len(self) == data_len, (
                    "Adding a field of length {} to a Instances of length {}".format(data_len, len(self))
                )
            self._fields[name] = value

        def has(self, name: str) -> bool:
            return name in self._fields

        def remove(self, name: str) -> None:
            del self._fields[name]

        def get(self, name: str) -> Any:
            return self._fields[name]

        def get_fields(self) -> Dict[str, Any]:
            return self._fields

        def to(self, *args: Any, **kwargs: Any) -> "Instances":
            ret = Instances(self._image_size)
            for k, v in self._fields.items():
                if hasattr(v, "to"):
                    v = v.to(*args, **kwargs)
                ret.set(k, v)
            return ret

        def __getitem__(self, item: Any) -> "Instances":
            if type(item) == int:
                if item >= len(self) or item < -len(self):
                    raise IndexError("Instances index out of range!")
                item = slice(item, None, len(self))
            ret = Instances(self._image_size)
            for k, v in self._fields.items():
                ret.set(k, v[item])
            return ret

        def __len__(self) -> int:
            for v in self._fields.values():
                return v.__len__()
            raise NotImplementedError("Empty Instances does not support __len__!")

        def __iter__(self):
            raise NotImplementedError("`Instances` object is not iterable!")

        def __repr__(self) -> str:
            s = self.__class__.__name__ + "("
            s += "num_instances={}, ".format(len(self) if self._fields else 0)
            s += "image_height={}, image_width={}, ".format(self._image_size[0], self._image_size[1])
            s += "fields=[{}])".format(", ".join("{}: {}".format(k, type(v).__name__) for k, v in self._fields.items()))
            return s

    def _called_with_cfg(*args: Any, **kwargs: Any) -> bool:
        from .config import CfgNode

        if len(args) and isinstance(args[0], CfgNode):
            return True
        if isinstance(kwargs.pop("cfg", None), CfgNode):
            return True
        return False

    def _get_args_from_config(from_config_func: Any, *args: Any, **kwargs: Any) -> Dict[str, Any]:
        signature = inspect.signature(from_config_func)
        if list(signature.parameters.keys())[0] != "cfg":
            raise TypeError("from_config() must take 'cfg' as the first argument!")
        support_var_arg = any(
            p.kind in [p.VAR_POSITIONAL, p.VAR_KEYWORD] for p in signature.parameters.values()
        )
        if support_var_arg:
            return from_config_func(*args, **kwargs)
        supported = set(signature.parameters.keys())
        extra = {k: kwargs.pop(k) for k in list(kwargs.keys()) if k not in supported}
        ret = from_config_func(*args, **kwargs)
        ret.update(extra)
        return ret

    def configurable(init_func: Any = None, *, from_config: Any = None) -> Any:
        """`__init__` decorator: `Cls(cfg, ...)` is routed through `Cls.from_config(cfg, ...)`."""
        assert init_func is not None and from_config is None, "only the __init__ form is supported"
        assert inspect.isfunction(init_func) and init_func.__name__ == "__init__"

        @functools.wraps(init_func)
        def wrapped(self: Any, *args: Any, **kwargs: Any) -> None:
            try:
                from_config_func = type(self).from_config
            except AttributeError as e:
                raise AttributeError("Class with @configurable must have a 'from_config' classmethod.") from e
            if not inspect.ismethod(from_config_func):
                raise TypeError("Class with @configurable must have a 'from_config' classmethod.")
            if _called_with_cfg(*args, **kwargs):
                explicit_args = _get_args_from_config(from_config_func, *args, **kwargs)
                init_func(self, **explicit_args)
            else:
                init_func(self, *args, **kwargs)

        return wrapped


ROI_CUBE_HEAD_REGISTRY = Registry("ROI_CUBE_HEAD_B200")
"""Registry for the cube head (`cube_head.py:18` keeps its own, separate from detectron2's)."""

__all__ = [
    "HAVE_DETECTRON2",
    "Registry",
    "ROI_HEADS_REGISTRY",
    "ROI_CUBE_HEAD_REGISTRY",
    "ShapeSpec",
    "Boxes",
    "Instances",
    "configurable",
]
