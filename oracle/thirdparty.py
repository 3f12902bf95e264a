"""ORACLE (test infrastructure only) -- CPU restatement of the un-vendored third-party pieces.

The reference's lift path calls into two dependencies whose sources are absent from
`/root/reference` (empty submodule directories, `.SUBMODULES.json:9-43`):

* detectron2, fork `yaojin17/detectron2` @ 8ece38992c987b0877a45deeb662ec1d1dd069a4 (`setup.sh:2`):
  `modeling/poolers.py` (`ROIPooler`, `assign_boxes_to_levels`, `convert_boxes_to_pooler_format`)
  and `layers/roi_align.py` (`ROIAlign`, a wrapper over `torchvision.ops.roi_align`).
  Call sites: `cubercnn/modeling/roi_heads/roi_heads.py:12,166-171,366`.
* pytorch3d @ 055ab3a2e3e611dff66fa82f632e62a315f3b5e7 (`setup.sh:1`):
  `transforms/rotation_conversions.py` (`rotation_6d_to_matrix`, `axis_angle_to_matrix`,
  `axis_angle_to_quaternion`, `quaternion_to_matrix`, `euler_angles_to_matrix`, `_copysign`).
  Call sites: `cube_head.py:10-15,177-186`, `cubercnn/util/math_util.py:34,676`.

Their published algorithms are restated here in plain torch, in the same op order so fp32
rounding matches.  The one deliberate deviation: upstream `ROIPooler.__init__` asserts that the
strides are powers of two; the shipped DINOv2 pyramid uses 1/7, 1/14, 1/28 and the logged model
proves a 3-level pooler was built with those scales (`nohup.out:676-682`), so the fork must have
dropped that assert -- `min_level = int(log2 7) = 2`, `max_level = int(log2 28) = 4`.

PARITY: unpinned by the reference's own tests (it has none).  Nothing outside
`tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline leg may import this module.
"""
from __future__ import annotations

import math
from typing import List, Sequence

import torch
import torch.nn.functional as F
import torchvision  # noqa: F401  (registers torch.ops.torchvision.roi_align)


# --------------------------------------------------------------------------- detectron2: poolers
def boxes_area(boxes: torch.Tensor) -> torch.Tensor:
    """d2 `Boxes.area`: (x2-x1)*(y2-y1)."""
    return (boxes[:, 2] - boxes[:, 0]) * (boxes[:, 3] - boxes[:, 1])


def assign_boxes_to_levels(
    box_tensors: Sequence[torch.Tensor],
    min_level: int,
    max_level: int,
    canonical_box_size: int = 224,
    canonical_level: int = 4,
) -> torch.Tensor:
    """d2 `assign_boxes_to_levels`: FPN eqn (1) with the 1e-8 *inside* the log, result 0-based."""
    box_sizes = torch.sqrt(torch.cat([boxes_area(b) for b in box_tensors]))
    level_assignments = torch.floor(canonical_level + torch.log2(box_sizes / canonical_box_size + 1e-8))
    level_assignments = torch.clamp(level_assignments, min=min_level, max=max_level)
    return level_assignments.to(torch.int64) - min_level


def convert_boxes_to_pooler_format(box_tensors: Sequence[torch.Tensor]) -> torch.Tensor:
    """d2 `convert_boxes_to_pooler_format`: [n,5] = (batch index, x1, y1, x2, y2)."""
    boxes = torch.cat(list(box_tensors), dim=0)
    sizes = torch.tensor([len(b) for b in box_tensors], device=boxes.device)
    indices = torch.repeat_interleave(torch.arange(len(box_tensors), dtype=boxes.dtype, device=boxes.device), sizes)
    return torch.cat([indices[:, None], boxes], dim=1)


def roi_align_v2(feat: torch.Tensor, rois: torch.Tensor, output_size: int, spatial_scale: float,
                 sampling_ratio: int) -> torch.Tensor:
    """d2 `ROIAlign(aligned=True)` == torchvision `roi_align(..., aligned=True)`."""
    return torch.ops.torchvision.roi_align(
        feat, rois.to(dtype=feat.dtype), spatial_scale, output_size, output_size, sampling_ratio, True
    )


class ROIPoolerOracle:
    """d2 `ROIPooler(output_size, scales, sampling_ratio, "ROIAlignV2")`, inference semantics."""

    def __init__(self, output_size: int, scales: Sequence[float], sampling_ratio: int,
                 pooler_type: str = "ROIAlignV2", canonical_box_size: int = 224, canonical_level: int = 4):
        assert pooler_type == "ROIAlignV2", "only the pooler type the reference ships is restated"
        self.output_size = int(output_size)
        self.scales = tuple(float(s) for s in scales)
        self.sampling_ratio = int(sampling_ratio)
        min_level = -(math.log2(self.scales[0]))
        max_level = -(math.log2(self.scales[-1]))
        self.min_level = int(min_level)
        self.max_level = int(max_level)
        assert len(self.scales) == self.max_level - self.min_level + 1, \
            "[ROIPooler] Sizes of input featuremaps do not form a pyramid!"
        assert 0 <= self.min_level <= self.max_level
        self.canonical_level = canonical_level
        assert canonical_box_size > 0
        self.canonical_box_size = canonical_box_size

    def levels(self, box_tensors: Sequence[torch.Tensor]) -> torch.Tensor:
        return assign_boxes_to_levels(box_tensors, self.min_level, self.max_level,
                                      self.canonical_box_size, self.canonical_level)

    def __call__(self, x: List[torch.Tensor], box_tensors: Sequence[torch.Tensor]) -> torch.Tensor:
        num_level_assignments = len(self.scales)
        assert len(x) == num_level_assignments and len(box_tensors) == x[0].size(0)
        if len(box_tensors) == 0:
            return x[0].new_zeros((0, x[0].shape[1], self.output_size, self.output_size))
        pooler_fmt_boxes = convert_boxes_to_pooler_format(box_tensors)
        if num_level_assignments == 1:
            return roi_align_v2(x[0], pooler_fmt_boxes, self.output_size, self.scales[0], self.sampling_ratio)
        level_assignments = self.levels(box_tensors)
        output = x[0].new_zeros((pooler_fmt_boxes.size(0), x[0].shape[1], self.output_size, self.output_size))
        for level, scale in enumerate(self.scales):
            inds = torch.nonzero(level_assignments == level, as_tuple=True)[0]
            pooled = roi_align_v2(x[level], pooler_fmt_boxes[inds], self.output_size, scale, self.sampling_ratio)
            output.index_put_((inds,), pooled)
        return output


# --------------------------------------------------------------------------- pytorch3d: transforms
def rotation_6d_to_matrix(d6: torch.Tensor) -> torch.Tensor:
    """p3d `rotation_6d_to_matrix` (Zhou et al. 6D -> R by Gram-Schmidt, basis vectors as rows)."""
    a1, a2 = d6[..., :3], d6[..., 3:]
    b1 = F.normalize(a1, dim=-1)
    b2 = a2 - (b1 * a2).sum(-1, keepdim=True) * b1
    b2 = F.normalize(b2, dim=-1)
    b3 = torch.cross(b1, b2, dim=-1)
    return torch.stack((b1, b2, b3), dim=-2)


def quaternion_to_matrix(quaternions: torch.Tensor) -> torch.Tensor:
    """p3d `quaternion_to_matrix` (real part first)."""
    r, i, j, k = torch.unbind(quaternions, -1)
    two_s = 2.0 / (quaternions * quaternions).sum(-1)
    o = torch.stack(
        (
            1 - two_s * (j * j + k * k), two_s * (i * j - k * r), two_s * (i * k + j * r),
            two_s * (i * j + k * r), 1 - two_s * (i * i + k * k), two_s * (j * k - i * r),
            two_s * (i * k - j * r), two_s * (j * k + i * r), 1 - two_s * (i * i + j * j),
        ),
        -1,
    )
    return o.reshape(quaternions.shape[:-1] + (3, 3))


def axis_angle_to_quaternion(axis_angle: torch.Tensor) -> torch.Tensor:
    """p3d `axis_angle_to_quaternion`: q = (cos(t/2), axis*sin(t/2)), Taylor branch for |t|<1e-6."""
    angles = torch.norm(axis_angle, p=2, dim=-1, keepdim=True)
    half_angles = angles * 0.5
    eps = 1e-6
    small_angles = angles.abs() < eps
    sin_half_angles_over_angles = torch.empty_like(angles)
    sin_half_angles_over_angles[~small_angles] = torch.sin(half_angles[~small_angles]) / angles[~small_angles]
    sin_half_angles_over_angles[small_angles] = 0.5 - (angles[small_angles] * angles[small_angles]) / 48
    return torch.cat([torch.cos(half_angles), axis_angle * sin_half_angles_over_angles], dim=-1)


def axis_angle_to_matrix(axis_angle: torch.Tensor) -> torch.Tensor:
    """p3d `axis_angle_to_matrix` (default path: through the quaternion)."""
    return quaternion_to_matrix(axis_angle_to_quaternion(axis_angle))


def _copysign(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """p3d `_copysign`: |a| with b's sign, where sign(0) counts as positive."""
    signs_differ = (a < 0) != (b < 0)
    return torch.where(signs_differ, -a, a)


def _axis_angle_rotation(axis: str, angle: torch.Tensor) -> torch.Tensor:
    cos, sin = torch.cos(angle), torch.sin(angle)
    one, zero = torch.ones_like(angle), torch.zeros_like(angle)
    if axis == "X":
        flat = (one, zero, zero, zero, cos, -sin, zero, sin, cos)
    elif axis == "Y":
        flat = (cos, zero, sin, zero, one, zero, -sin, zero, cos)
    elif axis == "Z":
        flat = (cos, -sin, zero, sin, cos, zero, zero, zero, one)
    else:
        raise ValueError("letter must be either X, Y or Z.")
    return torch.stack(flat, -1).reshape(angle.shape + (3, 3))


def euler_angles_to_matrix(euler_angles: torch.Tensor, convention: str) -> torch.Tensor:
    """p3d `euler_angles_to_matrix`: R = R_c0(a0) @ R_c1(a1) @ R_c2(a2)."""
    mats = [_axis_angle_rotation(c, e) for c, e in zip(convention, torch.unbind(euler_angles, -1))]
    return torch.matmul(torch.matmul(mats[0], mats[1]), mats[2])
