"""ORACLE (test infrastructure only) -- CPU fp32 restatement of the reference's lift path.

Follows, op for op and in the same order (so fp32 rounding matches):
  * `ROIHeads3D._forward_cube`, inference slice  -- cubercnn/modeling/roi_heads/roi_heads.py:329-549, 795-848
  * `ROIHeads3D.scale_proposals`                 -- roi_heads.py:310-327 (incl. the width-for-height bug, :317)
  * `CubeHead.forward`                           -- cubercnn/modeling/roi_heads/cube_head.py:148-204
  * `get_cuboid_verts_faces`                     -- cubercnn/util/math_util.py:116-219
  * `compute_virtual_scale_from_focal_spaces`    -- math_util.py:581-592
  * `R_from_allocentric` (tensor branch)         -- math_util.py:651-679
  * `scaled_sigmoid`                             -- math_util.py:969-978
with the detectron2 / pytorch3d pieces taken from `oracle/thirdparty.py`.

Checked against the reference's own code run in this container with stubbed third-party imports
(`tests/golden/make_golden.py` -> `tests/golden/*.npz`) and against the detections logged in
`err.txt:5` (`tests/golden/logged_detections.json`).  PARITY: unpinned by the reference's own
tests -- it has none for this path.

Nothing outside `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline leg may import
this module; the product path never routes through it.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import torch
import torch.nn.functional as F

from . import thirdparty as tp


@dataclass
class LiftConfig:
    """The cfg keys that change the lift's arithmetic."""
    scales: Sequence[float] = (1 / 7, 1 / 14, 1 / 28)
    pooler_resolution: int = 7
    sampling_ratio: int = 0
    num_classes: int = 50
    z_type: str = "direct"
    pose_type: str = "6d"
    cluster_bins: int = 1
    shared_fc: bool = True
    dims_priors_enabled: bool = False
    dims_priors_func: str = "exp"
    use_confidence: float = 1.0
    virtual_depth: bool = True
    virtual_focal: float = 512.0
    allocentric_pose: bool = True
    scale_roi_boxes: float = 0.0


# --------------------------------------------------------------------------- math_util.py
def get_cuboid_verts(box3d: torch.Tensor, R: Optional[torch.Tensor]) -> torch.Tensor:
    """math_util.py:116-193 -- [n,6] (X,Y,Z,W,H,L) + [n,3,3] -> [n,8,3]; x<->L, y<->H, z<->W."""
    n = len(box3d)
    x3d, y3d, z3d = (box3d[:, i].unsqueeze(1) for i in range(3))
    w3d, h3d, l3d = (box3d[:, i].unsqueeze(1) for i in range(3, 6))
    verts = torch.zeros([n, 3, 8], device=box3d.device).float()
    verts[:, 0, [0, 3, 4, 7]] = -l3d / 2
    verts[:, 0, [1, 2, 5, 6]] = l3d / 2
    verts[:, 1, [0, 1, 4, 5]] = -h3d / 2
    verts[:, 1, [2, 3, 6, 7]] = h3d / 2
    verts[:, 2, [0, 1, 2, 3]] = -w3d / 2
    verts[:, 2, [4, 5, 6, 7]] = w3d / 2
    if R is not None:
        verts = R @ verts
    verts[:, 0, :] += x3d
    verts[:, 1, :] += y3d
    verts[:, 2, :] += z3d
    return verts.transpose(1, 2)


def virtual_scale(f, H, f0, H0):
    """math_util.py:581-592."""
    return (H0 * f) / (f0 * H)


def R_from_allocentric(K: torch.Tensor, R_view: torch.Tensor, u: torch.Tensor, v: torch.Tensor) -> torch.Tensor:
    """math_util.py:657-679 (tensor branch)."""
    fx, fy, sx, sy = K[:, 0, 0], K[:, 1, 1], K[:, 0, 2], K[:, 1, 2]
    oray = torch.stack(((u - sx) / fx, (v - sy) / fy, torch.ones_like(u))).T
    oray = oray / torch.linalg.norm(oray, dim=1).unsqueeze(1)
    angle = torch.acos(oray[:, -1])
    axis = torch.zeros_like(oray)
    axis[:, 0] = axis[:, 0] - oray[:, 1]
    axis[:, 1] = axis[:, 1] + oray[:, 0]
    norms = torch.linalg.norm(axis, dim=1)
    valid_angle = angle > 0
    M = tp.axis_angle_to_matrix(angle.unsqueeze(1) * axis / norms.unsqueeze(1))
    R = R_view.clone()
    R[valid_angle] = torch.bmm(M[valid_angle], R_view[valid_angle])
    return R


def scaled_sigmoid(vals, min=0.0, max=1.0):
    """math_util.py:969-978."""
    return min + (max - min) * torch.sigmoid(vals)


# --------------------------------------------------------------------------- cube_head.py
def _fc_stack(w: Dict[str, torch.Tensor], prefix: str, x: torch.Tensor) -> torch.Tensor:
    k = 1
    while f"{prefix}.fc{k}.weight" in w:
        x = F.relu(F.linear(x, w[f"{prefix}.fc{k}.weight"], w[f"{prefix}.fc{k}.bias"]))
        k += 1
    return x


def cube_head_forward(w: Dict[str, torch.Tensor], x: torch.Tensor, cfg: LiftConfig):
    """cube_head.py:148-204.  `w` uses the reference's state-dict keys without the
    `roi_heads.cube_head.` prefix (SURVEY.md §5 "Checkpoint / resume")."""
    n = x.shape[0]
    lin = lambda name, feats: F.linear(feats, w[f"{name}.weight"], w[f"{name}.bias"])  # noqa: E731
    box_uncert = None
    if cfg.shared_fc:
        feats = _fc_stack(w, "feature_generator", x)
        box_2d_deltas = lin("bbox_3D_center_deltas", feats)
        box_dims = lin("bbox_3D_dims", feats)
        box_pose = lin("bbox_3D_pose", feats)
        box_z = lin("bbox_3D_center_depth", feats)
        if cfg.use_confidence:
            box_uncert = lin("bbox_3D_uncertainty", feats).clip(0.01)
    else:
        box_2d_deltas = lin("bbox_3D_center_deltas", _fc_stack(w, "feature_generator_XY", x))
        box_dims = lin("bbox_3D_dims", _fc_stack(w, "feature_generator_dims", x))
        box_pose = lin("bbox_3D_pose", _fc_stack(w, "feature_generator_pose", x))
        box_z = lin("bbox_3D_center_depth", _fc_stack(w, "feature_generator_Z", x))
        if cfg.use_confidence:
            box_uncert = lin("bbox_3D_uncertainty", _fc_stack(w, "feature_generator_conf", x)).clip(0.01)

    if cfg.pose_type == "6d":
        box_pose = tp.rotation_6d_to_matrix(box_pose.view(-1, 6))
    elif cfg.pose_type == "quaternion":
        quats = box_pose.view(-1, 4)
        quats_scales = (quats * quats).sum(1)
        quats = quats / tp._copysign(torch.sqrt(quats_scales), quats[:, 0])[:, None]
        box_pose = tp.quaternion_to_matrix(quats)
    elif cfg.pose_type == "euler":
        box_pose = tp.euler_angles_to_matrix(box_pose.view(-1, 3), "XYZ")
    else:
        raise ValueError("Cuboid pose type {} is not recognized".format(cfg.pose_type))

    C = cfg.num_classes
    if cfg.dims_priors_enabled:
        box_2d_deltas = box_2d_deltas.view(n, C, 2)
        box_dims = box_dims.view(n, C, 3)
        box_pose = box_pose.view(n, C, 3, 3)
    if cfg.cluster_bins > 1:
        box_z = box_z.view(n, cfg.cluster_bins, C, -1) if cfg.dims_priors_enabled else box_z.view(n, cfg.cluster_bins, -1)
    else:
        box_z = box_z.view(n, C, -1) if cfg.dims_priors_enabled else box_z.view(n, -1)
    return box_2d_deltas, box_z, box_dims, box_pose, box_uncert


# --------------------------------------------------------------------------- roi_heads.py
def scale_proposals(boxes: List[torch.Tensor], scale_roi_boxes: float) -> List[torch.Tensor]:
    """roi_heads.py:310-327.  Heights use x2-x1 exactly as the reference does (:317)."""
    if not scale_roi_boxes > 0:
        return boxes
    out = []
    for b in boxes:
        centers = (b[:, :2] + b[:, 2:]) / 2
        widths = b[:, 2] - b[:, 0]
        heights = b[:, 2] - b[:, 0]
        x1 = centers[:, 0] - 0.5 * widths * scale_roi_boxes
        x2 = centers[:, 0] + 0.5 * widths * scale_roi_boxes
        y1 = centers[:, 1] - 0.5 * heights * scale_roi_boxes
        y2 = centers[:, 1] + 0.5 * heights * scale_roi_boxes
        out.append(torch.stack([x1, y1, x2, y2], dim=1))
    return out


@dataclass
class LiftOutput:
    levels: torch.Tensor = None            # [n] int64, 0-based pyramid level per box
    pooled: torch.Tensor = None            # [n, C, P, P]
    head_raw: Dict[str, torch.Tensor] = field(default_factory=dict)
    scores: List[torch.Tensor] = field(default_factory=list)
    pred_bbox3D: List[torch.Tensor] = field(default_factory=list)
    pred_center_cam: List[torch.Tensor] = field(default_factory=list)
    pred_center_2D: List[torch.Tensor] = field(default_factory=list)
    pred_dimensions: List[torch.Tensor] = field(default_factory=list)
    pred_pose: List[torch.Tensor] = field(default_factory=list)


def forward_cube(
    features: List[torch.Tensor],
    boxes: List[torch.Tensor],
    classes: List[torch.Tensor],
    scores: Optional[List[torch.Tensor]],
    Ks: List[torch.Tensor],
    im_current_dims: List[Sequence[int]],
    im_scales_ratio: List[float],
    weights: Dict[str, torch.Tensor],
    cfg: LiftConfig,
    priors_dims_per_cat: Optional[torch.Tensor] = None,   # [1, num_classes, 2, 3]
    priors_z_scales: Optional[torch.Tensor] = None,       # [num_classes, cluster_bins]
    priors_z_stats: Optional[torch.Tensor] = None,        # [num_classes, cluster_bins, 2]
) -> LiftOutput:
    """roi_heads.py:329-549 + 795-848, eval branch."""
    out = LiftOutput()
    device = features[0].device
    box_classes = torch.cat(classes)
    pooler = tp.ROIPoolerOracle(cfg.pooler_resolution, cfg.scales, cfg.sampling_ratio, "ROIAlignV2")
    boxes_scaled = scale_proposals(boxes, cfg.scale_roi_boxes)
    pooled = pooler(features, boxes_scaled)
    out.pooled = pooled
    if len(cfg.scales) > 1 and sum(len(b) for b in boxes):
        out.levels = pooler.levels(boxes_scaled)
    else:
        out.levels = torch.zeros(sum(len(b) for b in boxes), dtype=torch.int64)
    cube_features = pooled.flatten(1)
    n = cube_features.shape[0]
    if n == 0:
        return out
    num_boxes_per_image = [len(b) for b in boxes]

    # roi_heads.py:378-400
    Ks_scaled_per_box = torch.cat([
        (Ks[i] / im_scales_ratio[i]).unsqueeze(0).repeat([num, 1, 1]) for (i, num) in enumerate(num_boxes_per_image)
    ]).to(device)
    Ks_scaled_per_box[:, -1, -1] = 1
    focal_lengths_per_box = torch.cat([
        (Ks[i][1, 1]).unsqueeze(0).repeat([num]) for (i, num) in enumerate(num_boxes_per_image)
    ]).to(device)
    im_ratios_per_box = torch.cat([
        torch.FloatTensor([im_scales_ratio[i]]).repeat(num) for (i, num) in enumerate(num_boxes_per_image)
    ]).to(device)
    im_scales_per_box = torch.cat([
        torch.FloatTensor([im_current_dims[i][0]]).repeat(num) for (i, num) in enumerate(num_boxes_per_image)
    ]).to(device)
    im_scales_original_per_box = im_scales_per_box * im_ratios_per_box
    if cfg.virtual_depth:
        virtual_to_real = virtual_scale(focal_lengths_per_box, im_scales_original_per_box,
                                        cfg.virtual_focal, im_scales_per_box)
    else:
        virtual_to_real = 1.0

    # roi_heads.py:414-419
    src_boxes = torch.cat(boxes, dim=0)
    src_widths = src_boxes[:, 2] - src_boxes[:, 0]
    src_heights = src_boxes[:, 3] - src_boxes[:, 1]
    src_scales = (src_heights ** 2 + src_widths ** 2).sqrt()
    src_ctr_x = src_boxes[:, 0] + 0.5 * src_widths
    src_ctr_y = src_boxes[:, 1] + 0.5 * src_heights

    cube_2d_deltas, cube_z, cube_dims, cube_pose, cube_uncert = cube_head_forward(weights, cube_features, cfg)
    out.head_raw = dict(deltas=cube_2d_deltas, z=cube_z, dims=cube_dims, pose=cube_pose, uncert=cube_uncert)
    fg_inds = torch.arange(n)

    # roi_heads.py:436-477
    assignments = None
    if cube_z is not None and cfg.cluster_bins > 1:
        scales_diff = (priors_z_scales.detach().T.unsqueeze(0) - src_scales.unsqueeze(1).unsqueeze(2)).abs()
        assignments = scales_diff.argmin(1)
        if cfg.dims_priors_enabled:
            cube_z = cube_z[fg_inds, :, box_classes, :][fg_inds, assignments[fg_inds, box_classes]]
        else:
            cube_z = cube_z[fg_inds, :, :][fg_inds, assignments[fg_inds]]
    elif cube_z is not None:
        cube_z = cube_z[fg_inds, box_classes, :] if cfg.dims_priors_enabled else cube_z[fg_inds, :]
    if cfg.dims_priors_enabled:
        cube_dims = cube_dims[fg_inds, box_classes, :]
        cube_pose = cube_pose[fg_inds, box_classes, :, :]
    if cfg.use_confidence:
        cube_uncert = cube_uncert[fg_inds, box_classes] if cfg.dims_priors_enabled else cube_uncert[fg_inds, 0]
    if cfg.dims_priors_enabled:
        cube_2d_deltas = cube_2d_deltas[fg_inds, box_classes, :]

    # roi_heads.py:480-507
    cube_x = src_ctr_x + src_widths * cube_2d_deltas[:, 0]
    cube_y = src_ctr_y + src_heights * cube_2d_deltas[:, 1]
    cube_xy = torch.cat((cube_x.unsqueeze(1), cube_y.unsqueeze(1)), dim=1)
    cube_dims_norm = cube_dims
    if cfg.dims_priors_enabled:
        prior_dims = priors_dims_per_cat.detach().repeat([n, 1, 1, 1])[fg_inds, box_classes]
        prior_dims_mean = prior_dims[:, 0, :]
        prior_dims_std = prior_dims[:, 1, :]
        if cfg.dims_priors_func == "sigmoid":
            prior_dims_min = (prior_dims_mean - 3 * prior_dims_std).clip(0.0)
            prior_dims_max = (prior_dims_mean + 3 * prior_dims_std)
            cube_dims = scaled_sigmoid(cube_dims_norm, min=prior_dims_min, max=prior_dims_max)
        elif cfg.dims_priors_func == "exp":
            cube_dims = torch.exp(cube_dims_norm.clip(max=5)) * prior_dims_mean
    else:
        cube_dims = torch.exp(cube_dims_norm.clip(max=5))

    # roi_heads.py:509-549
    if cfg.allocentric_pose:
        cube_pose = R_from_allocentric(Ks_scaled_per_box, cube_pose, u=cube_x.detach(), v=cube_y.detach())
    cube_z = cube_z.squeeze()
    if cfg.z_type == "sigmoid":
        cube_z = torch.sigmoid(cube_z) * 100
    elif cfg.z_type == "log":
        cube_z = torch.exp(cube_z)
    elif cfg.z_type == "clusters":
        z_means = priors_z_stats[:, :, 0].T.unsqueeze(0).repeat([n, 1, 1])
        z_means = torch.gather(z_means, 1, assignments.unsqueeze(1)).squeeze(1)
        z_stds = priors_z_stats[:, :, 1].T.unsqueeze(0).repeat([n, 1, 1])
        z_stds = torch.gather(z_stds, 1, assignments.unsqueeze(1)).squeeze(1)
        z_means = z_means[fg_inds, box_classes]
        z_stds = z_stds[fg_inds, box_classes]
        z_mins = (z_means - 3 * z_stds).clip(0)
        z_maxs = (z_means + 3 * z_stds)
        cube_z = scaled_sigmoid(cube_z, min=z_mins, max=z_maxs)
    if cfg.virtual_depth:
        cube_z = cube_z * virtual_to_real

    # roi_heads.py:798-808
    if len(cube_z.shape) == 0:
        cube_z = cube_z.unsqueeze(0)
    cube_x3d = cube_z * (cube_x - Ks_scaled_per_box[:, 0, 2]) / Ks_scaled_per_box[:, 0, 0]
    cube_y3d = cube_z * (cube_y - Ks_scaled_per_box[:, 1, 2]) / Ks_scaled_per_box[:, 1, 1]
    cube_3D = torch.cat((torch.stack((cube_x3d, cube_y3d, cube_z)).T, cube_dims,
                         cube_xy * im_ratios_per_box.unsqueeze(1)), dim=1)
    if cfg.use_confidence:
        cube_conf = torch.exp(-cube_uncert)
        cube_3D = torch.cat((cube_3D, cube_conf.unsqueeze(1)), dim=1)

    # roi_heads.py:811-843
    cube_3D_split = cube_3D.split(num_boxes_per_image)
    cube_pose_split = cube_pose.split(num_boxes_per_image)
    for i, (cube_3D_i, cube_pose_i) in enumerate(zip(cube_3D_split, cube_pose_split)):
        if scores is not None:
            out.scores.append((scores[i] * cube_3D_i[:, -1]) ** (1 / 2))
        else:
            out.scores.append(cube_3D_i[:, -1])
        out.pred_bbox3D.append(get_cuboid_verts(cube_3D_i[:, :6], cube_pose_i))
        out.pred_center_cam.append(cube_3D_i[:, :3])
        out.pred_center_2D.append(cube_3D_i[:, 6:8])
        out.pred_dimensions.append(cube_3D_i[:, 3:6])
        out.pred_pose.append(cube_pose_i)
    return out
