"""ORACLE (test infrastructure only) -- exact 3-D IoU of cuboids as the reference's evaluator computes it.

Restates
  * `box3d_overlap`, `_check_coplanar`, `_check_nonzero`  -- cubercnn/evaluation/omni3d_evaluation.py:68-169 (op for op,
    torch CPU ops in the reference's order, including the quirk that `_check_coplanar` sums the out-of-plane distances of
    the six faces before taking the absolute value, :86-89);
  * `pytorch3d._C.iou_box3d` (omni3d_evaluation.py:38,158) -- pytorch3d @ 055ab3a (setup.sh:1) is an un-vendored
    submodule, so this is the PUBLISHED algorithm of pytorch3d/csrc/iou_box3d/{iou_utils.h, iou_box3d_cpu.cpp} restated
    from its description (Mesh R-CNN "exact IoU of oriented boxes"): every triangle of one box is clipped by the six face
    planes of the other (triangle in / out / split into one or two triangles), triangles of the second box that are
    coplanar with a surviving triangle of the first are dropped, the polyhedron made of all surviving triangles is summed
    as |tetrahedra| around its vertex mean, iou = vol / (vol1 + vol2 - vol).

PARITY: unpinned -- the reference has no tests or vectors for this function and pytorch3d cannot be built here.  What
pins this file instead (tests/test_iou3d.py): analytic known answers (identical, disjoint, face-touching, nested, half
overlap, 45-degree rotation = 1/sqrt(2)) and, on random oriented boxes, agreement to 1e-5 with an independent exact
construction (scipy's qhull: half-space intersection -> convex hull volume).  Known property of the published
algorithm, reproduced here: its coplanarity tolerance dEpsilon = 1e-3 makes it keep whole a sliver triangle that lies
within 1e-3 of a clipping plane, so on about one random pair in 700 its IoU is off the exact value by up to a few
1e-3 (tests/test_iou3d.py checks those pairs against qhull).

Nothing outside `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline leg may import this module.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

# pytorch3d/ops/iou_box3d.py (imported by the reference at omni3d_evaluation.py:41)
BOX_PLANES = [[0, 1, 2, 3], [3, 2, 6, 7], [0, 1, 5, 4], [0, 3, 7, 4], [1, 2, 6, 5], [4, 5, 6, 7]]
BOX_TRIANGLES = [[0, 1, 2], [0, 3, 2], [4, 5, 6], [4, 6, 7], [1, 5, 6], [1, 6, 2], [0, 4, 7], [0, 7, 3], [3, 2, 6], [3, 6, 7],
                 [0, 1, 5], [0, 4, 5]]
K_EPS = 1e-8      # kEpsilon
D_EPS = 1e-3      # dEpsilon: coplanarity
A_EPS = 1e-4      # aEpsilon: area of a triangle that still counts


# --------------------------------------------------------------------------- omni3d_evaluation.py:68-107
def check_coplanar(boxes: torch.Tensor, eps: float = 1e-4) -> torch.Tensor:
    faces = torch.tensor(BOX_PLANES, dtype=torch.int64, device=boxes.device)
    verts = boxes.index_select(index=faces.view(-1), dim=1)
    B = boxes.shape[0]
    P, V = faces.shape
    v0, v1, v2, v3 = verts.reshape(B, P, V, 3).unbind(2)
    e0 = F.normalize(v1 - v0, dim=-1)
    e1 = F.normalize(v2 - v0, dim=-1)
    normal = F.normalize(torch.cross(e0, e1, dim=-1), dim=-1)
    mat1 = (v3 - v0).view(B, 1, -1)
    mat2 = normal.view(B, -1, 1)
    return (mat1.bmm(mat2).abs() < eps).view(B)


def check_nonzero(boxes: torch.Tensor, eps: float = 1e-8) -> torch.Tensor:
    faces = torch.tensor(BOX_TRIANGLES, dtype=torch.int64, device=boxes.device)
    verts = boxes.index_select(index=faces.view(-1), dim=1)
    B = boxes.shape[0]
    T, V = faces.shape
    v0, v1, v2 = verts.reshape(B, T, V, 3).unbind(2)
    normals = torch.cross(v1 - v0, v2 - v0, dim=-1)
    face_areas = normals.norm(dim=-1) / 2
    return (face_areas > eps).all(1).view(B)


# --------------------------------------------------------------------------- pytorch3d iou_utils.h (published algorithm)
def _unit(v: np.ndarray) -> np.ndarray:
    return v / max(float(np.linalg.norm(v)), K_EPS)


def _tri_normal(tri) -> np.ndarray:
    return _unit(np.cross(tri[1] - tri[0], tri[2] - tri[0]))


def _plane_normal(plane) -> np.ndarray:
    """Unit normal of a (planar) quad: of its two edge pairs the one closer to orthogonal (robust for skewed quads)."""
    v0, v1, v2, v3 = plane
    best, n = 2.0, None
    for a, b in ((v1 - v0, v3 - v0), (v1 - v0, v2 - v0), (v2 - v0, v3 - v0)):
        c = abs(float(np.dot(_unit(a), _unit(b))))
        if c < best:
            best, n = c, np.cross(a, b)
    return _unit(n)


def _plane_center(plane) -> np.ndarray:
    return (plane[0] + plane[1] + plane[2] + plane[3]) / 4.0


def _plane_inward_normal(plane, box_center) -> np.ndarray:
    n = _plane_normal(plane)
    if float(np.dot(box_center - _plane_center(plane), n)) < 0.0:
        n = -n
    return n


def _is_inside(plane_ctr, n, p) -> bool:
    return float(np.dot(p - plane_ctr, n)) >= 0.0


def _edge_plane(plane_ctr, n, p0, p1) -> np.ndarray:
    direc = _unit(p1 - p0)
    top = -float(np.dot(p0 - plane_ctr, n))
    bot = float(np.dot(direc, n))
    return p0 + (top / bot) * direc if abs(bot) >= K_EPS else p1.copy()


def _tri_coplanar_with_plane(tri, plane_ctr, n) -> bool:
    return all(abs(float(np.dot(v - plane_ctr, n))) < D_EPS for v in tri)


def _clip_tri(plane_ctr, n, tri) -> List[Tuple[np.ndarray, np.ndarray, np.ndarray]]:
    if _tri_coplanar_with_plane(tri, plane_ctr, n):
        return [tri]
    v0, v1, v2 = tri
    i0, i1, i2 = (_is_inside(plane_ctr, n, v) for v in tri)
    if i0 and i1 and i2:
        return [tri]
    if not (i0 or i1 or i2):
        return []

    def one_out(vout, vin1, vin2):
        p1, p2 = _edge_plane(plane_ctr, n, vin1, vout), _edge_plane(plane_ctr, n, vin2, vout)
        return [(vin1, p1, p2), (vin1, p2, vin2)]

    def two_out(vin, vout1, vout2):
        return [(vin, _edge_plane(plane_ctr, n, vin, vout1), _edge_plane(plane_ctr, n, vin, vout2))]

    if i0 and i1:
        return one_out(v2, v0, v1)
    if i0 and i2:
        return one_out(v1, v0, v2)
    if i1 and i2:
        return one_out(v0, v1, v2)
    if i0:
        return two_out(v0, v1, v2)
    if i1:
        return two_out(v1, v0, v2)
    return two_out(v2, v0, v1)


def _tri_area(tri) -> float:
    return float(np.linalg.norm(np.cross(tri[1] - tri[0], tri[2] - tri[0]))) / 2.0


def _tris_coplanar(t1, t2) -> bool:
    n1, n2 = _tri_normal(t1), _tri_normal(t2)
    if not abs(float(np.dot(n1, n2))) > 1.0 - D_EPS:
        return False
    best, pair = -1.0, None
    for a in t1:
        for b in t2:
            d = float(np.linalg.norm(a - b))
            if d > best:
                best, pair = d, (a, b)
    m = _unit(pair[0] - pair[1])
    return abs(float(np.dot(m, n1))) < D_EPS or abs(float(np.dot(m, n2))) < D_EPS


def _box_tris(box: np.ndarray):
    return [tuple(box[i] for i in t) for t in BOX_TRIANGLES]


def _box_intersections(tris, other_box: np.ndarray, other_center: np.ndarray):
    out = list(tris)
    for idx in BOX_PLANES:
        plane = [other_box[i] for i in idx]
        ctr, n = _plane_center(plane), _plane_inward_normal(plane, other_center)
        nxt = []
        for t in out:
            nxt.extend(_clip_tri(ctr, n, t))
        out = nxt
    return out


def _volume(tris, center: np.ndarray) -> float:
    return sum(abs(float(np.dot(np.cross(t[0] - center, t[1] - center), t[2] - center))) for t in tris) / 6.0


def iou_box3d_pair(box1: np.ndarray, box2: np.ndarray) -> Tuple[float, float]:
    """(intersection volume, IoU) of two [8,3] cuboids in pytorch3d's corner order; float64 arithmetic."""
    box1, box2 = np.asarray(box1, np.float64), np.asarray(box2, np.float64)
    c1, c2 = box1.mean(0), box2.mean(0)
    tris1, tris2 = _box_tris(box1), _box_tris(box2)
    in1 = _box_intersections(tris1, box2, c2)
    in2 = _box_intersections(tris2, box1, c1)
    if in2:
        keep = [True] * len(in2)
        for a in in1:
            if _tri_area(a) <= A_EPS:
                continue
            for j, b in enumerate(in2):
                if keep[j] and _tris_coplanar(a, b):
                    keep[j] = False
        in1 = in1 + [b for b, k in zip(in2, keep) if k]
    if not in1:
        return 0.0, 0.0
    poly_center = np.mean([v for t in in1 for v in t], axis=0)
    vol = _volume(in1, poly_center)
    v1, v2 = _volume(tris1, c1), _volume(tris2, c2)
    return vol, vol / (v1 + v2 - vol)


def iou_box3d(boxes1: torch.Tensor, boxes2: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """`pytorch3d._C.iou_box3d`: ([N,M] volumes, [N,M] IoUs), fp32 tensors."""
    b1, b2 = boxes1.detach().cpu().double().numpy(), boxes2.detach().cpu().double().numpy()
    vol = np.zeros((len(b1), len(b2)))
    iou = np.zeros((len(b1), len(b2)))
    for i in range(len(b1)):
        for j in range(len(b2)):
            vol[i, j], iou[i, j] = iou_box3d_pair(b1[i], b2[j])
    return torch.from_numpy(vol).float(), torch.from_numpy(iou).float()


# --------------------------------------------------------------------------- omni3d_evaluation.py:109-169
def box3d_overlap(boxes_dt: torch.Tensor, boxes_gt: torch.Tensor, eps_coplanar: float = 1e-4,
                  eps_nonzero: float = 1e-8) -> torch.Tensor:
    invalid_coplanar = ~check_coplanar(boxes_dt, eps=eps_coplanar)
    invalid_nonzero = ~check_nonzero(boxes_dt, eps=eps_nonzero)
    ious = iou_box3d(boxes_dt, boxes_gt)[1]
    if invalid_coplanar.any():
        ious[invalid_coplanar] = 0
    if invalid_nonzero.any():
        ious[invalid_nonzero] = 0
    return ious


# --------------------------------------------------------------------------- independent exact construction (qhull)
def iou_box3d_qhull(box1: np.ndarray, box2: np.ndarray) -> Tuple[float, float]:
    """Intersection volume / IoU through scipy: half-space intersection of the 12 face planes, convex hull volume."""
    from scipy.spatial import ConvexHull, HalfspaceIntersection
    from scipy.optimize import linprog

    hs = []
    for box in (np.asarray(box1, np.float64), np.asarray(box2, np.float64)):
        c = box.mean(0)
        for idx in BOX_PLANES:
            plane = [box[i] for i in idx]
            n = -_plane_inward_normal(plane, c)                 # outward
            hs.append(np.append(n, -float(np.dot(n, _plane_center(plane)))))   # n.x + d <= 0
    hs = np.asarray(hs)
    v1, v2 = ConvexHull(box1).volume, ConvexHull(box2).volume
    # Chebyshev centre: an interior point, or proof that the intersection has no interior
    res = linprog([0, 0, 0, -1], A_ub=np.hstack([hs[:, :3], np.ones((len(hs), 1))]), b_ub=-hs[:, 3],
                  bounds=[(None, None)] * 3 + [(0, None)])
    if not res.success or res.x[3] < 1e-9:
        return 0.0, 0.0
    pts = HalfspaceIntersection(hs, res.x[:3]).intersections
    vol = ConvexHull(pts).volume
    return vol, vol / (v1 + v2 - vol)


def make_box(center: Sequence[float], dims: Sequence[float], R: np.ndarray) -> np.ndarray:
    """[8,3] corners in pytorch3d's order for a cuboid with edge lengths dims = (dx, dy, dz) and rotation R."""
    unit = np.array([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0], [0, 0, 1], [1, 0, 1], [1, 1, 1], [0, 1, 1]], np.float64) - 0.5
    return (unit * np.asarray(dims, np.float64)) @ np.asarray(R, np.float64).T + np.asarray(center, np.float64)
