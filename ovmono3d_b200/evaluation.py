"""Evaluation-side pieces that sit right after the lift.

`box3d_overlap` mirrors the reference's function of the same name (cubercnn/evaluation/omni3d_evaluation.py:109-169):
same arguments, same result (an [N, M] IoU matrix with the rows of non-coplanar / zero-area detections zeroed), same
warnings -- computed by `ovm_iou_box3d` on the GPU instead of pytorch3d's CPU loop.
"""
from __future__ import annotations

import torch

from . import ops


def box3d_overlap(boxes_dt: torch.Tensor, boxes_gt: torch.Tensor, eps_coplanar: float = 1e-4,
                  eps_nonzero: float = 1e-8) -> torch.Tensor:
    """IoU of 3-D boxes given by their 8 corners, [N,8,3] x [M,8,3] -> [N,M] (corner order: omni3d_evaluation.py:116-143).

    The tensors must live on a CUDA device (the evaluator of the reference moves them to the CPU first, :65); there is
    no CPU path here."""
    iou, flags, _ = ops.iou_box3d(boxes_dt.float(), boxes_gt.float(), eps_coplanar, eps_nonzero, check_dt=True)
    if boxes_dt.shape[0]:
        f = flags.cpu()          # the reference prints its warnings from host-side .any() tests as well (:161-167)
        n_coplanar, n_zero = int((f & 1).ne(0).sum()), int((f & 2).ne(0).sum())
        if n_coplanar:
            print('Warning: skipping {:d} non-coplanar boxes at eval.'.format(n_coplanar))
        if n_zero:
            print('Warning: skipping {:d} zero volume boxes at eval.'.format(n_zero))
    return iou
