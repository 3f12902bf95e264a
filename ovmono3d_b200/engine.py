"""LiftEngine: owns the device-side state of the lift (bf16 FC weights, workspaces, C structs) and issues
`ovm_lift_forward` -- ROIAlign -> FC stack -> heads + decode -- on the current CUDA stream.

Host side of the inference slice of `ROIHeads3D._forward_cube`
(reference cubercnn/modeling/roi_heads/roi_heads.py:329-848); the arithmetic is in ovmono3d_b200/csrc/*.cu.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import torch
from torch import nn

from . import ops


@dataclass
class DecodeSettings:
    """The cfg keys that change the decode arithmetic (reference cubercnn/config/config.py:118-214)."""
    z_type: str = "direct"
    pose_type: str = "6d"
    cluster_bins: int = 1
    num_classes: int = 50
    dims_priors_enabled: bool = False
    dims_priors_func: str = "exp"
    use_confidence: float = 1.0
    virtual_depth: bool = True
    virtual_focal: float = 512.0
    allocentric_pose: bool = True
    scale_roi_boxes: float = 0.0


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class LiftEngine:
    # FC groups in feature-column order; index = position in ovm_decode_cfg_t.feat_offset
    GROUPS = ("XY", "dims", "pose", "Z", "conf")

    def __init__(self, cube_head: nn.Module, scales: Sequence[float], pooler_resolution: int, sampling_ratio: int,
                 settings: DecodeSettings, canonical_box_size: int = 224, canonical_level: int = 4):
        self.cube_head = cube_head
        self.scales = tuple(float(s) for s in scales)
        self.P = int(pooler_resolution)
        self.sampling_ratio = int(sampling_ratio)
        self.settings = settings
        self.canonical_box_size = canonical_box_size
        self.canonical_level = canonical_level
        if settings.pose_type not in ops.POSE_TYPES:
            raise ValueError("Cuboid pose type {} is not recognized".format(settings.pose_type))
        if settings.z_type not in ops.Z_TYPES:
            raise ValueError("Z type {} is not recognized".format(settings.z_type))
        # operand format of the stack's last FC layer when it runs fused with the heads and the decode (ovm_fc_tail):
        # "f16" (default: TF32's 10-bit mantissa in two bytes), "tf32" (fp32 storage) or "bf16"
        self.tail_precision = os.environ.get("OVM_TAIL_PRECISION", "f16").lower()
        if self.tail_precision not in ("f16", "tf32", "bf16"):
            raise ValueError("OVM_TAIL_PRECISION must be f16, tf32 or bf16")
        self._weights_key = None
        self._params = None
        self._fc_weights: List[torch.Tensor] = []
        self._fc_layers = None
        self._fc_stack = None
        self._heads = None
        self._keep: List[torch.Tensor] = []
        self._cap = 0
        self._ws_tensors: Dict[str, torch.Tensor] = {}
        self._ws = None
        self._pooler_cache: Dict[tuple, tuple] = {}

    # ------------------------------------------------------------------ weights
    def _generators(self) -> List[nn.Module]:
        h = self.cube_head
        if getattr(h, "shared_fc", True):
            return [h.feature_generator]
        gens = [h.feature_generator_XY, h.feature_generator_dims, h.feature_generator_pose, h.feature_generator_Z]
        if self.settings.use_confidence:
            gens.append(h.feature_generator_conf)
        return gens

    def _state_key(self, priors: Sequence[Optional[torch.Tensor]]):
        if self._params is None:   # the module structure is fixed after construction
            self._params = list(self.cube_head.parameters())
        ts = self._params + [p for p in priors if p is not None]
        return tuple((t.data_ptr(), t._version) for t in ts)

    @property
    def ready(self) -> bool:
        """`prepare` has built the descriptors at least once (they may be stale: `prepare` checks)."""
        return self._weights_key is not None

    def invalidate(self) -> None:
        """Forget the bf16 / TF32 copies of the FC weights: the next `prepare` rebuilds them.  `prepare` notices
        re-assigned parameters and in-place updates that bump the tensor version (`load_state_dict`, `copy_`, optimiser
        steps); writes through `param.data` or raw pointers do not bump it -- call this after such an update (the
        plugin does so from a `load_state_dict` post-hook)."""
        self._weights_key = None

    def prepare(self, device: torch.device, priors_dims_per_cat: Optional[torch.Tensor],
                priors_z_scales: Optional[torch.Tensor], priors_z_stats: Optional[torch.Tensor]) -> None:
        """(Re)build the bf16 copies of the FC weights and the C descriptors when the parameters changed."""
        key = self._state_key((priors_dims_per_cat, priors_z_scales, priors_z_stats))
        if key == self._weights_key:
            return
        h, s = self.cube_head, self.settings
        gens = self._generators()
        self._fc_weights, keep = [], []
        layers = []
        num_fc = None
        for g in gens:
            fcs = [m for m in g if isinstance(m, nn.Linear)]
            num_fc = len(fcs) if num_fc is None else num_fc
            assert len(fcs) == num_fc and num_fc >= 1
            for fc in fcs:
                w = fc.weight.detach()
                if w.device != device:
                    raise ops.LiftError("cube_head lives on %s but the features are on %s" % (w.device, device))
                N, K = w.shape
                ldw = _round_up(K, 8)
                wf = w.float().contiguous()
                wb = ops.cast_bf16(wf, ldw)                                 # [N, K] view of [N, ldw] bf16
                b = fc.bias.detach().float().contiguous()
                self._fc_weights.append(wb)
                keep.append(b)
                walt, ldalt, alt_dtype = None, 0, 0
                if fc is fcs[-1] and len(gens) == 1 and num_fc >= 2 and self.tail_precision != "bf16":
                    # the stack's last layer may run fused with the heads and the decode (ovm_fc_tail) on fp16 / TF32 operands
                    if self.tail_precision == "f16":
                        ldalt, alt_dtype = _round_up(K, 8), ops.F16
                        walt = ops.cast_f16(wf, ldalt)
                    else:
                        ldalt, alt_dtype = _round_up(K, 4), ops.TF32
                        walt = ops.round_tf32(wf, ldalt)
                    keep.append(walt)
                layers.append(ops.ovm_fc_layer_t(wb.data_ptr(), b.data_ptr(), N, K, ldw,
                                                 None if walt is None else walt.data_ptr(), ldalt, alt_dtype, 0))
        self._fc_layers = (ops.ovm_fc_layer_t * len(layers))(*layers)
        self._fc_stack = ops.ovm_fc_stack_t(len(gens), num_fc, self._fc_layers)
        self.fc_dim = layers[-1].N
        self.in_dim = layers[0].K

        def f32(t: Optional[torch.Tensor]) -> Optional[int]:
            if t is None:
                return None
            t = t.detach()
            if t.dtype != torch.float32 or not t.is_contiguous() or t.device != device:
                t = t.to(device=device, dtype=torch.float32).contiguous()
            keep.append(t)
            return t.data_ptr()

        hw = ops.ovm_head_weights_t()
        hw.w_deltas, hw.b_deltas = f32(h.bbox_3D_center_deltas.weight), f32(h.bbox_3D_center_deltas.bias)
        hw.w_dims, hw.b_dims = f32(h.bbox_3D_dims.weight), f32(h.bbox_3D_dims.bias)
        hw.w_pose, hw.b_pose = f32(h.bbox_3D_pose.weight), f32(h.bbox_3D_pose.bias)
        hw.w_z, hw.b_z = f32(h.bbox_3D_center_depth.weight), f32(h.bbox_3D_center_depth.bias)
        if s.use_confidence:
            hw.w_uncert, hw.b_uncert = f32(h.bbox_3D_uncertainty.weight), f32(h.bbox_3D_uncertainty.bias)
        hw.priors_dims_per_cat = f32(priors_dims_per_cat) if s.dims_priors_enabled else None
        hw.priors_z_scales = f32(priors_z_scales) if s.cluster_bins > 1 else None
        hw.priors_z_stats = f32(priors_z_stats) if s.z_type == "clusters" else None
        self._heads = hw

        cfg = ops.ovm_decode_cfg_t()
        cfg.z_type, cfg.pose_type = ops.Z_TYPES[s.z_type], ops.POSE_TYPES[s.pose_type]
        cfg.dims_priors_func = ops.DIMS_FUNCS[s.dims_priors_func]
        cfg.cluster_bins, cfg.num_classes = max(int(s.cluster_bins), 1), int(s.num_classes)
        cfg.dims_priors_enabled = int(bool(s.dims_priors_enabled))
        cfg.use_confidence = int(bool(s.use_confidence))
        cfg.virtual_depth, cfg.allocentric_pose = int(bool(s.virtual_depth)), int(bool(s.allocentric_pose))
        cfg.virtual_focal = float(s.virtual_focal)
        cfg.fc_dim = self.fc_dim
        shared = len(gens) == 1
        for i in range(5):
            cfg.feat_offset[i] = 0 if shared else min(i, len(gens) - 1) * self.fc_dim
        self._cfg = cfg
        # mirrors fc_tail_supported() (csrc/fc_tail.cu): does ovm_lift_forward end in the fused fc-last + heads + decode launch?
        self.tail_fused = (shared and num_fc >= 2 and not cfg.dims_priors_enabled and cfg.cluster_bins == 1
                           and self.fc_dim in (128, 256, 512, 1024) and os.environ.get("OVM_FUSE_TAIL", "1") != "0")
        self._keep = keep
        self._weights_key = key
        self._cap = 0  # workspace widths depend on the stack

    # ------------------------------------------------------------------ workspace
    def _workspace(self, n: int, device: torch.device) -> ops.ovm_workspace_t:
        if n <= self._cap and self._ws is not None:
            return self._ws
        cap = max(n, int(self._cap * 1.5), 128)
        groups = self._fc_stack.num_groups
        ld_pooled = _round_up(self.in_dim, 8)
        t = dict(
            pooled=torch.empty((cap, ld_pooled), device=device, dtype=torch.bfloat16),
            acc=torch.zeros((cap, self.fc_dim), device=device, dtype=torch.float32),   # zero on entry / exit of every lift
            act_a=torch.empty((cap, self.fc_dim), device=device, dtype=torch.float32),   # bf16 or TF32-in-fp32 activations
            act_b=torch.empty((cap, self.fc_dim), device=device, dtype=torch.float32),
            feats=torch.empty((cap, groups * self.fc_dim), device=device, dtype=torch.float32),
            levels=torch.empty((cap,), device=device, dtype=torch.int32),
            counters=torch.zeros((_round_up(cap, 128) // 128 * (_round_up(self.fc_dim, 256) // 256),), device=device,
                                 dtype=torch.int32),   # split-K tile counters: zero on entry / exit
            order=torch.zeros((ops.order_ws_len(cap),), device=device, dtype=torch.int32),   # zero once: ticket word
            # split-K exchange of the first FC layer: per-cluster partial tiles + their flags (zero between calls)
            slots=torch.empty((ops.SPLITK_SLOTS_LEN,), device=device, dtype=torch.float32),
            flags=torch.zeros((ops.SPLITK_FLAGS_LEN,), device=device, dtype=torch.int32),
        )
        ws = ops.ovm_workspace_t(t["pooled"].data_ptr(), ld_pooled, t["acc"].data_ptr(), t["act_a"].data_ptr(),
                                 t["act_b"].data_ptr(), t["feats"].data_ptr(), t["levels"].data_ptr(),
                                 t["counters"].data_ptr(), t["counters"].numel(), t["flags"].numel(),
                                 t["slots"].data_ptr(), t["slots"].numel(), t["flags"].data_ptr())
        self._ws_tensors, self._ws, self._cap = t, ws, cap
        return ws

    # ------------------------------------------------------------------ forward
    def forward(self, features: Sequence[torch.Tensor], boxes: torch.Tensor, classes: torch.Tensor,
                scores: Optional[torch.Tensor], batch_idx: torch.Tensor, cams: torch.Tensor,
                want_raw: bool = False, out: Optional[torch.Tensor] = None,
                peer_ptrs: Optional[Sequence[int]] = None) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
        """-> (records [n, 48] fp32, raw [n, 16] or None).  `prepare` must have been called.  `out` (optional,
        contiguous fp32 [n, 48] on the same device) receives the records instead of a fresh tensor.  `peer_ptrs`
        (optional): device-visible addresses of other ranks' gather slots for THIS call's records (shard.RecordExchange);
        the decode stores every record there as well (`ovm_lift_forward_peers`)."""
        lib = ops.load_library()
        n = boxes.shape[0]
        device = boxes.device
        if out is not None:
            if out.shape != (n, ops.RECORD_WIDTH) or not out.is_contiguous() or out.device != device:
                raise ops.LiftError("out must be a contiguous [n, %d] tensor on %s" % (ops.RECORD_WIDTH, device))
            records = ops._need_cuda(out, "out", torch.float32)
        else:
            records = torch.empty((n, ops.RECORD_WIDTH), device=device, dtype=torch.float32)
        raw = torch.empty((n, 16), device=device, dtype=torch.float32) if want_raw else None
        if n == 0:
            return records, raw
        pooler, keep, ws = self._describe(features, n, device)
        ops._need_cuda(boxes, "boxes", torch.float32)
        ops._need_cuda(classes, "classes", torch.int64)
        ops._need_cuda(batch_idx, "batch_idx", torch.int32)
        ops._need_cuda(cams, "cams", torch.float32)
        if scores is not None:
            ops._need_cuda(scores, "scores", torch.float32)
        if peer_ptrs:
            if len(peer_ptrs) > ops.OVM_MAX_PEERS:
                raise ops.LiftError("at most %d peers" % ops.OVM_MAX_PEERS)
            peers = ops.ovm_peer_records_t(len(peer_ptrs), 0, (C.c_void_p * ops.OVM_MAX_PEERS)(*peer_ptrs))
            rc = lib.ovm_lift_forward_peers(C.byref(pooler), C.byref(self._fc_stack), C.byref(self._cfg),
                                            C.byref(self._heads), boxes.data_ptr(), classes.data_ptr(), ops._ptr(scores),
                                            batch_idx.data_ptr(), cams.data_ptr(), n, C.byref(ws), records.data_ptr(),
                                            ops._ptr(raw), C.byref(peers), ops._stream())
        else:
            rc = lib.ovm_lift_forward(C.byref(pooler), C.byref(self._fc_stack), C.byref(self._cfg), C.byref(self._heads),
                                      boxes.data_ptr(), classes.data_ptr(), ops._ptr(scores), batch_idx.data_ptr(),
                                      cams.data_ptr(), n, C.byref(ws), records.data_ptr(), ops._ptr(raw), ops._stream())
        ops._check(rc, "ovm_lift_forward")
        del keep
        return records, raw

    def begin(self, features: Sequence[torch.Tensor], boxes: torch.Tensor, batch_idx: torch.Tensor):
        """First half of `forward` for a host-driven caller (`ovm_lift_forward_phase`, OVM_PHASE_POOL): plan and pool as
        soon as the boxes are on the device, so that the GPU works while the host is still staging classes, scores and
        cameras.  Returns the state `finish` needs; n must be > 0."""
        lib = ops.load_library()
        n, device = boxes.shape[0], boxes.device
        pooler, keep, ws = self._describe(features, n, device)
        ops._need_cuda(boxes, "boxes", torch.float32)
        ops._need_cuda(batch_idx, "batch_idx", torch.int32)
        rc = lib.ovm_lift_forward_phase(C.byref(pooler), C.byref(self._fc_stack), C.byref(self._cfg), C.byref(self._heads),
                                        boxes.data_ptr(), None, None, batch_idx.data_ptr(), None, n, C.byref(ws), None, None,
                                        None, ops.PHASE_POOL, ops._stream())
        ops._check(rc, "ovm_lift_forward_phase(pool)")
        return (pooler, keep, ws, boxes, batch_idx)

    def finish(self, state, classes: torch.Tensor, scores: Optional[torch.Tensor], cams: torch.Tensor,
               want_raw: bool = False) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
        """Second half (OVM_PHASE_HEAD): FC stack + decode on what `begin` pooled.  Same stream as `begin`."""
        lib = ops.load_library()
        pooler, keep, ws, boxes, batch_idx = state
        n, device = boxes.shape[0], boxes.device
        records = torch.empty((n, ops.RECORD_WIDTH), device=device, dtype=torch.float32)
        raw = torch.empty((n, 16), device=device, dtype=torch.float32) if want_raw else None
        ops._need_cuda(classes, "classes", torch.int64)
        ops._need_cuda(cams, "cams", torch.float32)
        if scores is not None:
            ops._need_cuda(scores, "scores", torch.float32)
        rc = lib.ovm_lift_forward_phase(C.byref(pooler), C.byref(self._fc_stack), C.byref(self._cfg), C.byref(self._heads),
                                        boxes.data_ptr(), classes.data_ptr(), ops._ptr(scores), batch_idx.data_ptr(),
                                        cams.data_ptr(), n, C.byref(ws), records.data_ptr(), ops._ptr(raw), None,
                                        ops.PHASE_HEAD, ops._stream())
        ops._check(rc, "ovm_lift_forward_phase(head)")
        del keep
        return records, raw

    def _describe(self, features: Sequence[torch.Tensor], n: int, device: torch.device):
        """-> (pooler descriptor, tensors it needs kept alive or None, workspace) for n boxes on `device`."""
        # the pyramid descriptor depends only on the maps' addresses, shapes and layouts: a steady-state caller (the same
        # backbone output buffers call after call) re-describes nothing
        pkey = tuple((f.data_ptr(), f.shape, f.stride()) for f in features)
        cached = self._pooler_cache.get(pkey)
        if cached is None:
            cached = ops.make_pooler(features, self.scales, self.P, self.sampling_ratio, self.canonical_box_size,
                                     self.canonical_level, self.settings.scale_roi_boxes)
            if cached[0].channels * self.P * self.P != self.in_dim:
                raise ops.LiftError("fc1 expects %d inputs but the pyramid gives C*P*P = %d"
                                    % (self.in_dim, cached[0].channels * self.P * self.P))
            if any(k is not f for k, f in zip(cached[1], features)):
                cached = (cached[0], cached[1], False)     # a level was copied to make it contiguous: do not cache the copy
            else:
                cached = (cached[0], None, True)           # (do not keep the caller's tensors alive)
                if len(self._pooler_cache) >= 8:
                    self._pooler_cache.clear()
                self._pooler_cache[pkey] = cached
        pooler, keep = cached[0], cached[1]
        ws = self._workspace(n, device)
        pooler.order_ws = self._ws_tensors["order"].data_ptr()
        pooler.order_ws_len = self._ws_tensors["order"].numel()
        return pooler, keep, ws

    @property
    def last_levels(self) -> Optional[torch.Tensor]:
        return self._ws_tensors.get("levels")

    @property
    def last_pooled(self) -> Optional[torch.Tensor]:
        return self._ws_tensors.get("pooled")
