"""ctypes binding of `libovmono3d_lift.so` (include/ovmono3d_lift.h) for torch CUDA tensors.

Every function here takes CUDA tensors, passes raw device pointers + the current CUDA stream through the C ABI
and raises `LiftError` on a non-zero status.  There is no CPU path: a missing library or a CPU tensor is an error.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence, Tuple

import torch

OVM_ABI_VERSION = 6
OVM_MAX_LEVELS = 8
RECORD_WIDTH = 48
LAYOUT_NCHW, LAYOUT_NHWC = 0, 1
F32, BF16, TF32, F16 = 0, 1, 2, 3
Z_TYPES = {"direct": 0, "sigmoid": 1, "log": 2, "clusters": 3}
POSE_TYPES = {"6d": 0, "quaternion": 1, "euler": 2}
DIMS_FUNCS = {"exp": 0, "sigmoid": 1}
GEMM_AUTO, GEMM_DIRECT, GEMM_STREAMK = 0, 1, 2

# record columns (include/ovmono3d_lift.h, ovm_head_decode)
REC_BBOX3D = slice(0, 24)
REC_CENTER_CAM = slice(24, 27)
REC_CENTER_2D = slice(27, 29)
REC_DIMENSIONS = slice(29, 32)
REC_POSE = slice(32, 41)
REC_BOX2D = slice(41, 45)
REC_SCORE = 45
REC_CLASS = 46

EXPORTS = ("ovm_abi_version", "ovm_last_error", "ovm_launch_count", "ovm_struct_size", "ovm_roi_align", "ovm_gemm_bf16",
           "ovm_bias_act", "ovm_cast_bf16", "ovm_head_decode", "ovm_lift_forward", "ovm_set_stage_events", "ovm_gemm_set_trace",
           "ovm_roi_set_trace", "ovm_head_set_trace", "ovm_nms", "ovm_gdino_post", "ovm_box_inference", "ovm_post_set_trace",
           "ovm_fc_tail", "ovm_round_tf32", "ovm_iou_box3d", "ovm_lift_forward_peers",
           "ovm_records_postprocess", "ovm_peer_store_f32", "ovm_cast_f16", "ovm_roi_align_static",
           "ovm_lift_forward_phase")
PHASE_POOL, PHASE_HEAD = 1, 2
ROI_STATIC_MAX_BOXES = 16384               # OVM_ROI_STATIC_MAX_BOXES of the header
NMS_MAX_SEGMENT = 4096
SPLITK_SLOTS_LEN = 74 * 2 * 128 * 256      # OVM_SPLITK_SLOTS_LEN / OVM_SPLITK_FLAGS_LEN of the header
SPLITK_FLAGS_LEN = 74 * 2 * 8
NMS_AGNOSTIC, NMS_COORD_TRICK, NMS_PER_CLASS = 0, 1, 2


class LiftError(RuntimeError):
    pass


class ovm_level_t(C.Structure):
    _fields_ = [("data", C.c_void_p), ("height", C.c_int32), ("width", C.c_int32), ("spatial_scale", C.c_double),
                ("layout", C.c_int32), ("_pad", C.c_int32)]


class ovm_pooler_t(C.Structure):
    _fields_ = [("num_levels", C.c_int32), ("levels", ovm_level_t * OVM_MAX_LEVELS), ("num_images", C.c_int32),
                ("channels", C.c_int32), ("pooler_resolution", C.c_int32), ("sampling_ratio", C.c_int32),
                ("canonical_box_size", C.c_int32), ("canonical_level", C.c_int32), ("scale_roi_boxes", C.c_float),
                ("order_ws_len", C.c_int32), ("order_ws", C.c_void_p)]


class ovm_decode_cfg_t(C.Structure):
    _fields_ = [("z_type", C.c_int32), ("pose_type", C.c_int32), ("dims_priors_func", C.c_int32),
                ("cluster_bins", C.c_int32), ("num_classes", C.c_int32), ("dims_priors_enabled", C.c_int32),
                ("use_confidence", C.c_int32), ("virtual_depth", C.c_int32), ("allocentric_pose", C.c_int32),
                ("virtual_focal", C.c_float), ("fc_dim", C.c_int32), ("feat_offset", C.c_int32 * 5)]


class ovm_head_weights_t(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in (
        "w_deltas", "b_deltas", "w_dims", "b_dims", "w_pose", "b_pose", "w_z", "b_z", "w_uncert", "b_uncert",
        "priors_dims_per_cat", "priors_z_scales", "priors_z_stats")]


class ovm_fc_layer_t(C.Structure):
    _fields_ = [("weight", C.c_void_p), ("bias", C.c_void_p), ("N", C.c_int32), ("K", C.c_int32), ("ldw", C.c_int64),
                ("weight_alt", C.c_void_p), ("ldw_alt", C.c_int64), ("alt_dtype", C.c_int32), ("_pad", C.c_int32)]


class ovm_fc_stack_t(C.Structure):
    _fields_ = [("num_groups", C.c_int32), ("num_fc", C.c_int32), ("layers", C.POINTER(ovm_fc_layer_t))]


class ovm_workspace_t(C.Structure):
    _fields_ = [("pooled", C.c_void_p), ("ld_pooled", C.c_int64), ("acc", C.c_void_p), ("act_a", C.c_void_p),
                ("act_b", C.c_void_p), ("feats", C.c_void_p), ("levels", C.c_void_p),
                ("tile_counters", C.c_void_p), ("num_tile_counters", C.c_int32), ("splitk_flags_len", C.c_int32),
                ("splitk_slots", C.c_void_p), ("splitk_slots_len", C.c_int64), ("splitk_flags", C.c_void_p)]


class ovm_box_inference_cfg_t(C.Structure):
    _fields_ = [("num_classes", C.c_int32), ("cls_agnostic_bbox_reg", C.c_int32), ("reg_weights", C.c_float * 4),
                ("scale_clamp", C.c_float), ("score_thresh", C.c_float), ("nms_thresh", C.c_double),
                ("topk_per_image", C.c_int32), ("coord_trick_max", C.c_int32)]


class ovm_box_inference_ws_t(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in ("probs", "cand_boxes", "class_keep", "class_count", "image_stats")]


OVM_MAX_PEERS = 8


class ovm_peer_records_t(C.Structure):
    _fields_ = [("num_peers", C.c_int32), ("_pad", C.c_int32), ("records", C.c_void_p * OVM_MAX_PEERS)]


ABI_STRUCTS = (ovm_level_t, ovm_pooler_t, ovm_decode_cfg_t, ovm_head_weights_t, ovm_fc_layer_t, ovm_fc_stack_t,
               ovm_workspace_t, ovm_box_inference_cfg_t, ovm_box_inference_ws_t, ovm_peer_records_t)
_LIB: Optional[C.CDLL] = None
LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libovmono3d_lift.so")


def load_library() -> C.CDLL:
    """Load the in-tree C-ABI library.  Fails loudly: the lift has no other implementation."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        raise LiftError(
            "%s is missing: build it with `python -m ovmono3d_b200.build` (needs nvcc, sm_100a). "
            "The lift path has no CPU or PyTorch fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name in EXPORTS:
        if not hasattr(lib, name):
            raise LiftError("%s does not export %s" % (LIB_PATH, name))
    lib.ovm_abi_version.restype = C.c_int
    lib.ovm_last_error.restype = C.c_char_p
    lib.ovm_launch_count.restype = C.c_int64
    if lib.ovm_abi_version() != OVM_ABI_VERSION:
        raise LiftError("ABI version mismatch: library %d, binding %d" % (lib.ovm_abi_version(), OVM_ABI_VERSION))
    vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
    lib.ovm_roi_align.argtypes = [C.POINTER(ovm_pooler_t), vp, vp, i32, vp, i32, i64, vp, vp]
    lib.ovm_roi_align_static.argtypes = [C.POINTER(ovm_pooler_t), vp, vp, i32, vp, i32, i64, vp, vp]
    lib.ovm_gemm_bf16.argtypes = [vp, i64, vp, i64, i32, i32, i32, vp, i32, vp, i32, i64, i32, C.POINTER(i32), vp]
    lib.ovm_bias_act.argtypes = [vp, i64, vp, i32, i32, i32, vp, i32, i64, vp]
    lib.ovm_cast_bf16.argtypes = [vp, i64, i32, i32, vp, i64, vp]
    lib.ovm_head_decode.argtypes = [C.POINTER(ovm_decode_cfg_t), C.POINTER(ovm_head_weights_t), vp, i64, vp, vp, vp,
                                    vp, vp, i32, i32, vp, vp, vp]
    lib.ovm_lift_forward.argtypes = [C.POINTER(ovm_pooler_t), C.POINTER(ovm_fc_stack_t), C.POINTER(ovm_decode_cfg_t),
                                     C.POINTER(ovm_head_weights_t), vp, vp, vp, vp, vp, i32,
                                     C.POINTER(ovm_workspace_t), vp, vp, vp]
    lib.ovm_lift_forward_peers.argtypes = [C.POINTER(ovm_pooler_t), C.POINTER(ovm_fc_stack_t), C.POINTER(ovm_decode_cfg_t),
                                           C.POINTER(ovm_head_weights_t), vp, vp, vp, vp, vp, i32,
                                           C.POINTER(ovm_workspace_t), vp, vp, C.POINTER(ovm_peer_records_t), vp]
    lib.ovm_lift_forward_phase.argtypes = [C.POINTER(ovm_pooler_t), C.POINTER(ovm_fc_stack_t), C.POINTER(ovm_decode_cfg_t),
                                           C.POINTER(ovm_head_weights_t), vp, vp, vp, vp, vp, i32,
                                           C.POINTER(ovm_workspace_t), vp, vp, C.POINTER(ovm_peer_records_t), i32, vp]
    lib.ovm_fc_tail.argtypes = [vp, i64, i32, vp, i64, vp, i32, i32, C.POINTER(ovm_decode_cfg_t),
                                C.POINTER(ovm_head_weights_t), vp, vp, vp, vp, vp, i32, i32, vp, vp, vp]
    lib.ovm_round_tf32.argtypes = [vp, i64, i32, i32, vp, i64, vp]
    lib.ovm_cast_f16.argtypes = [vp, i64, i32, i32, vp, i64, vp]
    lib.ovm_iou_box3d.argtypes = [vp, vp, i32, i32, C.c_float, C.c_float, vp, vp, vp, vp]
    lib.ovm_records_postprocess.argtypes = [vp, vp, vp, i32, vp, vp, vp]
    lib.ovm_peer_store_f32.argtypes = [C.POINTER(C.c_void_p), i32, C.c_float, vp]
    lib.ovm_gemm_set_trace.argtypes = [vp]
    lib.ovm_roi_set_trace.argtypes = [vp]
    lib.ovm_head_set_trace.argtypes = [vp]
    lib.ovm_post_set_trace.argtypes = [vp]
    lib.ovm_set_stage_events.argtypes = [C.POINTER(C.c_void_p), i32]
    lib.ovm_nms.argtypes = [vp, vp, vp, vp, i32, i32, C.c_double, i32, vp, vp, vp]
    lib.ovm_box_inference.argtypes = [C.POINTER(ovm_box_inference_cfg_t), vp, i64, vp, i64, vp, vp, vp, i32, i32, i32,
                                      C.POINTER(ovm_box_inference_ws_t), vp, vp, vp, vp, vp, vp, vp]
    lib.ovm_gdino_post.argtypes = [vp, vp, vp, i32, i32, i32, i32, i32, i32, C.c_float, C.c_double, vp, vp, vp, vp,
                                   vp, vp]
    for name in EXPORTS[3:]:
        getattr(lib, name).restype = C.c_int
    lib.ovm_struct_size.argtypes = [C.c_int]
    for i, st in enumerate(ABI_STRUCTS):
        if lib.ovm_struct_size(i) != C.sizeof(st):
            raise LiftError("ABI struct %s: library says %d bytes, binding %d" % (
                st.__name__, lib.ovm_struct_size(i), C.sizeof(st)))
    _LIB = lib
    return lib


def launch_count() -> int:
    return int(load_library().ovm_launch_count())


def set_stage_events(events: Optional[Sequence["torch.cuda.Event"]]) -> None:
    """bench.py's per-stage timing hook: `ovm_lift_forward` records these events between its stages (see the
    header).  The events must have been recorded once already so that their CUDA handles exist."""
    lib = load_library()
    if not events:
        _check(lib.ovm_set_stage_events(None, 0), "ovm_set_stage_events")
        return
    arr = (C.c_void_p * len(events))(*[e.cuda_event for e in events])
    _check(lib.ovm_set_stage_events(arr, len(events)), "ovm_set_stage_events")


def _check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load_library().ovm_last_error().decode("utf-8", "replace")
        raise LiftError("%s failed (%d): %s" % (what, rc, msg))


def _stream() -> C.c_void_p:
    # the raw handle of the current stream of the current device, through torch's C binding: torch.cuda.current_stream()
    # builds a Stream object through three Python layers (18 us per call in the plugin's host profile)
    return C.c_void_p(torch._C._cuda_getCurrentRawStream(torch._C._cuda_getDevice()))


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _need_cuda(t: torch.Tensor, name: str, dtype: torch.dtype) -> torch.Tensor:
    if not t.is_cuda:
        raise LiftError("%s must be a CUDA tensor (the lift has no CPU path)" % name)
    if t.dtype != dtype:
        raise LiftError("%s must be %s, got %s" % (name, dtype, t.dtype))
    return t


def dtype_code(dt: torch.dtype) -> int:
    if dt == torch.float32:
        return F32
    if dt == torch.bfloat16:
        return BF16
    if dt == torch.float16:
        return F16
    raise LiftError("unsupported output dtype %s" % dt)


def cast_f16(x: torch.Tensor, ld_out: Optional[int] = None) -> torch.Tensor:
    """fp32 [rows, cols] -> fp16 [rows, ld_out] (saturated, zero-padded), returned as the [rows, cols] view."""
    lib = load_library()
    _need_cuda(x, "x", torch.float32)
    rows, cols = x.shape
    ld_out = ld_out or (cols + 7) // 8 * 8
    out = torch.empty((rows, ld_out), device=x.device, dtype=torch.float16)
    if rows:
        _check(lib.ovm_cast_f16(x.data_ptr(), x.stride(0), rows, cols, out.data_ptr(), ld_out, _stream()), "ovm_cast_f16")
    return out[:, :cols]


def round_tf32(x: torch.Tensor, ld_out: Optional[int] = None) -> torch.Tensor:
    """fp32 [rows, cols] -> fp32 [rows, ld_out] rounded to TF32 (zero-padded), returned as the [rows, cols] view: the
    weight operand of a `kind::tf32` layer (`ovm_fc_tail`)."""
    lib = load_library()
    _need_cuda(x, "x", torch.float32)
    rows, cols = x.shape
    ld_out = ld_out or (cols + 3) // 4 * 4
    out = torch.empty((rows, ld_out), device=x.device, dtype=torch.float32)
    if rows:
        _check(lib.ovm_round_tf32(x.data_ptr(), x.stride(0), rows, cols, out.data_ptr(), ld_out, _stream()),
               "ovm_round_tf32")
    return out[:, :cols]


def order_ws_len(n: int) -> int:
    """OVM_ORDER_WS_LEN(n) of include/ovmono3d_lift.h."""
    return 48 * max(int(n), 1) + 2048


def make_pooler(features: Sequence[torch.Tensor], scales: Sequence[float], output_size: int, sampling_ratio: int,
                canonical_box_size: int = 224, canonical_level: int = 4, scale_roi_boxes: float = 0.0
                ) -> Tuple[ovm_pooler_t, List[torch.Tensor]]:
    """Describe a pyramid for `ovm_roi_align`.  Returns the struct and the tensors it points into (keep alive).

    NCHW-contiguous maps are used as they are; `torch.channels_last` maps are passed as NHWC without a copy."""
    if not 1 <= len(features) <= OVM_MAX_LEVELS or len(features) != len(scales):
        raise LiftError("need 1..%d feature maps with one scale each" % OVM_MAX_LEVELS)
    p = ovm_pooler_t()
    p.num_levels = len(features)
    keep = []
    N, Cc = features[0].shape[0], features[0].shape[1]
    for l, (f, s) in enumerate(zip(features, scales)):
        _need_cuda(f, "features[%d]" % l, torch.float32)
        if f.dim() != 4 or f.shape[0] != N or f.shape[1] != Cc:
            raise LiftError("features[%d] has shape %s, expected [%d, %d, H, W]" % (l, tuple(f.shape), N, Cc))
        if f.is_contiguous():
            layout = LAYOUT_NCHW
        elif f.is_contiguous(memory_format=torch.channels_last):
            layout = LAYOUT_NHWC
        else:
            f = f.contiguous()
            layout = LAYOUT_NCHW
        keep.append(f)
        p.levels[l] = ovm_level_t(f.data_ptr(), f.shape[2], f.shape[3], float(s), layout, 0)
    p.num_images, p.channels = N, Cc
    p.pooler_resolution, p.sampling_ratio = int(output_size), int(sampling_ratio)
    p.canonical_box_size, p.canonical_level = int(canonical_box_size), int(canonical_level)
    p.scale_roi_boxes = float(scale_roi_boxes)
    return p, keep


def roi_align(features: Sequence[torch.Tensor], scales: Sequence[float], boxes: torch.Tensor,
              batch_idx: torch.Tensor, output_size: int = 7, sampling_ratio: int = 0,
              out_dtype: torch.dtype = torch.float32, scale_roi_boxes: float = 0.0,
              canonical_box_size: int = 224, canonical_level: int = 4) -> Tuple[torch.Tensor, torch.Tensor]:
    """Multi-level ROIAlignV2 -> (pooled [n, C, P, P], 0-based level per box [n] int32)."""
    lib = load_library()
    pooler, keep = make_pooler(features, scales, output_size, sampling_ratio, canonical_box_size, canonical_level,
                               scale_roi_boxes)
    boxes = _need_cuda(boxes, "boxes", torch.float32).contiguous()
    batch_idx = _need_cuda(batch_idx, "batch_idx", torch.int32).contiguous()
    n = boxes.shape[0]
    Cc, P = pooler.channels, output_size
    out = torch.empty((n, Cc, P, P), device=boxes.device, dtype=out_dtype)
    levels = torch.empty((n,), device=boxes.device, dtype=torch.int32)
    order = torch.zeros((order_ws_len(n),), device=boxes.device, dtype=torch.int32)     # zero-filled once (ticket word)
    pooler.order_ws, pooler.order_ws_len = order.data_ptr(), order.numel()
    if n:
        _check(lib.ovm_roi_align(C.byref(pooler), boxes.data_ptr(), batch_idx.data_ptr(), n, out.data_ptr(),
                                 dtype_code(out_dtype), Cc * P * P, levels.data_ptr(), _stream()), "ovm_roi_align")
    del keep
    return out, levels


def roi_align_static(pooler: ovm_pooler_t, boxes: torch.Tensor, batch_idx: torch.Tensor, out: torch.Tensor,
                     levels: Optional[torch.Tensor] = None) -> None:
    """`ovm_roi_align_static`: the same ROIAlign with the planner's scratch inside the library (a file-scope device array),
    for callers that may not allocate.  Everything is supplied by the caller -- `pooler` from `make_pooler` (keep its
    tensors alive), `out` [n, C, P, P] fp32 / bf16 contiguous, `levels` [n] int32 -- and nothing is allocated here, by
    torch or by the library.  n <= ROI_STATIC_MAX_BOXES; calls that overlap on the GPU race (one shared array)."""
    lib = load_library()
    n = boxes.shape[0]
    _need_cuda(boxes, "boxes", torch.float32)
    _need_cuda(batch_idx, "batch_idx", torch.int32)
    if not (boxes.is_contiguous() and batch_idx.is_contiguous() and out.is_contiguous()):
        raise LiftError("roi_align_static takes contiguous tensors (it copies nothing)")
    cpp = pooler.channels * pooler.pooler_resolution ** 2
    if out.device != boxes.device or out.numel() != n * cpp:
        raise LiftError("out must hold [n, C, P, P] on %s" % boxes.device)
    if levels is not None:
        _need_cuda(levels, "levels", torch.int32)
    _check(lib.ovm_roi_align_static(C.byref(pooler), boxes.data_ptr(), batch_idx.data_ptr(), n, out.data_ptr(),
                                    dtype_code(out.dtype), cpp, _ptr(levels), _stream()), "ovm_roi_align_static")


def gemm_bf16(A: torch.Tensor, B: torch.Tensor, bias: Optional[torch.Tensor] = None, relu: bool = False,
              out_dtype: torch.dtype = torch.float32, mode: int = GEMM_AUTO,
              out: Optional[torch.Tensor] = None, tf32_out: bool = False) -> Tuple[torch.Tensor, int]:
    """act(A[M,K] @ B[N,K]^T + bias) with bf16 operands and fp32 accumulation.  Returns (out, mode used).

    In STREAMK mode (chosen by AUTO for long-K, few-tile shapes when no bias/relu is requested) `out` is the raw
    fp32 accumulation and must start at zero (it is zeroed here when allocated here).  `tf32_out`: fp32 results are
    rounded to TF32 (the input format of a `kind::tf32` layer; DIRECT mode only)."""
    lib = load_library()
    _need_cuda(A, "A", torch.bfloat16)
    _need_cuda(B, "B", torch.bfloat16)
    if A.stride(-1) != 1 or B.stride(-1) != 1 or A.dim() != 2 or B.dim() != 2 or A.shape[1] != B.shape[1]:
        raise LiftError("gemm_bf16: A [M,K] and B [N,K] must be row-major with the same K")
    M, K = A.shape
    N = B.shape[0]
    ldo = (N + 7) // 8 * 8
    if out is None:
        store = torch.zeros((M, ldo), device=A.device, dtype=out_dtype)
        out = store[:, :N]
    else:
        ldo = out.stride(0)
    used = C.c_int32(0)
    if bias is not None:
        _need_cuda(bias, "bias", torch.float32)
    _check(lib.ovm_gemm_bf16(A.data_ptr(), A.stride(0), B.data_ptr(), B.stride(0), M, N, K, _ptr(bias), int(relu),
                             out.data_ptr(), TF32 if (tf32_out and out.dtype == torch.float32) else dtype_code(out.dtype),
                             ldo, mode, C.byref(used), _stream()),
           "ovm_gemm_bf16")
    return out, used.value


def bias_act(x: torch.Tensor, bias: Optional[torch.Tensor], relu: bool, out_dtype: torch.dtype) -> torch.Tensor:
    lib = load_library()
    _need_cuda(x, "x", torch.float32)
    M, N = x.shape
    out = torch.empty((M, N), device=x.device, dtype=out_dtype)
    _check(lib.ovm_bias_act(x.data_ptr(), x.stride(0), _ptr(bias), int(relu), M, N, out.data_ptr(),
                            dtype_code(out_dtype), N, _stream()), "ovm_bias_act")
    return out


def cast_bf16(x: torch.Tensor, ld_out: Optional[int] = None) -> torch.Tensor:
    """fp32 [rows, cols] -> bf16 [rows, ld_out] (zero-padded), returned as the [rows, cols] view."""
    lib = load_library()
    _need_cuda(x, "x", torch.float32)
    rows, cols = x.shape
    ld_out = ld_out or (cols + 7) // 8 * 8
    out = torch.empty((rows, ld_out), device=x.device, dtype=torch.bfloat16)
    if rows:
        _check(lib.ovm_cast_bf16(x.data_ptr(), x.stride(0), rows, cols, out.data_ptr(), ld_out, _stream()),
               "ovm_cast_bf16")
    return out[:, :cols]


def head_decode(cfg: "ovm_decode_cfg_t", heads: "ovm_head_weights_t", feats: torch.Tensor, boxes: torch.Tensor,
                classes: torch.Tensor, scores: Optional[torch.Tensor], batch_idx: torch.Tensor, cams: torch.Tensor,
                want_raw: bool = False) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """`ovm_head_decode`: fp32 FC features [n, ld] -> (records [n, 48], raw [n, 16] or None)."""
    lib = load_library()
    _need_cuda(feats, "feats", torch.float32)
    n = boxes.shape[0]
    rec = torch.empty((n, RECORD_WIDTH), device=feats.device, dtype=torch.float32)
    raw = torch.empty((n, 16), device=feats.device, dtype=torch.float32) if want_raw else None
    _check(lib.ovm_head_decode(C.byref(cfg), C.byref(heads), feats.data_ptr(), feats.stride(0), boxes.data_ptr(),
                               classes.data_ptr(), _ptr(scores), batch_idx.data_ptr(), cams.data_ptr(), cams.shape[0], n,
                               rec.data_ptr(), _ptr(raw), _stream()), "ovm_head_decode")
    return rec, raw


def fc_tail(cfg: "ovm_decode_cfg_t", heads: "ovm_head_weights_t", A: torch.Tensor, W: torch.Tensor, bias: torch.Tensor,
            boxes: torch.Tensor, classes: torch.Tensor, scores: Optional[torch.Tensor], batch_idx: torch.Tensor,
            cams: torch.Tensor, want_raw: bool = False) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """`ovm_fc_tail`: last FC layer + heads + decode in one launch.  A [n, K] and W [N, K] are both bf16, both fp16
    (`cast_f16`), or both fp32 holding TF32-rounded values (`round_tf32`, or `gemm_bf16(..., tf32_out=True)`)."""
    lib = load_library()
    if A.dtype != W.dtype or A.dtype not in (torch.float32, torch.bfloat16, torch.float16):
        raise LiftError("fc_tail: A and W must both be bf16, both fp16, or both fp32 (TF32-rounded)")
    _need_cuda(A, "A", A.dtype)
    _need_cuda(W, "W", W.dtype)
    _need_cuda(bias, "bias", torch.float32)
    if A.dim() != 2 or W.dim() != 2 or A.shape[1] != W.shape[1] or A.stride(1) != 1 or W.stride(1) != 1:
        raise LiftError("fc_tail: A [n,K] and W [N,K] must be row-major with the same K")
    n, K = A.shape
    N = W.shape[0]
    rec = torch.empty((n, RECORD_WIDTH), device=A.device, dtype=torch.float32)
    raw = torch.empty((n, 16), device=A.device, dtype=torch.float32) if want_raw else None
    _check(lib.ovm_fc_tail(A.data_ptr(), A.stride(0), {torch.float32: TF32, torch.float16: F16, torch.bfloat16: BF16}[A.dtype],
                           W.data_ptr(),
                           W.stride(0), bias.data_ptr(), N, K, C.byref(cfg), C.byref(heads), boxes.data_ptr(),
                           classes.data_ptr(), _ptr(scores), batch_idx.data_ptr(), cams.data_ptr(), cams.shape[0], n,
                           rec.data_ptr(), _ptr(raw), _stream()), "ovm_fc_tail")
    return rec, raw


def peer_store_f32(ptrs: Sequence[int], value: float) -> None:
    """`ovm_peer_store_f32`: one fp32 value to each of the (peer-mapped) addresses, one launch on the current stream."""
    lib = load_library()
    arr = (C.c_void_p * len(ptrs))(*ptrs)
    _check(lib.ovm_peer_store_f32(arr, len(ptrs), float(value), _stream()), "ovm_peer_store_f32")


def records_postprocess(records: torch.Tensor, offsets: torch.Tensor, geom: torch.Tensor):
    """`ovm_records_postprocess`: records [n,48], offsets [N+1] int32, geom [N,4] (scale_x, scale_y, out_w, out_h), all
    CUDA -> (out [n,48] with image i's surviving rows compacted from offsets[i], XYWH boxes; count [N] int32)."""
    lib = load_library()
    _need_cuda(records, "records", torch.float32)
    _need_cuda(offsets, "offsets", torch.int32)
    _need_cuda(geom, "geom", torch.float32)
    if records.dim() != 2 or records.shape[1] != RECORD_WIDTH or not records.is_contiguous():
        raise LiftError("records must be a contiguous [n, %d] tensor" % RECORD_WIDTH)
    N = offsets.numel() - 1
    if geom.shape != (N, 4) or not geom.is_contiguous():
        raise LiftError("geom must be a contiguous [N, 4] tensor")
    out = torch.empty_like(records)
    count = torch.zeros((max(N, 1),), dtype=torch.int32, device=records.device)
    _check(lib.ovm_records_postprocess(_ptr(records), _ptr(offsets), _ptr(geom), N, _ptr(out), _ptr(count), _stream()),
           "ovm_records_postprocess")
    return out, count[:N]


def iou_box3d(boxes_dt: torch.Tensor, boxes_gt: torch.Tensor, eps_coplanar: float = 1e-4, eps_nonzero: float = 1e-8,
              check_dt: bool = True, want_vol: bool = False):
    """`ovm_iou_box3d`: [N,8,3] x [M,8,3] corners (pytorch3d order) -> (iou [N,M], flags [N] int32 or None, vol or None).
    With `check_dt` the reference's coplanarity / zero-area guards run on the detections and zero their rows."""
    lib = load_library()
    _need_cuda(boxes_dt, "boxes_dt", torch.float32)
    _need_cuda(boxes_gt, "boxes_gt", torch.float32)
    if boxes_dt.dim() != 3 or boxes_dt.shape[1:] != (8, 3) or boxes_gt.dim() != 3 or boxes_gt.shape[1:] != (8, 3):
        raise LiftError("iou_box3d expects [N,8,3] and [M,8,3] corner tensors")
    boxes_dt, boxes_gt = boxes_dt.contiguous(), boxes_gt.contiguous()
    N, M = boxes_dt.shape[0], boxes_gt.shape[0]
    dev = boxes_dt.device
    iou = torch.empty((N, M), dtype=torch.float32, device=dev)
    vol = torch.empty((N, M), dtype=torch.float32, device=dev) if want_vol else None
    flags = torch.zeros((N,), dtype=torch.int32, device=dev) if check_dt else None
    _check(lib.ovm_iou_box3d(_ptr(boxes_dt), _ptr(boxes_gt), N, M, float(eps_coplanar), float(eps_nonzero), _ptr(iou),
                             _ptr(vol), _ptr(flags), _stream()), "ovm_iou_box3d")
    return iou, flags, vol


def nms(boxes: torch.Tensor, scores: torch.Tensor, iou_threshold: float, classes: Optional[torch.Tensor] = None,
        seg_offsets: Optional[torch.Tensor] = None, max_segment_len: Optional[int] = None,
        mode: Optional[int] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """Segmented greedy NMS (`ovm_nms`): torchvision.ops.nms / batched_nms per segment, one CTA each, no host
    round trip.  Returns (keep [n] int32, keep_count [S] int32): `keep[off[s] : off[s] + keep_count[s]]` are the
    rows of segment s (relative to its start) that survive, in decreasing score order."""
    lib = load_library()
    _need_cuda(boxes, "boxes", torch.float32)
    _need_cuda(scores, "scores", torch.float32)
    n = boxes.shape[0]
    if boxes.dim() != 2 or boxes.shape[1] != 4 or scores.shape != (n,):
        raise LiftError("nms expects boxes [n,4] and scores [n]")
    boxes, scores = boxes.contiguous(), scores.contiguous()
    if mode is None:
        mode = NMS_AGNOSTIC if classes is None else NMS_COORD_TRICK
    if classes is not None:
        classes = _need_cuda(classes, "classes", torch.int64).contiguous()
    if seg_offsets is None:
        seg_offsets = torch.tensor([0, n], dtype=torch.int32, device=boxes.device)
        max_segment_len = n
    else:
        _need_cuda(seg_offsets, "seg_offsets", torch.int32)
        if max_segment_len is None:
            max_segment_len = n
    S = seg_offsets.numel() - 1
    if n == 0:                      # nothing to launch (torchvision returns an empty keep list as well)
        return (torch.zeros((0,), dtype=torch.int32, device=boxes.device),
                torch.zeros((S,), dtype=torch.int32, device=boxes.device))
    keep = torch.empty((max(n, 1),), dtype=torch.int32, device=boxes.device)
    count = torch.empty((max(S, 1),), dtype=torch.int32, device=boxes.device)
    _check(lib.ovm_nms(_ptr(boxes), _ptr(scores), _ptr(classes), _ptr(seg_offsets), S, int(max_segment_len),
                       float(iou_threshold), int(mode), _ptr(keep), _ptr(count), _stream()), "ovm_nms")
    return keep[:n], count[:S]


def gdino_post(logits: torch.Tensor, boxes_cxcywh: torch.Tensor, phrase_ranges: torch.Tensor, image_hw: Tuple[int, int],
               box_threshold: float, nms_threshold: float):
    """`ovm_gdino_post`: raw GroundingDINO outputs [B,nq,T] / [B,nq,4] -> (boxes [B,nq,4] xyxy px, scores [B,nq],
    classes [B,nq] int64, query [B,nq] int32, count [B] int32); rows past count[b] are undefined."""
    lib = load_library()
    _need_cuda(logits, "logits", torch.float32)
    _need_cuda(boxes_cxcywh, "boxes_cxcywh", torch.float32)
    _need_cuda(phrase_ranges, "phrase_ranges", torch.int32)
    if logits.dim() != 3 or boxes_cxcywh.shape != (logits.shape[0], logits.shape[1], 4):
        raise LiftError("gdino_post expects logits [B,nq,T] and boxes [B,nq,4]")
    if phrase_ranges.dim() != 2 or phrase_ranges.shape[1] != 2 or phrase_ranges.shape[0] == 0:
        raise LiftError("phrase_ranges must be [P,2] int32 with P >= 1")
    logits, boxes_cxcywh, phrase_ranges = logits.contiguous(), boxes_cxcywh.contiguous(), phrase_ranges.contiguous()
    B, nq, T = logits.shape
    dev = logits.device
    out_boxes = torch.empty((B, max(nq, 1), 4), dtype=torch.float32, device=dev)
    out_scores = torch.empty((B, max(nq, 1)), dtype=torch.float32, device=dev)
    out_classes = torch.empty((B, max(nq, 1)), dtype=torch.int64, device=dev)
    out_query = torch.empty((B, max(nq, 1)), dtype=torch.int32, device=dev)
    count = torch.empty((max(B, 1),), dtype=torch.int32, device=dev)
    _check(lib.ovm_gdino_post(_ptr(logits), _ptr(boxes_cxcywh), _ptr(phrase_ranges), B, nq, T, phrase_ranges.shape[0],
                              int(image_hw[0]), int(image_hw[1]), float(box_threshold), float(nms_threshold),
                              _ptr(out_boxes), _ptr(out_scores), _ptr(out_classes), _ptr(out_query), _ptr(count),
                              _stream()), "ovm_gdino_post")
    return out_boxes[:, :nq], out_scores[:, :nq], out_classes[:, :nq], out_query[:, :nq], count[:B]


def box_inference(logits: torch.Tensor, deltas: torch.Tensor, proposals: torch.Tensor, seg_offsets: torch.Tensor,
                  image_sizes: torch.Tensor, max_per_image: int, num_classes: int, reg_weights: Sequence[float],
                  score_thresh: float, nms_thresh: float, topk_per_image: int, cls_agnostic_bbox_reg: bool = False,
                  coord_trick_max: int = 1000, want_scores_full: bool = True,
                  scale_clamp: float = 4.135166556742356):
    """`ovm_box_inference` (fast_rcnn.py:57-143).  logits [R, >=K+1], deltas [R, >=4K or 4] fp32 (row strides are
    passed through), proposals [R,4], seg_offsets [N+1] int32, image_sizes [N,2] int32 (h, w), all CUDA.
    Returns (boxes [N,cap,4], scores [N,cap], classes [N,cap] int64, index [N,cap] int32, scores_full [N,cap,K] or
    None, count [N] int32)."""
    lib = load_library()
    for t, nm in ((logits, "logits"), (deltas, "deltas"), (proposals, "proposals")):
        _need_cuda(t, nm, torch.float32)
    _need_cuda(seg_offsets, "seg_offsets", torch.int32)
    _need_cuda(image_sizes, "image_sizes", torch.int32)
    if logits.stride(-1) != 1 or deltas.stride(-1) != 1:
        raise LiftError("box_inference: logits / deltas rows must be contiguous")
    proposals, image_sizes = proposals.contiguous(), image_sizes.contiguous()
    R, K, N = proposals.shape[0], int(num_classes), seg_offsets.numel() - 1
    dev = proposals.device
    Kb = 1 if cls_agnostic_bbox_reg else K
    per_class = min(topk_per_image, max_per_image) if topk_per_image >= 0 else max_per_image
    cap = max(topk_per_image if topk_per_image >= 0 else K * per_class, 1)
    cfg = ovm_box_inference_cfg_t(K, int(bool(cls_agnostic_bbox_reg)), (C.c_float * 4)(*[float(w) for w in reg_weights]),
                                  float(scale_clamp), float(score_thresh), float(nms_thresh), int(topk_per_image),
                                  int(coord_trick_max))
    scratch = dict(probs=torch.empty((max(R, 1), K), dtype=torch.float32, device=dev),
                   cand_boxes=torch.empty((max(R, 1), Kb, 4), dtype=torch.float32, device=dev),
                   class_keep=torch.empty((max(N, 1), K, max(max_per_image, 1)), dtype=torch.int32, device=dev),
                   class_count=torch.empty((max(N, 1), K), dtype=torch.int32, device=dev),
                   image_stats=torch.empty((max(N, 1), 2), dtype=torch.int32, device=dev))
    ws = ovm_box_inference_ws_t(*[scratch[k].data_ptr() for k in ("probs", "cand_boxes", "class_keep", "class_count",
                                                                  "image_stats")])
    out_boxes = torch.empty((max(N, 1), cap, 4), dtype=torch.float32, device=dev)
    out_scores = torch.empty((max(N, 1), cap), dtype=torch.float32, device=dev)
    out_classes = torch.empty((max(N, 1), cap), dtype=torch.int64, device=dev)
    out_index = torch.empty((max(N, 1), cap), dtype=torch.int32, device=dev)
    out_full = torch.empty((max(N, 1), cap, K), dtype=torch.float32, device=dev) if want_scores_full else None
    count = torch.zeros((max(N, 1),), dtype=torch.int32, device=dev)
    _check(lib.ovm_box_inference(C.byref(cfg), _ptr(logits) if R else None, logits.stride(0) if R else K + 1,
                                 _ptr(deltas) if R else None, deltas.stride(0) if R else 4 * Kb,
                                 _ptr(proposals) if R else None, _ptr(seg_offsets), _ptr(image_sizes), N, R,
                                 int(max_per_image), C.byref(ws), _ptr(out_boxes), _ptr(out_scores), _ptr(out_classes),
                                 _ptr(out_index), _ptr(out_full), _ptr(count), _stream()), "ovm_box_inference")
    return out_boxes[:N], out_scores[:N], out_classes[:N], out_index[:N], (out_full[:N] if want_scores_full else None), count[:N]
