/* ovmono3d_lift.h -- C ABI of the B200-native 2D->3D lift (libovmono3d_lift.so).
 *
 * Every entry point takes raw device pointers, sizes and a cudaStream_t (passed as void*),
 * returns 0 on success or a negative OVM_E* code (text through ovm_last_error()), never throws,
 * never allocates or frees device memory, and never synchronises the stream.  The caller
 * (ovmono3d_b200/ops.py through ctypes) owns every buffer.
 *
 * What each call replaces in the reference (paths relative to the reference checkout):
 *   ovm_roi_align    detectron2 ROIPooler.forward + assign_boxes_to_levels + torchvision roi_align,
 *                    constructed at cubercnn/modeling/roi_heads/roi_heads.py:166-171, called at :366;
 *                    optional zoom = ROIHeads3D.scale_proposals, roi_heads.py:310-327.
 *   ovm_gemm_bf16    one nn.Linear(+ReLU) of CubeHead's FC stack, cube_head.py:64-74,157
 *                    (cuBLAS SGEMM in the reference; tcgen05 bf16 x bf16 -> fp32 here).
 *   ovm_bias_act     the bias + ReLU that follows a split-K ovm_gemm_bf16 (same lines).
 *   ovm_head_decode  the five linear heads, cube_head.py:158-164,176-204, the class/bin selection
 *                    roi_heads.py:433-477 and the whole decode roi_heads.py:378-419,480-549,795-843
 *                    with math_util.py:116-219,581-592,651-679,969-978.
 *   ovm_fc_tail      the last nn.Linear of the FC stack, the heads and the decode fused (same lines as the two above).
 *   ovm_lift_forward the inference slice of ROIHeads3D._forward_cube, roi_heads.py:329-848:
 *                    the kernels above on one stream.
 *   ovm_nms          torchvision.ops.nms / batched_nms as the reference calls them:
 *                    roi_heads_gdino.py:6,272 (class-agnostic, 0.5) and fast_rcnn.py:105 (per class).
 *   ovm_gdino_post   GroundingDINO post-processing, the step right before the lift:
 *                    get_grounding_output roi_heads_gdino.py:186-202 (sigmoid, phrase logits :296-317, box
 *                    threshold, per-query max / argmax) + grounding_dino_inference_detector :269-275
 *                    (cxcywh -> xyxy in pixels :283-287, NMS 0.5, gather in keep order).
 *   ovm_iou_box3d    box3d_overlap at evaluation: omni3d_evaluation.py:68-169 with pytorch3d's
 *                    _C.iou_box3d.
 *   ovm_box_inference the inference tail of the 2-D box head: FastRCNNOutputLayers
 *                    predict_probs / predict_boxes (d2 Box2BoxTransform.apply_deltas) and
 *                    fast_rcnn_inference_single_image, fast_rcnn.py:57-116.
 */
#ifndef OVMONO3D_LIFT_H_
#define OVMONO3D_LIFT_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OVM_ABI_VERSION 6
#define OVM_MAX_LEVELS 8
#define OVM_RECORD_WIDTH 48   /* floats per detection record, see ovm_head_decode */

enum {
  OVM_OK = 0,
  OVM_EINVAL = -1,      /* bad argument (shape, alignment, enum value) */
  OVM_ECUDA = -2,       /* a CUDA runtime / driver call failed */
  OVM_EUNSUPPORTED = -3 /* valid in the reference, not implemented here (message says what) */
};

/* Memory format of a pyramid level.  NCHW: [N, C, H, W] contiguous (what the reference's backbone emits).  NHWC:
 * torch.channels_last, [N, H, W, C] in memory.  A pyramid whose levels are ALL NHWC (16-byte aligned, C % 4 == 0) runs
 * the channels-last ROIAlign kernel; NHWC levels of a mixed pyramid are sampled by the slower direct path. */
enum { OVM_LAYOUT_NCHW = 0, OVM_LAYOUT_NHWC = 1 };
enum {
  OVM_F32 = 0,
  OVM_BF16 = 1,
  OVM_TF32 = 2, /* fp32 storage, values rounded to TF32 (10 mantissa bits, nearest): operands of a kind::tf32 layer */
  OVM_F16 = 3   /* IEEE fp16, saturated to +-65504 when produced here: TF32's mantissa in bf16's bytes (ovm_fc_tail) */
};
enum { OVM_Z_DIRECT = 0, OVM_Z_SIGMOID = 1, OVM_Z_LOG = 2, OVM_Z_CLUSTERS = 3 };
enum { OVM_POSE_6D = 0, OVM_POSE_QUATERNION = 1, OVM_POSE_EULER = 2 };
enum { OVM_DIMS_EXP = 0, OVM_DIMS_SIGMOID = 1 };
enum { OVM_GEMM_AUTO = 0, OVM_GEMM_DIRECT = 1, OVM_GEMM_STREAMK = 2 };

/* int32 elements of ovm_pooler_t.order_ws that are always enough for n boxes */
#define OVM_ORDER_WS_LEN(n) (48 * (int64_t)(n) + 2048)

/* One pyramid level: fp32 feature map of N images, C channels, contiguous in `layout`. */
typedef struct {
  const float* data;      /* device */
  int32_t height, width;
  double spatial_scale;   /* 1 / stride, as d2 ROIPooler receives it (roi_heads.py:161) */
  int32_t layout;         /* OVM_LAYOUT_* */
  int32_t _pad;
} ovm_level_t;

typedef struct {
  int32_t num_levels;
  ovm_level_t levels[OVM_MAX_LEVELS];
  int32_t num_images, channels;
  int32_t pooler_resolution;     /* MODEL.ROI_CUBE_HEAD.POOLER_RESOLUTION (7) */
  int32_t sampling_ratio;        /* POOLER_SAMPLING_RATIO (0 = adaptive) */
  int32_t canonical_box_size;    /* d2 default 224 */
  int32_t canonical_level;       /* d2 default 4 */
  float scale_roi_boxes;         /* SCALE_ROI_BOXES; <= 0 disables (roi_heads.py:310-327) */
  int32_t order_ws_len;          /* int32 elements in order_ws; OVM_ORDER_WS_LEN(n) is enough for n boxes */
  int32_t* order_ws;             /* optional device scratch (16-byte aligned), ZERO-FILLED by the caller once when it is
                                    allocated (word 1 is a ticket counter that every call hands back cleared): when
                                    given, one small launch ahead of the pooling kernel plans every box once (level,
                                    sampling geometry, bin ranges, stage geometry: a warp per box, all SMs) and its last
                                    CTA turns the boxes into work items (large footprints are cut along the channels,
                                    largest item first) so that no single CTA sets the kernel time; NULL = one CTA per
                                    box in input order, planning its box itself */
} ovm_pooler_t;

/* Multi-level ROIAlignV2.  boxes [n,4] xyxy fp32, batch_idx [n] int32 (image of each box), both
 * device.  out: [n, out_ld] of out_dtype, row = box, columns (c, ph, pw)-major exactly like
 * `cube_pooler(...).flatten(1)` (roi_heads.py:366); out_ld >= C*P*P, columns past C*P*P untouched.
 * level_out (optional, device [n] int32): 0-based level per box, -1 where the reference's float->int
 * conversion of a NaN size is undefined (such boxes pool to zeros in both).                       */
int ovm_roi_align(const ovm_pooler_t* pooler, const float* boxes, const int32_t* batch_idx, int32_t n,
                  void* out, int32_t out_dtype, int64_t out_ld, int32_t* level_out, void* stream);

/* ovm_roi_align without caller-provided scratch, for callers that cannot allocate device memory (code under CUDA graph
 * capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same arguments, results and errors as ovm_roi_align; pooler->order_ws / order_ws_len are ignored:
 * the planner's scratch (work items, per-box headers) is a file-scope `static __device__` array of the library,
 * OVM_ROI_STATIC_WS_BYTES (3.0 MiB) that every GPU which loads the library holds until the process exits.  Limits and
 * rules: n <= OVM_ROI_STATIC_MAX_BOXES, larger batches are refused with OVM_EINVAL (use ovm_roi_align); the call
 * allocates nothing on the device (no cudaMalloc / cudaFree, no allocating library routine); every call on a device shares the one
 * array, which keeps the previous call's contents (the words that must start at zero are cleared on `stream` by the
 * call itself); therefore two calls that overlap on the GPU (different streams, or graph branches that may run
 * concurrently) RACE -- order them on one stream.                                                                */
#define OVM_ROI_STATIC_MAX_BOXES 16384
#define OVM_ROI_STATIC_WS_BYTES (4 * OVM_ORDER_WS_LEN(OVM_ROI_STATIC_MAX_BOXES))
int ovm_roi_align_static(const ovm_pooler_t* pooler, const float* boxes, const int32_t* batch_idx, int32_t n,
                         void* out, int32_t out_dtype, int64_t out_ld, int32_t* level_out, void* stream);

/* D[M,N] = act(A[M,K] . B[N,K]^T + bias), A and B bf16 row-major with leading dimensions lda, ldb
 * (elements, multiples of 8; base pointers 16-byte aligned), fp32 accumulation in TMEM.
 * mode DIRECT: one CTA per 128 x BN tile, epilogue applies bias (may be NULL) and ReLU (relu != 0)
 *   and stores out_dtype.
 * mode STREAMK: K is cut so that every SM gets the same number of 64-wide K blocks; partial tiles
 *   are added into `out` (fp32, MUST be zeroed by the caller, out_dtype must be OVM_F32) with
 *   red.global.add; bias/relu are NOT applied -- follow with ovm_bias_act.
 * mode AUTO picks STREAMK when the tile count cannot fill the SMs and K is long.  The chosen mode
 * is returned through mode_used (optional).                                                     */
int ovm_gemm_bf16(const void* A, int64_t lda, const void* B, int64_t ldb, int32_t M, int32_t N, int32_t K,
                  const float* bias, int32_t relu, void* out, int32_t out_dtype, int64_t ldo,
                  int32_t mode, int32_t* mode_used, void* stream);

/* Tracing hook (tools/gemm_trace.py): while buf != NULL every ovm_gemm_bf16 launched from this host thread writes
 * sixteen %globaltimer stamps (ns) per CTA into buf[cta*16 + k], buf sized for the largest grid (<= 4096 CTAs):
 * 0 kernel entry, 1 set-up done (barriers, TMEM), 2 last TMA issued, 3 first operand tile landed, 4 last MMA
 * issued, 5 last accumulator ready, 6 epilogue done, 7 before TMEM release, 8..15 the first eight 32-column
 * chunks of the last accumulator read out of TMEM.  NULL switches it off.                                     */
int ovm_gemm_set_trace(uint64_t* buf);

/* Same for ovm_roi_align (tools/roi_trace.py): sixteen slots per CTA, buf[cta*16 + k]: 0 entry, 1 bin ranges known,
 * 2 tap tables built, 3 first stage landed, 4..11 channel chunk k written, 12 exit, 15 packed
 * (rows << 48 | footprint width << 32 | tile-width index << 24 | SM id).  CTAs on the direct path stamp slot 0 only. */
int ovm_roi_set_trace(uint64_t* buf);

/* Same for the heads + decode kernel: [grid, 8] stamps of each CTA's warp 0 -- 0 entry, 1 head rows staged,
 * 2 FC features available (programmatic-launch wait returned), 3 box / camera loads done, 4 dot products done,
 * 5 record written.  NULL switches it off.                                                                  */
int ovm_head_set_trace(uint64_t* buf);

/* Same for ovm_gdino_post (tests/perf_gdino_post.py --trace): [batch, 8] stamps of each CTA's thread 0 -- 0 entry,
 * 1 query scores loaded, 2 sorted, 3 boxes staged, 4 suppression done, 5 exit, 6 / 7 SM cycles inside the suppression
 * (pair tests << 32 | serial resolve; later positions).  NULL switches it off.               */
int ovm_post_set_trace(uint64_t* buf);

/* out[m, j] = act(in[m, j] + bias[j]) for j < N; in fp32 [M, ldi], out out_dtype [M, ldo]. */
int ovm_bias_act(const float* in, int64_t ldi, const float* bias, int32_t relu, int32_t M, int32_t N,
                 void* out, int32_t out_dtype, int64_t ldo, void* stream);

/* x (fp32 [rows, ld_in]) -> bf16 [rows, ld_out], columns >= cols zero-filled up to ld_out. */
int ovm_cast_bf16(const float* in, int64_t ld_in, int32_t rows, int32_t cols, void* out, int64_t ld_out,
                  void* stream);

typedef struct {
  int32_t z_type, pose_type, dims_priors_func;  /* OVM_Z_*, OVM_POSE_*, OVM_DIMS_* */
  int32_t cluster_bins, num_classes;
  int32_t dims_priors_enabled, use_confidence, virtual_depth, allocentric_pose;
  float virtual_focal;
  int32_t fc_dim;          /* width of one FC feature vector (MODEL.ROI_CUBE_HEAD.FC_DIM) */
  int32_t feat_offset[5];  /* column of `feats` where the features feeding deltas, dims, pose, z,
                              uncertainty start: all 0 with SHARED_FC, g*fc_dim otherwise */
} ovm_decode_cfg_t;

typedef struct {
  /* Linear heads, fp32, reference layout [out, fc_dim] (cube_head.py:109-143); uncertainty may be
   * NULL when use_confidence == 0. */
  const float *w_deltas, *b_deltas, *w_dims, *b_dims, *w_pose, *b_pose, *w_z, *b_z, *w_uncert, *b_uncert;
  const float* priors_dims_per_cat; /* [num_classes, 2, 3] (roi_heads.py:117-120) or NULL */
  const float* priors_z_scales;     /* [num_classes, cluster_bins] (roi_heads.py:123-130) or NULL */
  const float* priors_z_stats;      /* [num_classes, cluster_bins, 2] (roi_heads.py:133-143) or NULL */
} ovm_head_weights_t;

/* Heads + decode, one warp per box.
 * feats [n, ld_feats] fp32: output of the last FC layer(s).  boxes [n,4] (unscaled 2-D boxes),
 * classes [n] int64, scores [n] fp32 or NULL, batch_idx [n] int32.
 * cams [num_images, 12] fp32: K row-major (9), im_scales_ratio, H of the padded network input, 0
 * (roi_heads.py:378-400).
 * A class id outside [0, num_classes) (only looked at with per-class heads / depth bins) or an image index outside
 * [0, num_images) never indexes device memory: such a box gets a record whose decoded fields are NaN (the reference
 * raises IndexError there); the plugin validates class ids on the host when it has them there.
 * records [n, 48] fp32: bbox3D 8x3 (0..23) | center_cam (24..26) | center_2D (27..28) |
 *   dimensions W,H,L (29..31) | pose 3x3 (32..40) | 2-D box (41..44) | score (45) | class (46) | 0.
 * raw (optional) [n,16] fp32: the selected head outputs before decode: deltas(2) z(1) dims(3)
 *   uncertainty(1) pose(<=9, row-major rotation is NOT here; these are the raw 6/4/3 numbers).  */
int ovm_head_decode(const ovm_decode_cfg_t* cfg, const ovm_head_weights_t* w, const float* feats,
                    int64_t ld_feats, const float* boxes, const int64_t* classes, const float* scores,
                    const int32_t* batch_idx, const float* cams, int32_t num_images, int32_t n,
                    float* records, float* raw, void* stream);

/* One FC layer of the stack as the lift runs it. */
typedef struct {
  const void* weight;   /* bf16 [N, ldw], K-major (nn.Linear layout), device */
  const float* bias;    /* fp32 [N] */
  int32_t N, K;
  int64_t ldw;
  const void* weight_alt;   /* optional: the same weight in another operand format -- fp16 [N, ldw_alt] (ovm_cast_f16) or fp32
                               rounded to TF32 (ovm_round_tf32).  Used by the LAST layer of a stack that ends in
                               ovm_fc_tail: the layer then runs on that format, and the layer before writes its
                               activations in it, instead of bf16 x bf16 (see ovm_fc_tail) */
  int64_t ldw_alt;          /* elements; the row pitch in bytes must be a multiple of 16 */
  int32_t alt_dtype;        /* OVM_F16 or OVM_TF32 (ignored when weight_alt is NULL) */
  int32_t _pad;
} ovm_fc_layer_t;

typedef struct {
  int32_t num_groups;          /* 1 with SHARED_FC, 5 (dims, XY, pose, Z, conf) or 4 otherwise */
  int32_t num_fc;              /* layers per group (MODEL.ROI_CUBE_HEAD.NUM_FC) */
  const ovm_fc_layer_t* layers; /* host array [num_groups * num_fc], group-major */
} ovm_fc_stack_t;

typedef struct {
  void* pooled;        /* bf16 [n, ld_pooled], ld_pooled = roundup(C*P*P, 8) */
  int64_t ld_pooled;
  float* acc;          /* fp32 [n, fc_dim] split-K accumulator: ZERO-FILLED by the caller once, when it is
                          allocated; ovm_lift_forward leaves it zero-filled again (no per-call memset) */
  void* act_a;         /* [n, fc_dim] x 4 bytes: activations between FC layers, ping (bf16, or TF32 in fp32 storage) */
  void* act_b;         /* same, pong */
  float* feats;        /* fp32 [n, num_groups * fc_dim]: output of the last FC layer; not touched when the stack ends
                          in the fused ovm_fc_tail (the last layer's output never leaves the SM) */
  int32_t* levels;     /* int32 [n] */
  int32_t* tile_counters;      /* optional int32 [num_tile_counters], ZERO-FILLED once like acc (left zeroed): with
                                  it the split-K FC layer finishes (bias, ReLU, cast) inside the GEMM launch; NULL:
                                  a separate bias/ReLU launch follows.  ceil(n/128) * ceil(fc_dim/256) are needed */
  int32_t num_tile_counters;
  int32_t splitk_flags_len;    /* int32 elements in splitk_flags */
  float* splitk_slots;         /* optional fp32 [splitk_slots_len], 16-byte aligned: with it (and splitk_flags), and
                                  OVM_GEMM_SLOTS=1 in the environment, the split-K partial tiles of the first FC layer are
                                  exchanged through per-cluster slots with plain stores and summed in a fixed order
                                  (bit-reproducible, no L2 atomics) instead of red.global.add into acc; measured 5 us
                                  slower at C2, hence opt-in.  OVM_SPLITK_SLOTS_LEN floats serve any n. */
  int64_t splitk_slots_len;
  int32_t* splitk_flags;       /* int32 [splitk_flags_len], ZERO-FILLED once (left zeroed); OVM_SPLITK_FLAGS_LEN */
} ovm_workspace_t;
#define OVM_SPLITK_SLOTS_LEN (74 * 2 * 128 * 256)
#define OVM_SPLITK_FLAGS_LEN (74 * 2 * 8)

/* Last FC layer + the five linear heads + the cube decode in ONE launch (cube_head.py:157-164 last Linear + heads,
 * roi_heads.py:378-549,798-843): a cluster of N/128 CTAs per 128 boxes, the layer's output stays in tensor memory /
 * registers, per-box head partial sums are exchanged through distributed shared memory, one thread decodes a box.
 * A [n, lda]: the layer's input -- bf16 (a_dtype OVM_BF16), fp16 (OVM_F16) or fp32 values rounded to TF32 (OVM_TF32), e.g.
 * written by ovm_gemm_bf16 with that out_dtype; W [N, ldw] the layer's weight in the SAME type (nn.Linear layout), bias
 * [N] fp32.  Needs class-agnostic heads (DIMS_PRIORS_ENABLED off, CLUSTER_BINS 1), one shared FC stack and N in
 * {128, 256, 512, 1024}; returns OVM_EUNSUPPORTED otherwise (use ovm_gemm_bf16 + ovm_head_decode).  The remaining
 * arguments are those of ovm_head_decode.  Class ids are not used to index anything in this configuration.        */
int ovm_fc_tail(const void* A, int64_t lda, int32_t a_dtype, const void* W, int64_t ldw, const float* bias, int32_t N,
                int32_t K, const ovm_decode_cfg_t* cfg, const ovm_head_weights_t* w, const float* boxes,
                const int64_t* classes, const float* scores, const int32_t* batch_idx, const float* cams,
                int32_t num_images, int32_t n, float* records, float* raw, void* stream);

/* x (fp32 [rows, ld_in]) -> fp16 [rows, ld_out] (nearest even, saturated to +-65504), columns >= cols zero-filled. */
int ovm_cast_f16(const float* in, int64_t ld_in, int32_t rows, int32_t cols, void* out, int64_t ld_out, void* stream);

/* x (fp32 [rows, ld_in]) -> fp32 [rows, ld_out] rounded to TF32 (nearest, ties away), columns >= cols zero-filled. */
int ovm_round_tf32(const float* in, int64_t ld_in, int32_t rows, int32_t cols, float* out, int64_t ld_out, void* stream);

/* ROIAlign (bf16 out) -> FC stack -> heads + decode on `stream`.  With one shared FC stack of >= 2 layers and
 * class-agnostic heads the last layer, the heads and the decode run as one ovm_fc_tail launch (in the format of the
 * last layer's weight_alt -- fp16 or TF32 --, bf16 without one; OVM_FUSE_TAIL=0 in the environment selects the unfused
 * sequence). */
int ovm_lift_forward(const ovm_pooler_t* pooler, const ovm_fc_stack_t* fc, const ovm_decode_cfg_t* cfg,
                     const ovm_head_weights_t* heads, const float* boxes, const int64_t* classes,
                     const float* scores, const int32_t* batch_idx, const float* cams, int32_t n,
                     const ovm_workspace_t* ws, float* records, float* raw, void* stream);

/* The same, with the cross-rank exchange of the detections fused into the decode (SURVEY.md sections 5 / 8e; the
 * reference gathers pickled per-image lists over Gloo once per dataset, omni3d_evaluation.py:717-720).
 * peers->records[p] (p < num_peers <= OVM_MAX_PEERS) are device-visible addresses -- NVLink peer mappings of the other
 * ranks' exchange buffers (CUDA IPC / torch symmetric memory), each already offset to the row where THIS call's first
 * record belongs; the decode stores every 192-byte record to `records` and to each of them.  The caller completes the
 * exchange with one barrier across the ranks after its last call (a kernel boundary orders the posted writes).  16-byte
 * aligned pointers; NULL or num_peers == 0 = ovm_lift_forward.                                                      */
#define OVM_MAX_PEERS 8
typedef struct {
  int32_t num_peers;
  int32_t _pad;
  float* records[OVM_MAX_PEERS];
} ovm_peer_records_t;
int ovm_lift_forward_peers(const ovm_pooler_t* pooler, const ovm_fc_stack_t* fc, const ovm_decode_cfg_t* cfg,
                           const ovm_head_weights_t* heads, const float* boxes, const int64_t* classes,
                           const float* scores, const int32_t* batch_idx, const float* cams, int32_t n,
                           const ovm_workspace_t* ws, float* records, float* raw, const ovm_peer_records_t* peers,
                           void* stream);

/* Stores `value` to n <= OVM_MAX_PEERS device-visible addresses (host array of pointers) in one launch: how a rank
 * announces its record count to every rank of the fused exchange (shard.RecordExchange, "peer" transport). */
int ovm_peer_store_f32(float* const* ptrs, int32_t n, float value, void* stream);

/* ovm_lift_forward_peers in two calls, so that a host-driven caller can start the GPU before it has finished staging
 * its inputs (the reference stages everything first, roi_heads.py:378-400, then launches): OVM_PHASE_POOL plans and
 * pools (needs pooler, boxes, batch_idx, n, ws; classes / scores / cams / records may be NULL), OVM_PHASE_HEAD runs the
 * FC stack and the decode on what the pooling phase of the SAME n, ws left in ws->pooled (needs everything).  Both
 * bits = ovm_lift_forward_peers.  The two calls go to the same stream, pooling first.                               */
#define OVM_PHASE_POOL 1
#define OVM_PHASE_HEAD 2
int ovm_lift_forward_phase(const ovm_pooler_t* pooler, const ovm_fc_stack_t* fc, const ovm_decode_cfg_t* cfg,
                           const ovm_head_weights_t* heads, const float* boxes, const int64_t* classes,
                           const float* scores, const int32_t* batch_idx, const float* cams, int32_t n,
                           const ovm_workspace_t* ws, float* records, float* raw, const ovm_peer_records_t* peers,
                           int32_t phases, void* stream);

/* Measurement hook (bench.py): while set, ovm_lift_forward records events[0] before its first launch and the
 * next event after each stage -- ROIAlign, every FC layer in issue order (memset + split-K GEMM + bias/ReLU
 * count as one layer), heads+decode -- on the stream it was given, up to `count` events.  Events are
 * cudaEvent_t handles owned by the caller; count = 0 (or events = NULL) switches the hook off.  The hook is
 * per host thread and adds no synchronisation.                                                           */
int ovm_set_stage_events(void* const* events, int32_t count);

/* ---- detections either side of the lift ------------------------------------------------------------ */

#define OVM_NMS_MAX_SEGMENT 4096  /* boxes of one segment that one CTA sorts and suppresses in shared memory */
enum { OVM_NMS_AGNOSTIC = 0, OVM_NMS_COORD_TRICK = 1, OVM_NMS_PER_CLASS = 2 };

/* Greedy NMS, one cluster of 8 CTAs per segment (image; single CTAs with OVM_NMS_CLUSTER=1 in the environment); no host
 * round trip (torchvision's CUDA nms copies its mask to the host).
 * boxes [n,4] xyxy fp32, scores [n] fp32, classes [n] int64 (NULL with OVM_NMS_AGNOSTIC), seg_offsets [S+1] int32
 * (device): segment s = rows [seg_offsets[s], seg_offsets[s+1]), each at most max_segment_len (<= OVM_NMS_MAX_SEGMENT)
 * long.  Arithmetic is torchvision's CPU kernel (fp32 areas, inter / (a_i + a_j - inter) > threshold compared in
 * double, stable descending score order), so the keep list is bit-identical.
 * mode OVM_NMS_COORD_TRICK: torchvision batched_nms' `boxes + class * (max_coordinate + 1)` per segment;
 * mode OVM_NMS_PER_CLASS: its `_batched_nms_vanilla` (boxes of different classes never suppress each other).
 * keep [n] int32: keep[seg_offsets[s] + k], k < keep_count[s] = row (relative to the segment start) of the k-th
 * kept box in decreasing score order; keep_count [S] int32 (-1 for a segment longer than max_segment_len). */
int ovm_nms(const float* boxes, const float* scores, const int64_t* classes, const int32_t* seg_offsets,
            int32_t num_segments, int32_t max_segment_len, double iou_threshold, int32_t mode, int32_t* keep,
            int32_t* keep_count, void* stream);

/* GroundingDINO outputs -> prompt boxes of the lift in two launches: query scores (a warp per query, all SMs), then
 * one CTA per image for everything else.
 * logits [B, nq, T] fp32 raw `pred_logits` (sigmoid is applied here), boxes_cxcywh [B, nq, 4] fp32 `pred_boxes`
 * (normalised), phrase_ranges [P, 2] int32 (device): token columns [begin, end) of phrase p in the tokenised caption
 * (roi_heads_gdino.py:300-311).  A query survives when max_p sum_t sigmoid(logit[q, t in phrase p]) > box_threshold;
 * its score / class are that max / argmax (first on ties); boxes become xyxy pixels with image_width / image_height
 * (the padded network input, :269); class-agnostic NMS at nms_threshold; survivors are written in decreasing score
 * order: out_boxes [B, nq, 4], out_scores [B, nq], out_classes [B, nq] int64, out_query [B, nq] int32 (source query,
 * optional), out_count [B] int32.  Rows past out_count[b] are undefined (out_scores / out_classes carry the per-query
 * scores between the two launches).  nq <= OVM_NMS_MAX_SEGMENT, T <= 512. */
int ovm_gdino_post(const float* logits, const float* boxes_cxcywh, const int32_t* phrase_ranges, int32_t batch,
                   int32_t num_queries, int32_t num_tokens, int32_t num_phrases, int32_t image_height,
                   int32_t image_width, float box_threshold, double nms_threshold, float* out_boxes,
                   float* out_scores, int64_t* out_classes, int32_t* out_query, int32_t* out_count, void* stream);

typedef struct {
  int32_t num_classes;           /* K foreground classes: logits have K + 1 columns, background last */
  int32_t cls_agnostic_bbox_reg; /* MODEL.ROI_BOX_HEAD.CLS_AGNOSTIC_BBOX_REG: deltas have 4 columns instead of 4K */
  float reg_weights[4];          /* MODEL.ROI_BOX_HEAD.BBOX_REG_WEIGHTS (10, 10, 5, 5) */
  float scale_clamp;             /* d2 Box2BoxTransform: log(1000 / 16) */
  float score_thresh;            /* MODEL.ROI_HEADS.SCORE_THRESH_TEST */
  double nms_thresh;             /* MODEL.ROI_HEADS.NMS_THRESH_TEST */
  int32_t topk_per_image;        /* TEST.DETECTIONS_PER_IMAGE; < 0 keeps everything */
  int32_t coord_trick_max;       /* candidates per image up to which torchvision's batched_nms offsets the boxes by
                                    class (1000 on the CPU, 5000 on CUDA); above it runs class by class */
} ovm_box_inference_cfg_t;

typedef struct {                 /* device scratch owned by the caller; nothing needs initialising */
  float* probs;                  /* [R, K] */
  float* cand_boxes;             /* [R, Kb, 4], Kb = 1 when class-agnostic else K; 16-byte aligned */
  int32_t* class_keep;           /* [N, K, max_per_image] */
  int32_t* class_count;          /* [N, K] */
  int32_t* image_stats;          /* [N, 2] */
} ovm_box_inference_ws_t;

/* Inference tail of the 2-D box head: softmax over logits [R, ld_logits] (K + 1 used), Box2BoxTransform.apply_deltas
 * of deltas [R, ld_deltas] on proposals [R, 4] (xyxy), clip to image_sizes [N, 2] int32 (height, width), rows with a
 * non-finite number dropped, score threshold, per-class NMS, first top-k per image in decreasing score order
 * (fast_rcnn.py:57-116).  seg_offsets [N + 1] int32 (device): proposals of image i are rows
 * [seg_offsets[i], seg_offsets[i+1]), at most max_per_image (<= OVM_NMS_MAX_SEGMENT) of them.  Three launches.
 * Outputs, cap = topk_per_image (or K * max_per_image when < 0) rows per image: out_boxes [N, cap, 4],
 * out_scores [N, cap], out_classes [N, cap] int64, out_index [N, cap] int32 (proposal row inside the image; the
 * reference's kept_indices count only the finite rows), out_scores_full [N, cap, K] (optional), out_count [N]. */
int ovm_box_inference(const ovm_box_inference_cfg_t* cfg, const float* logits, int64_t ld_logits,
                      const float* deltas, int64_t ld_deltas, const float* proposals, const int32_t* seg_offsets,
                      const int32_t* image_sizes, int32_t num_images, int32_t num_proposals, int32_t max_per_image,
                      const ovm_box_inference_ws_t* ws, float* out_boxes, float* out_scores, int64_t* out_classes,
                      int32_t* out_index, float* out_scores_full, int32_t* out_count, void* stream);

/* Detection records -> the rows of the reference's result file: d2 detector_postprocess as
 * GeneralizedRCNN._postprocess runs it (cubercnn/modeling/meta_arch/rcnn3d.py:113-115: 2-D boxes scaled from the network
 * resolution to the original image, clipped, empty ones dropped) and the XYXY -> XYWH conversion of
 * Omni3DEvaluator.instances_to_coco_json (cubercnn/evaluation/omni3d_evaluation.py:1218-1219).
 * records [n,48] (ovm_head_decode layout), offsets [N+1] int32 (records of image i are rows [offsets[i], offsets[i+1])),
 * geom [N,4] fp32: scale_x = out_w / net_w and scale_y = out_h / net_h (computed in double, rounded to fp32, as torch
 * does for a Python-float factor), out_w, out_h.  out [n,48] (must not alias records): the surviving rows of image i,
 * in order, from row offsets[i]; columns 41..44 hold x, y, w, h in original-resolution pixels, the rest is copied.
 * out_count [N] int32.  One CTA per image, one launch; bit-identical to the reference's fp32 arithmetic.         */
int ovm_records_postprocess(const float* records, const int32_t* offsets, const float* geom, int32_t num_images,
                            float* out, int32_t* out_count, void* stream);

/* ---- evaluation side ---------------------------------------------------------------------- */

/* Exact 3-D IoU of oriented cuboids: `box3d_overlap` of the reference's evaluator (cubercnn/evaluation/
 * omni3d_evaluation.py:109-169) = pytorch3d `_C.iou_box3d` (:158, forced onto the CPU there by
 * MAX_DTS_CROSS_GTS_FOR_IOU3D = 0, :65) + the `_check_coplanar` / `_check_nonzero` guards on the detections (:68-107).
 * boxes_dt [N,8,3], boxes_gt [M,8,3] fp32 (device): the 8 corners in pytorch3d's order (the order of the docstring at
 * :116-143; `get_cuboid_verts_faces` emits it).  iou [N,M] = vol / (vol_dt + vol_gt - vol); vol (optional) [N,M] the
 * intersection volumes.  dt_flags (optional, device int32 [N]): when given, the guards run first -- flags[i] bit 0 =
 * detection i fails the coplanarity test (eps_coplanar, reference default 1e-4), bit 1 = it has a face triangle of
 * area <= eps_nonzero (1e-8) -- and the IoUs of flagged detections are zero, as in the reference (:160-167).
 * One warp per pair, two launches.                                                                               */
int ovm_iou_box3d(const float* boxes_dt, const float* boxes_gt, int32_t N, int32_t M, float eps_coplanar,
                  float eps_nonzero, float* iou, float* vol, int32_t* dt_flags, void* stream);

int ovm_abi_version(void);
/* sizeof() of the ABI structs as this library was compiled, for binding self-checks:
 * 0 ovm_level_t, 1 ovm_pooler_t, 2 ovm_decode_cfg_t, 3 ovm_head_weights_t, 4 ovm_fc_layer_t,
 * 5 ovm_fc_stack_t, 6 ovm_workspace_t, 7 ovm_box_inference_cfg_t, 8 ovm_box_inference_ws_t, 9 ovm_peer_records_t;
 * -1 for any other index. */
int ovm_struct_size(int which);
const char* ovm_last_error(void);
/* Number of kernel launches (incl. memsets) the library has issued in this process. */
int64_t ovm_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* OVMONO3D_LIFT_H_ */
