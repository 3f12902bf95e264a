// C ABI of libovmono3d_lift.so (declared in include/ovmono3d_lift.h) and the three-kernel lift sequence.
//
// ovm_lift_forward is the inference slice of ROIHeads3D._forward_cube (reference
// cubercnn/modeling/roi_heads/roi_heads.py:329-848): cube_pooler (:366) -> cube_head (:430) -> decode (:378-549,
// 798-843), issued back to back on the caller's stream with no host synchronisation in between.
#include <cstdarg>
#include <cstring>

#include "common.cuh"

namespace ovm {

std::atomic<int64_t> g_launches{0};

namespace {
thread_local char g_error[512] = "";
constexpr int kMaxStageEvents = 32;
thread_local cudaEvent_t g_stage_events[kMaxStageEvents];
thread_local int g_num_stage_events = 0;

}

// bench.py's per-stage timing hook (ovm_set_stage_events): record the idx-th event if one was supplied
int stage_mark(int& idx, cudaStream_t stream) {
  if (idx < g_num_stage_events) OVM_CUDA(cudaEventRecord(g_stage_events[idx], stream));
  ++idx;
  return OVM_OK;
}

EncodeTiledFn tensormap_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess || !p)
    return nullptr;
  fn = reinterpret_cast<EncodeTiledFn>(p);
  return fn;
}

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}

namespace {

// One FC stack (cube_head.py:64-74): every layer is Linear + ReLU.  The last layer's output is the fp32
// feature block `feats[:, col0 : col0 + N]`; earlier layers ping-pong through bf16 activations.
// `mid_dtype`: storage of the activations between layers (OVM_BF16, or OVM_TF32 for a layer that feeds a kind::tf32
// tail).  `num_fc` may stop short of the stack's last layer (fused tail): then every layer run here is a middle one
// and *out_last / *ld_last describe the activations handed to the tail.
int run_fc_group(const ovm_fc_layer_t* layers, int num_fc, bool ends_stack, int last_mid_dtype, const void* x, int64_t ldx,
                 int32_t n, const ovm_workspace_t* ws, float* feats, int64_t ld_feats, cudaStream_t stream, int& stage,
                 const void** out_last = nullptr, int64_t* ld_last = nullptr, TailPending* defer_last = nullptr) {
  const void* in = x;
  int64_t ld_in = ldx;
  for (int k = 0; k < num_fc; ++k) {
    const ovm_fc_layer_t& L = layers[k];
    const bool last = ends_stack && k == num_fc - 1;
    OVM_REQUIRE(L.weight && L.bias && L.N > 0 && L.K > 0, "ovm_lift_forward: FC layer %d is incomplete", k);
    OVM_REQUIRE(L.N % 8 == 0, "ovm_lift_forward: FC_DIM must be a multiple of 8 (got %d)", L.N);
    void* out = last ? static_cast<void*>(feats) : (k % 2 == 0 ? ws->act_a : ws->act_b);
    const int64_t ldo = last ? ld_feats : L.N;
    const int out_dtype = last ? OVM_F32 : (k == num_fc - 1 ? last_mid_dtype : OVM_BF16);
    // Ask the GEMM which mode it would pick for a raw fp32 accumulation; a long-K layer whose tiles cannot fill
    // the SMs runs split-K into ws->acc and gets bias+ReLU from the small follow-up kernel.
    int mode = OVM_GEMM_DIRECT;
    {
      const int BN = L.N >= 256 ? 256 : (L.N >= 128 ? 128 : 64);
      const int tiles = ((n + 127) / 128) * ((L.N + BN - 1) / BN);
      const int kb = (L.K + 63) / 64;
      const bool ragged = tiles < kNumSMs || (tiles % kNumSMs) * 2 < kNumSMs;
      if (kb >= 32 && tiles < 2 * kNumSMs && ragged && ws->acc) mode = OVM_GEMM_STREAMK;
    }
    if (mode == OVM_GEMM_STREAMK) {
      // ws->acc (and the tile counters) are zero on entry (workspace contract) and handed back zeroed
      if (ws->tile_counters && ws->num_tile_counters >= ((n + 127) / 128) * ((L.N + 255) / 256)) {
        // one launch: the CTA that completes a tile applies bias + ReLU and writes the layer's output
        GemmFinish fin{ws->tile_counters, ws->num_tile_counters, L.bias, 1, out, out_dtype, ldo,
                       ws->splitk_slots, ws->splitk_slots_len, ws->splitk_flags, ws->splitk_flags_len};
        if (defer_last && k == num_fc - 1 && !ends_stack) {
          // the fused tail finishes this layer on its way in (TailPending): the GEMM only accumulates into ws->acc
          fin.defer = 1;
          *defer_last = TailPending{ws->acc, L.N, L.bias};
        }
        int rc = gemm_bf16_launch(in, ld_in, L.weight, L.ldw, n, L.N, L.K, nullptr, 0, ws->acc, OVM_F32, L.N,
                                  OVM_GEMM_STREAMK, nullptr, stream, &fin);
        if (rc != OVM_OK) return rc;
      } else {
        int rc = gemm_bf16_launch(in, ld_in, L.weight, L.ldw, n, L.N, L.K, nullptr, 0, ws->acc, OVM_F32, L.N,
                                  OVM_GEMM_STREAMK, nullptr, stream);
        if (rc != OVM_OK) return rc;
        rc = bias_act_launch(ws->acc, L.N, L.bias, 1, n, L.N, out, out_dtype, ldo, stream, /*clear_in=*/true);
        if (rc != OVM_OK) return rc;
      }
    } else {
      int rc = gemm_bf16_launch(in, ld_in, L.weight, L.ldw, n, L.N, L.K, L.bias, 1, out, out_dtype, ldo,
                                OVM_GEMM_DIRECT, nullptr, stream);
      if (rc != OVM_OK) return rc;
    }
    in = out;
    ld_in = ldo;
    int rc = stage_mark(stage, stream);
    if (rc != OVM_OK) return rc;
  }
  if (out_last) *out_last = in;
  if (ld_last) *ld_last = ld_in;
  return OVM_OK;
}

}  // namespace
}  // namespace ovm

using namespace ovm;

#pragma GCC visibility push(default)
extern "C" {

int ovm_abi_version(void) { return OVM_ABI_VERSION; }

const char* ovm_last_error(void) { return g_error; }

int ovm_struct_size(int which) {
  switch (which) {
    case 0: return (int)sizeof(ovm_level_t);
    case 1: return (int)sizeof(ovm_pooler_t);
    case 2: return (int)sizeof(ovm_decode_cfg_t);
    case 3: return (int)sizeof(ovm_head_weights_t);
    case 4: return (int)sizeof(ovm_fc_layer_t);
    case 5: return (int)sizeof(ovm_fc_stack_t);
    case 6: return (int)sizeof(ovm_workspace_t);
    case 7: return (int)sizeof(ovm_box_inference_cfg_t);
    case 8: return (int)sizeof(ovm_box_inference_ws_t);
    case 9: return (int)sizeof(ovm_peer_records_t);
    default: return -1;
  }
}

int64_t ovm_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int ovm_set_stage_events(void* const* events, int32_t count) {
  OVM_REQUIRE(count >= 0 && count <= kMaxStageEvents, "ovm_set_stage_events: count %d not in [0,%d]", count,
              kMaxStageEvents);
  OVM_REQUIRE(count == 0 || events, "ovm_set_stage_events: events is NULL");
  for (int i = 0; i < count; ++i) g_stage_events[i] = static_cast<cudaEvent_t>(events[i]);
  g_num_stage_events = events ? count : 0;
  return OVM_OK;
}

int ovm_roi_align(const ovm_pooler_t* pooler, const float* boxes, const int32_t* batch_idx, int32_t n, void* out,
                  int32_t out_dtype, int64_t out_ld, int32_t* level_out, void* stream) {
  return roi_align_launch(pooler, boxes, batch_idx, n, out, out_dtype, out_ld, level_out,
                          static_cast<cudaStream_t>(stream));
}

int ovm_gemm_bf16(const void* A, int64_t lda, const void* B, int64_t ldb, int32_t M, int32_t N, int32_t K,
                  const float* bias, int32_t relu, void* out, int32_t out_dtype, int64_t ldo, int32_t mode,
                  int32_t* mode_used, void* stream) {
  return gemm_bf16_launch(A, lda, B, ldb, M, N, K, bias, relu, out, out_dtype, ldo, mode, mode_used,
                          static_cast<cudaStream_t>(stream));
}

int ovm_gemm_set_trace(uint64_t* buf) {
  gemm_set_trace(reinterpret_cast<unsigned long long*>(buf));
  return OVM_OK;
}

int ovm_roi_set_trace(uint64_t* buf) {
  roi_set_trace(reinterpret_cast<unsigned long long*>(buf));
  return OVM_OK;
}

int ovm_head_set_trace(uint64_t* buf) {
  head_set_trace(reinterpret_cast<unsigned long long*>(buf));
  return OVM_OK;
}

int ovm_post_set_trace(uint64_t* buf) {
  post_set_trace(reinterpret_cast<unsigned long long*>(buf));
  return OVM_OK;
}

int ovm_bias_act(const float* in, int64_t ldi, const float* bias, int32_t relu, int32_t M, int32_t N, void* out,
                 int32_t out_dtype, int64_t ldo, void* stream) {
  return bias_act_launch(in, ldi, bias, relu, M, N, out, out_dtype, ldo, static_cast<cudaStream_t>(stream));
}

int ovm_cast_bf16(const float* in, int64_t ld_in, int32_t rows, int32_t cols, void* out, int64_t ld_out,
                  void* stream) {
  return cast_bf16_launch(in, ld_in, rows, cols, out, ld_out, static_cast<cudaStream_t>(stream));
}

int ovm_head_decode(const ovm_decode_cfg_t* cfg, const ovm_head_weights_t* w, const float* feats, int64_t ld_feats,
                    const float* boxes, const int64_t* classes, const float* scores, const int32_t* batch_idx,
                    const float* cams, int32_t num_images, int32_t n, float* records, float* raw, void* stream) {
  return head_decode_launch(cfg, w, feats, ld_feats, boxes, classes, scores, batch_idx, cams, num_images, n, records,
                            raw, static_cast<cudaStream_t>(stream));
}

int ovm_fc_tail(const void* A, int64_t lda, int32_t a_dtype, const void* W, int64_t ldw, const float* bias, int32_t N,
                int32_t K, const ovm_decode_cfg_t* cfg, const ovm_head_weights_t* w, const float* boxes,
                const int64_t* classes, const float* scores, const int32_t* batch_idx, const float* cams,
                int32_t num_images, int32_t n, float* records, float* raw, void* stream) {
  return fc_tail_launch(A, lda, a_dtype, W, ldw, bias, N, K, cfg, w, boxes, classes, scores, batch_idx, cams, num_images,
                        n, records, raw, static_cast<cudaStream_t>(stream));
}

int ovm_cast_f16(const float* in, int64_t ld_in, int32_t rows, int32_t cols, void* out, int64_t ld_out, void* stream) {
  return cast_f16_launch(in, ld_in, rows, cols, out, ld_out, static_cast<cudaStream_t>(stream));
}

int ovm_round_tf32(const float* in, int64_t ld_in, int32_t rows, int32_t cols, float* out, int64_t ld_out, void* stream) {
  return round_tf32_launch(in, ld_in, rows, cols, out, ld_out, static_cast<cudaStream_t>(stream));
}

int ovm_lift_forward(const ovm_pooler_t* pooler, const ovm_fc_stack_t* fc, const ovm_decode_cfg_t* cfg,
                     const ovm_head_weights_t* heads, const float* boxes, const int64_t* classes, const float* scores,
                     const int32_t* batch_idx, const float* cams, int32_t n, const ovm_workspace_t* ws,
                     float* records, float* raw, void* stream_) {
  return ovm_lift_forward_peers(pooler, fc, cfg, heads, boxes, classes, scores, batch_idx, cams, n, ws, records, raw,
                                nullptr, stream_);
}

int ovm_lift_forward_peers(const ovm_pooler_t* pooler, const ovm_fc_stack_t* fc, const ovm_decode_cfg_t* cfg,
                           const ovm_head_weights_t* heads, const float* boxes, const int64_t* classes,
                           const float* scores, const int32_t* batch_idx, const float* cams, int32_t n,
                           const ovm_workspace_t* ws, float* records, float* raw, const ovm_peer_records_t* peers,
                           void* stream_) {
  return ovm_lift_forward_phase(pooler, fc, cfg, heads, boxes, classes, scores, batch_idx, cams, n, ws, records, raw, peers,
                                OVM_PHASE_POOL | OVM_PHASE_HEAD, stream_);
}

int ovm_lift_forward_phase(const ovm_pooler_t* pooler, const ovm_fc_stack_t* fc, const ovm_decode_cfg_t* cfg,
                           const ovm_head_weights_t* heads, const float* boxes, const int64_t* classes,
                           const float* scores, const int32_t* batch_idx, const float* cams, int32_t n,
                           const ovm_workspace_t* ws, float* records, float* raw, const ovm_peer_records_t* peers,
                           int32_t phases, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  OVM_REQUIRE(phases >= 1 && phases <= 3, "ovm_lift_forward_phase: phases = %d (OVM_PHASE_POOL | OVM_PHASE_HEAD)", phases);
  if (peers && peers->num_peers == 0) peers = nullptr;
  OVM_REQUIRE(pooler && fc && cfg && heads && ws, "ovm_lift_forward: NULL descriptor");
  OVM_REQUIRE(fc->num_groups >= 1 && fc->num_groups <= 5 && fc->num_fc >= 1 && fc->layers,
              "ovm_lift_forward: bad FC stack (%d groups x %d layers)", fc->num_groups, fc->num_fc);
  if (n == 0) return OVM_OK;  // roi_heads.py:368-372
  OVM_REQUIRE(n > 0, "ovm_lift_forward: n=%d", n);
  const int PP = pooler->pooler_resolution * pooler->pooler_resolution;
  const int64_t k_in = (int64_t)pooler->channels * PP;
  OVM_REQUIRE(ws->pooled && ws->feats && ws->ld_pooled >= k_in && ws->ld_pooled % 8 == 0,
              "ovm_lift_forward: workspace.pooled must be [n, ld_pooled >= C*P*P], ld_pooled %% 8 == 0");
  OVM_REQUIRE(fc->num_fc == 1 || (ws->act_a && ws->act_b), "ovm_lift_forward: workspace.act_a/act_b missing");
  int stage = 0;
  int rc = OVM_OK;
  if (phases & OVM_PHASE_POOL) {
    rc = stage_mark(stage, stream);
    if (rc != OVM_OK) return rc;
    rc = roi_align_launch(pooler, boxes, batch_idx, n, ws->pooled, OVM_BF16, ws->ld_pooled, ws->levels, stream, &stage);
    if (rc != OVM_OK) return rc;
    rc = stage_mark(stage, stream);
    if (rc != OVM_OK) return rc;
  } else {
    stage = 3;     // the three events of the pooling phase belong to the earlier call
  }
  if (!(phases & OVM_PHASE_HEAD)) return OVM_OK;
  OVM_REQUIRE(classes && cams && records, "ovm_lift_forward: classes, cams and records are needed by the head phase");
  // fc1 reads only the first C*P*P columns of each pooled row (its TMA map is K wide), so the pad is never touched
  const int64_t ld_feats = (int64_t)fc->num_groups * cfg->fc_dim;
  const ovm_fc_layer_t& tail = fc->layers[fc->num_fc - 1];
  if (fc->num_groups == 1 && fc->num_fc >= 2 && fc_tail_supported(cfg, tail.N, tail.K) && tail.weight && tail.bias) {
    // one shared stack, class-agnostic heads: [fc1 .. fc(k-1)] -> {fc k + heads + decode} in one launch
    OVM_REQUIRE(fc->layers[0].K == k_in, "ovm_lift_forward: fc1 expects K=%d but the pooler produces %lld",
                fc->layers[0].K, (long long)k_in);
    const bool alt = tail.weight_alt != nullptr;
    OVM_REQUIRE(!alt || tail.alt_dtype == OVM_F16 || tail.alt_dtype == OVM_TF32,
                "ovm_lift_forward: weight_alt must be OVM_F16 or OVM_TF32 (got %d)", tail.alt_dtype);
    const int fmt = alt ? tail.alt_dtype : OVM_BF16;
    const void* h = nullptr;
    int64_t ldh = 0;
    // OVM_TAIL_DEFER=0: the split-K layer before the tail finishes itself (per-tile counters, round 2's first version)
    // (and OVM_GEMM_SLOTS=1, the bit-reproducible split-K exchange, keeps the finish inside the GEMM: its owners sum the
    // partials in a fixed order; the deferred finish reads sums that met in L2 atomics)
    static const bool defer_ok = [] {
      const char* v = getenv("OVM_TAIL_DEFER");
      const char* sl = getenv("OVM_GEMM_SLOTS");
      return (!v || atoi(v) != 0) && !(sl && atoi(sl) != 0);
    }();
    TailPending pending{nullptr, 0, nullptr};
    rc = run_fc_group(fc->layers, fc->num_fc - 1, /*ends_stack=*/false, fmt, ws->pooled, ws->ld_pooled, n, ws, nullptr, 0,
                      stream, stage, &h, &ldh, defer_ok && n >= 256 ? &pending : nullptr);   // (80 boxes: the finish is
                      // cheap and the tail's conversion is not hidden -- measured 69.4 vs 71.1 us per step at C1)
    if (rc != OVM_OK) return rc;
    rc = fc_tail_launch(h, ldh, fmt, alt ? tail.weight_alt : tail.weight, alt ? tail.ldw_alt : tail.ldw, tail.bias, tail.N,
                        tail.K, cfg, heads, boxes, classes, scores, batch_idx, cams, pooler->num_images, n, records, raw,
                        stream, peers, pending.acc ? &pending : nullptr);
    if (rc != OVM_OK) return rc;
    return stage_mark(stage, stream);
  }
  for (int g = 0; g < fc->num_groups; ++g) {
    const ovm_fc_layer_t* layers = fc->layers + (size_t)g * fc->num_fc;
    OVM_REQUIRE(layers[0].K == k_in, "ovm_lift_forward: fc1 expects K=%d but the pooler produces %lld",
                layers[0].K, (long long)k_in);
    OVM_REQUIRE(layers[fc->num_fc - 1].N == cfg->fc_dim, "ovm_lift_forward: last FC width != cfg.fc_dim");
    rc = run_fc_group(layers, fc->num_fc, /*ends_stack=*/true, OVM_BF16, ws->pooled, ws->ld_pooled, n, ws,
                      ws->feats + (size_t)g * cfg->fc_dim, ld_feats, stream, stage);
    if (rc != OVM_OK) return rc;
  }
  rc = head_decode_launch(cfg, heads, ws->feats, ld_feats, boxes, classes, scores, batch_idx, cams,
                          pooler->num_images, n, records, raw, stream, peers);
  if (rc != OVM_OK) return rc;
  return stage_mark(stage, stream);
}

int ovm_nms(const float* boxes, const float* scores, const int64_t* classes, const int32_t* seg_offsets,
            int32_t num_segments, int32_t max_segment_len, double iou_threshold, int32_t mode, int32_t* keep,
            int32_t* keep_count, void* stream) {
  return nms_launch(boxes, scores, classes, seg_offsets, num_segments, max_segment_len, iou_threshold, mode, keep,
                    keep_count, static_cast<cudaStream_t>(stream));
}

int ovm_gdino_post(const float* logits, const float* boxes_cxcywh, const int32_t* phrase_ranges, int32_t batch,
                   int32_t num_queries, int32_t num_tokens, int32_t num_phrases, int32_t image_height,
                   int32_t image_width, float box_threshold, double nms_threshold, float* out_boxes,
                   float* out_scores, int64_t* out_classes, int32_t* out_query, int32_t* out_count, void* stream) {
  return gdino_post_launch(logits, boxes_cxcywh, phrase_ranges, batch, num_queries, num_tokens, num_phrases,
                           image_height, image_width, box_threshold, nms_threshold, out_boxes, out_scores,
                           out_classes, out_query, out_count, static_cast<cudaStream_t>(stream));
}

int ovm_box_inference(const ovm_box_inference_cfg_t* cfg, const float* logits, int64_t ld_logits,
                      const float* deltas, int64_t ld_deltas, const float* proposals, const int32_t* seg_offsets,
                      const int32_t* image_sizes, int32_t num_images, int32_t num_proposals, int32_t max_per_image,
                      const ovm_box_inference_ws_t* ws, float* out_boxes, float* out_scores, int64_t* out_classes,
                      int32_t* out_index, float* out_scores_full, int32_t* out_count, void* stream) {
  return box_inference_launch(cfg, logits, ld_logits, deltas, ld_deltas, proposals, seg_offsets, image_sizes,
                              num_images, num_proposals, max_per_image, ws, out_boxes, out_scores, out_classes,
                              out_index, out_scores_full, out_count, static_cast<cudaStream_t>(stream));
}

int ovm_records_postprocess(const float* records, const int32_t* offsets, const float* geom, int32_t num_images,
                            float* out, int32_t* out_count, void* stream) {
  return records_post_launch(records, offsets, geom, num_images, out, out_count, static_cast<cudaStream_t>(stream));
}

int ovm_peer_store_f32(float* const* ptrs, int32_t n, float value, void* stream) {
  return peer_store_launch(ptrs, n, value, static_cast<cudaStream_t>(stream));
}

int ovm_iou_box3d(const float* boxes_dt, const float* boxes_gt, int32_t N, int32_t M, float eps_coplanar,
                  float eps_nonzero, float* iou, float* vol, int32_t* dt_flags, void* stream) {
  return iou_box3d_launch(boxes_dt, boxes_gt, N, M, eps_coplanar, eps_nonzero, iou, vol, dt_flags,
                          static_cast<cudaStream_t>(stream));
}

}  // extern "C"
#pragma GCC visibility pop
