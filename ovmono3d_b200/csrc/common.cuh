// Shared helpers for the lift kernels (sm_100a only).
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdio>

#include "../../include/ovmono3d_lift.h"

namespace ovm {

void set_error(const char* fmt, ...);
extern std::atomic<int64_t> g_launches;

inline int check_cuda(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return OVM_OK;
  set_error("%s: %s", what, cudaGetErrorString(e));
  return OVM_ECUDA;
}

#define OVM_CUDA(call)                                         \
  do {                                                         \
    int _rc = ::ovm::check_cuda((call), #call);                \
    if (_rc != OVM_OK) return _rc;                             \
  } while (0)

#define OVM_LAUNCHED(name)                                     \
  do {                                                         \
    ::ovm::g_launches.fetch_add(1, std::memory_order_relaxed); \
    int _rc = ::ovm::check_cuda(cudaGetLastError(), name);     \
    if (_rc != OVM_OK) return _rc;                             \
  } while (0)

#define OVM_REQUIRE(cond, ...)       \
  do {                               \
    if (!(cond)) {                   \
      ::ovm::set_error(__VA_ARGS__); \
      return OVM_EINVAL;             \
    }                                \
  } while (0)

constexpr int kNumSMs = 148;  // B200

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device (per-context) property of a kernel: remember, per device,
// the largest size already granted, so that a process driving several GPUs opts in on each of them and the
// steady state costs one cudaGetDevice.  Racing host threads at worst both set the attribute.
struct SmemOptIn {
  static constexpr int kMaxDevices = 64;
  std::atomic<int> granted[kMaxDevices];
  template <typename F>
  int ensure(F kernel, size_t bytes) {
    int dev = 0;
    OVM_CUDA(cudaGetDevice(&dev));
    std::atomic<int>* slot = (dev >= 0 && dev < kMaxDevices) ? &granted[dev] : nullptr;
    if (slot && (int)bytes <= slot->load(std::memory_order_relaxed)) return OVM_OK;
    OVM_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    if (slot) slot->store((int)bytes, std::memory_order_relaxed);
    return OVM_OK;
  }
};

// Programmatic dependent launch: every kernel of the lift is launched with the programmatic-stream-serialization
// attribute, tells the scheduler right away that its dependents may be launched (pdl_trigger), and waits for the
// kernel before it in the stream (pdl_wait: full completion, memory visible) only where it first touches that
// kernel's results.  What sits before the wait -- barrier / TMEM set-up, tensor-map prefetch, staging of weights --
// overlaps the tail of the previous kernel, and the launch latency disappears from the chain.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no -lcuda link dependency); NULL if absent.
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn tensormap_encode_fn();

// Entry points implemented per translation unit (host side, called from lift_api.cu).
// bench.py's per-stage timing hook (ovm_set_stage_events): records the idx-th event if one was supplied, ++idx
int stage_mark(int& idx, cudaStream_t stream);
// stage_idx (optional): a stage mark is recorded between the work-item kernel and the pooling kernel
int roi_align_launch(const ovm_pooler_t* pooler, const float* boxes, const int32_t* batch_idx, int32_t n,
                     void* out, int32_t out_dtype, int64_t out_ld, int32_t* level_out, cudaStream_t stream,
                     int* stage_idx = nullptr);
// Split-K with a fused finish (lift_api.cu): `out` of gemm_bf16_launch is the zeroed fp32 accumulator, the CTA that
// completes a tile writes act(acc + bias) to `out` here and zeroes accumulator and counter again.
#ifdef __CUDACC__
// round-to-nearest (ties away) to TF32's 10 mantissa bits; the value stays an fp32 bit pattern
__device__ __forceinline__ float round_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
__device__ __forceinline__ float4 round_tf32(float4 v) {
  return make_float4(round_tf32(v.x), round_tf32(v.y), round_tf32(v.z), round_tf32(v.w));
}
// four fp32 -> four 16-bit values: bf16 (round to nearest even) or fp16 (the same, saturated to +-65504 so that no
// infinity is made; NaN stays NaN).  fp16 carries TF32's 10 mantissa bits in bf16's 2 bytes: the format of the
// activations handed to a kind::f16 tail layer (fc_tail.cu).
__device__ __forceinline__ float sat_f16(float x) { return x > 65504.f ? 65504.f : (x < -65504.f ? -65504.f : x); }
__device__ __forceinline__ uint2 pack16(float4 v, int f16) {
  uint2 pk;
  if (f16) {
    __half2 lo = __floats2half2_rn(sat_f16(v.x), sat_f16(v.y)), hi = __floats2half2_rn(sat_f16(v.z), sat_f16(v.w));
    pk.x = *reinterpret_cast<uint32_t*>(&lo);
    pk.y = *reinterpret_cast<uint32_t*>(&hi);
  } else {
    __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
    pk.x = *reinterpret_cast<uint32_t*>(&lo);
    pk.y = *reinterpret_cast<uint32_t*>(&hi);
  }
  return pk;
}
#endif

struct GemmFinish {
  int32_t* tile_counters;   // device, zero on entry, >= num_counters entries
  int32_t num_counters;
  const float* bias;
  int32_t relu;
  void* out;
  int32_t out_dtype;
  int64_t ldo;
  // optional: deterministic split-K exchange (gemm_pair_kernel<true>), see ovm_workspace_t.splitk_slots
  float* slots = nullptr;
  int64_t slots_len = 0;
  int32_t* slot_flags = nullptr;
  int32_t flags_len = 0;
  int32_t defer = 0;        // accumulate only: no counters, no finish -- the consumer applies bias / ReLU (TailPending)
};
int gemm_bf16_launch(const void* A, int64_t lda, const void* B, int64_t ldb, int32_t M, int32_t N, int32_t K,
                     const float* bias, int32_t relu, void* out, int32_t out_dtype, int64_t ldo, int32_t mode,
                     int32_t* mode_used, cudaStream_t stream, const GemmFinish* fin = nullptr);
void gemm_set_trace(unsigned long long* buf);
void roi_set_trace(unsigned long long* buf);
void head_set_trace(unsigned long long* buf);
unsigned long long* head_get_trace();
void post_set_trace(unsigned long long* buf);
int bias_act_launch(const float* in, int64_t ldi, const float* bias, int32_t relu, int32_t M, int32_t N,
                    void* out, int32_t out_dtype, int64_t ldo, cudaStream_t stream, bool clear_in = false);
int cast_bf16_launch(const float* in, int64_t ld_in, int32_t rows, int32_t cols, void* out, int64_t ld_out,
                     cudaStream_t stream);
int head_decode_launch(const ovm_decode_cfg_t* cfg, const ovm_head_weights_t* w, const float* feats,
                       int64_t ld_feats, const float* boxes, const int64_t* classes, const float* scores,
                       const int32_t* batch_idx, const float* cams, int32_t num_images, int32_t n,
                       float* records, float* raw, cudaStream_t stream, const ovm_peer_records_t* peers = nullptr);
// fc_tail.cu: last FC layer + heads + decode in one launch (class-agnostic heads, shared FC stack)
// `pending`: the layer before was a split-K GEMM that only ACCUMULATED (GemmFinish.defer): its raw fp32 sums sit in
// acc [n, ld] (a workspace that is zero between calls) and A has not been written yet.  The tail then finishes that
// layer on the way in: every CTA reads the rows it is about to load, adds the bias, applies ReLU, rounds to A's format,
// writes A, and hands acc back zeroed.
struct TailPending {
  float* acc;
  int64_t ld;
  const float* bias;   // [K of the tail] = [N of the layer before]
};
bool fc_tail_supported(const ovm_decode_cfg_t* cfg, int32_t N, int32_t K);
int fc_tail_launch(const void* A, int64_t lda, int32_t a_dtype, const void* W, int64_t ldw, const float* bias, int32_t N,
                   int32_t K, const ovm_decode_cfg_t* cfg, const ovm_head_weights_t* w, const float* boxes,
                   const int64_t* classes, const float* scores, const int32_t* batch_idx, const float* cams,
                   int32_t num_images, int32_t n, float* records, float* raw, cudaStream_t stream,
                   const ovm_peer_records_t* peers = nullptr, const struct TailPending* pending = nullptr);
int cast_f16_launch(const float* in, int64_t ld_in, int32_t rows, int32_t cols, void* out, int64_t ld_out, cudaStream_t stream);
int round_tf32_launch(const float* in, int64_t ld_in, int32_t rows, int32_t cols, float* out, int64_t ld_out,
                      cudaStream_t stream);

// detect_post.cu
int nms_launch(const float* boxes, const float* scores, const int64_t* classes, const int32_t* seg_offsets,
               int32_t num_segments, int32_t max_segment_len, double iou_threshold, int32_t mode, int32_t* keep,
               int32_t* keep_count, cudaStream_t stream);
int gdino_post_launch(const float* logits, const float* boxes_cxcywh, const int32_t* phrase_ranges, int32_t batch,
                      int32_t nq, int32_t T, int32_t P, int32_t image_height, int32_t image_width,
                      float box_threshold, double nms_threshold, float* out_boxes, float* out_scores,
                      int64_t* out_classes, int32_t* out_query, int32_t* out_count, cudaStream_t stream);

int box_inference_launch(const ovm_box_inference_cfg_t* cfg, const float* logits, int64_t ld_logits,
                         const float* deltas, int64_t ld_deltas, const float* proposals, const int32_t* seg_offsets,
                         const int32_t* image_sizes, int32_t N, int32_t R, int32_t max_per_image,
                         const ovm_box_inference_ws_t* ws, float* out_boxes, float* out_scores, int64_t* out_classes,
                         int32_t* out_index, float* out_scores_full, int32_t* out_count, cudaStream_t stream);

// records.cu
int records_post_launch(const float* records, const int32_t* offsets, const float* geom, int32_t num_images, float* out,
                        int32_t* out_count, cudaStream_t stream);
int peer_store_launch(float* const* ptrs, int32_t n, float value, cudaStream_t stream);
// iou3d.cu
int iou_box3d_launch(const float* boxes_dt, const float* boxes_gt, int32_t N, int32_t M, float eps_coplanar,
                     float eps_nonzero, float* iou, float* vol, int32_t* dt_flags, cudaStream_t stream);

}  // namespace ovm
