// Detection records -> the reference's result rows, on the device.
//
// Replaces the arithmetic of d2 `detector_postprocess` as `GeneralizedRCNN._postprocess` calls it (reference
// cubercnn/modeling/meta_arch/rcnn3d.py:113-115: `Boxes.scale` to the original resolution, `Boxes.clip`, keep
// `Boxes.nonempty()`) and the XYXY -> XYWH conversion of `Omni3DEvaluator.instances_to_coco_json`
// (cubercnn/evaluation/omni3d_evaluation.py:1218-1219), which the reference runs per image on the host after a
// device -> host copy of every field.  One CTA per image: the image's records are filtered and compacted in order
// (ballot + prefix), so the host is left with one copy of the surviving rows and `tolist()`.
#include "common.cuh"

namespace ovm {

namespace {

constexpr int kRecThreads = 256;

// offsets [N + 1]: records of image i are rows [offsets[i], offsets[i + 1]).  geom [N, 4] fp32: scale_x, scale_y,
// output width, output height.  out: same row layout, image i's surviving rows compacted from offsets[i]; columns
// 41..44 become x, y, w, h in output-resolution pixels.  out_count [N].
__global__ void __launch_bounds__(kRecThreads)
records_post_kernel(const float* __restrict__ rec, const int* __restrict__ offsets, const float* __restrict__ geom,
                    float* __restrict__ out, int* __restrict__ out_count) {
  __shared__ int warp_cnt[kRecThreads / 32];
  __shared__ int base;
  const int img = blockIdx.x;
  const int r0 = offsets[img], n = offsets[img + 1] - r0;
  const float sx = geom[4 * img + 0], sy = geom[4 * img + 1], w = geom[4 * img + 2], h = geom[4 * img + 3];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) base = 0;
  __syncthreads();
  for (int start = 0; start < n; start += kRecThreads) {
    const int k = start + threadIdx.x;
    bool keep = false;
    float x1 = 0.f, y1 = 0.f, x2 = 0.f, y2 = 0.f;
    const float* src = rec + (size_t)(r0 + k) * OVM_RECORD_WIDTH;
    if (k < n) {
      // Boxes.scale (fp32 multiply by the fp32-rounded factor), Boxes.clip = clamp(min=0, max=size): NaN stays NaN
      x1 = __fmul_rn(src[41], sx); y1 = __fmul_rn(src[42], sy);
      x2 = __fmul_rn(src[43], sx); y2 = __fmul_rn(src[44], sy);
      x1 = x1 < 0.f ? 0.f : (x1 > w ? w : x1); x2 = x2 < 0.f ? 0.f : (x2 > w ? w : x2);
      y1 = y1 < 0.f ? 0.f : (y1 > h ? h : y1); y2 = y2 < 0.f ? 0.f : (y2 > h ? h : y2);
      keep = (__fsub_rn(x2, x1) > 0.f) && (__fsub_rn(y2, y1) > 0.f);      // Boxes.nonempty(threshold=0)
    }
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) warp_cnt[warp] = __popc(m);
    __syncthreads();
    int before = base;
    for (int wv = 0; wv < warp; ++wv) before += warp_cnt[wv];
    const int pos = before + __popc(m & ((1u << lane) - 1u));
    if (keep) {
      float* dst = out + (size_t)(r0 + pos) * OVM_RECORD_WIDTH;
      // dst may alias a row of an earlier (already read) chunk when out == rec: rows only ever move towards the front
      float v[OVM_RECORD_WIDTH];
#pragma unroll
      for (int c = 0; c < OVM_RECORD_WIDTH; c += 4) {
        const float4 t = *reinterpret_cast<const float4*>(src + c);
        v[c] = t.x; v[c + 1] = t.y; v[c + 2] = t.z; v[c + 3] = t.w;
      }
      v[41] = x1; v[42] = y1; v[43] = __fsub_rn(x2, x1); v[44] = __fsub_rn(y2, y1);   // BoxMode XYXY_ABS -> XYWH_ABS
#pragma unroll
      for (int c = 0; c < OVM_RECORD_WIDTH; c += 4)
        *reinterpret_cast<float4*>(dst + c) = make_float4(v[c], v[c + 1], v[c + 2], v[c + 3]);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      int t = 0;
      for (int wv = 0; wv < kRecThreads / 32; ++wv) t += warp_cnt[wv];
      base += t;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) out_count[img] = base;
}

// one value -> up to OVM_MAX_PEERS addresses (peer mappings): a rank announces its record count to every rank in one launch
struct PeerPtrs {
  float* p[OVM_MAX_PEERS];
};
__global__ void peer_store_kernel(const PeerPtrs ptrs, int n, float value) {
  if ((int)threadIdx.x < n) {
    *ptrs.p[threadIdx.x] = value;
    __threadfence_system();
  }
}

}  // namespace

int peer_store_launch(float* const* ptrs, int32_t n, float value, cudaStream_t stream) {
  OVM_REQUIRE(n >= 0 && n <= OVM_MAX_PEERS && (n == 0 || ptrs), "ovm_peer_store_f32: n=%d", n);
  if (n == 0) return OVM_OK;
  PeerPtrs pp{};
  for (int i = 0; i < n; ++i) {
    OVM_REQUIRE(ptrs[i] && !(reinterpret_cast<uintptr_t>(ptrs[i]) & 3), "ovm_peer_store_f32: bad pointer %d", i);
    pp.p[i] = ptrs[i];
  }
  peer_store_kernel<<<1, 32, 0, stream>>>(pp, n, value);
  OVM_LAUNCHED("peer_store_kernel");
  return OVM_OK;
}

int records_post_launch(const float* records, const int32_t* offsets, const float* geom, int32_t num_images, float* out,
                        int32_t* out_count, cudaStream_t stream) {
  OVM_REQUIRE(num_images >= 0, "ovm_records_postprocess: num_images %d", num_images);
  if (num_images == 0) return OVM_OK;
  OVM_REQUIRE(records && offsets && geom && out && out_count, "ovm_records_postprocess: NULL pointer");
  OVM_REQUIRE(out != records, "ovm_records_postprocess: out must not alias records");
  OVM_REQUIRE(!((reinterpret_cast<uintptr_t>(records) | reinterpret_cast<uintptr_t>(out)) & 15),
              "ovm_records_postprocess: records / out must be 16-byte aligned");
  records_post_kernel<<<num_images, kRecThreads, 0, stream>>>(records, offsets, geom, out, out_count);
  OVM_LAUNCHED("records_post_kernel");
  return OVM_OK;
}

}  // namespace ovm
