// Kernel 3: the five linear heads of CubeHead and the whole cube decode, one warp per box.
//
// Replaces (reference paths relative to the checkout):
//   * the 5 nn.Linear heads + uncertainty clip      cubercnn/modeling/roi_heads/cube_head.py:158-164 (shared FC)
//                                                    and :167-173 (one FC stack per head)
//   * pose -> rotation matrix                        cube_head.py:176-186 with pytorch3d rotation_6d_to_matrix /
//                                                    quaternion_to_matrix / euler_angles_to_matrix("XYZ")
//   * per-box camera / scale prep                    roi_heads.py:378-408, math_util.py:581-592
//   * class / depth-bin selection                    roi_heads.py:433-477
//   * xy, dims, allocentric->egocentric, z           roi_heads.py:414-419,480-549, math_util.py:651-679,969-978
//   * back-projection, confidence, score merge       roi_heads.py:798-829
//   * the 8 corners                                  math_util.py:116-193 (get_cuboid_verts_faces)
//
// Only the rows of the head matrices that the reference's advanced indexing keeps ([fg, box_classes] and the
// nearest depth bin) are evaluated: at most 2+3+6+1+1 = 13 dot products of length FC_DIM per box whatever
// NUM_CLASSES is.  The warp splits each dot product across lanes (float4 loads, coalesced over the feature row),
// reduces with shuffles, and every lane then runs the ~300 scalar FLOP of the decode redundantly in registers;
// lane 0 writes the 48-float detection record with 128-bit stores.  No fast-math: acosf/sinf/cosf/expf are the
// accurate versions, and NaNs propagate exactly where the reference lets them.
#include <cmath>

#include "common.cuh"
#include "decode_core.cuh"

namespace ovm {

namespace {

constexpr int kWarpsPerBlock = 12;   // 148 CTAs x 12 warps take 1776 boxes in one pass


thread_local unsigned long long* g_head_trace = nullptr;

__device__ __forceinline__ void head_stamp(const DecodeArgs& a, int slot) {
  if (a.trace && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    a.trace[(size_t)blockIdx.x * 8 + slot] = t;
  }
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

constexpr int kMaxRows = kMaxHeadRows;

// All selected head rows in one pass over the feature vector(s): out[r] = <w[r], f[r]> + bias[r].
// Lane i owns float4 i, i+32, ... of every row, so the loads of one row are coalesced across the warp and the
// kMaxRows accumulation chains are independent.  With SHARED the rows all read the same feature vector: it is
// loaded once, 8 float4 per lane at a time with all loads in flight together (the loop is load-latency bound
// otherwise).  With STAGED the weight rows are the CTA's shared-memory copy (class-agnostic heads: every box uses
// the same rows).  Every lane returns every sum.
template <bool SHARED, bool STAGED>
__device__ __forceinline__ void head_rows(const float* const (&w)[kMaxRows], const float* const (&f)[kMaxRows],
                                          const float (&bias)[kMaxRows], int nrows, int K, int lane, bool vec4,
                                          float (&out)[kMaxRows]) {
  float acc[kMaxRows];
#pragma unroll
  for (int r = 0; r < kMaxRows; ++r) acc[r] = 0.f;
  auto wload = [&](int r, int i) {
    return STAGED ? reinterpret_cast<const float4*>(w[r])[i] : __ldg(reinterpret_cast<const float4*>(w[r]) + i);
  };
  if (vec4 && SHARED) {
    const int K4 = K >> 2;
    for (int base = 0; base < K4; base += 256) {
      float4 fs[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int i = base + lane + 32 * j;
        fs[j] = i < K4 ? reinterpret_cast<const float4*>(f[0])[i] : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int r = 0; r < kMaxRows; ++r) {
        if (r < nrows) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int i = base + lane + 32 * j;
            if (i < K4) {
              const float4 a = wload(r, i);
              acc[r] = fmaf(a.x, fs[j].x, acc[r]);
              acc[r] = fmaf(a.y, fs[j].y, acc[r]);
              acc[r] = fmaf(a.z, fs[j].z, acc[r]);
              acc[r] = fmaf(a.w, fs[j].w, acc[r]);
            }
          }
        }
      }
    }
  } else if (vec4) {
    for (int i = lane; i < (K >> 2); i += 32) {
#pragma unroll
      for (int r = 0; r < kMaxRows; ++r) {
        if (r < nrows) {
          const float4 a = wload(r, i);
          const float4 b = reinterpret_cast<const float4*>(f[r])[i];
          acc[r] = fmaf(a.x, b.x, acc[r]);
          acc[r] = fmaf(a.y, b.y, acc[r]);
          acc[r] = fmaf(a.z, b.z, acc[r]);
          acc[r] = fmaf(a.w, b.w, acc[r]);
        }
      }
    }
  } else {
    for (int i = lane; i < K; i += 32) {
#pragma unroll
      for (int r = 0; r < kMaxRows; ++r)
        if (r < nrows) acc[r] = fmaf(STAGED ? w[r][i] : __ldg(w[r] + i), f[r][i], acc[r]);
    }
  }
#pragma unroll
  for (int r = 0; r < kMaxRows; ++r) out[r] = warp_sum(acc[r]) + bias[r];
}

// Sums of 16 per-lane values over the warp in 16 shuffles instead of 80: every step halves the number of values a
// lane carries (lanes with the step's bit set keep the upper half), so after the 16/8/4/2 steps lane L holds the
// partial of value (L >> 1) & 15 over its lane pair, and one more exchange completes it.  out[r] is then read from
// lane 2 r.
__device__ __forceinline__ void warp_sum16(float (&v)[16], int lane, float (&out)[kMaxRows]) {
#pragma unroll
  for (int half = 8, bit = 16; half >= 1; half >>= 1, bit >>= 1) {
    const bool up = lane & bit;
#pragma unroll
    for (int i = 0; i < half; ++i) {
      const float send = up ? v[i] : v[i + half];
      const float keep = up ? v[i + half] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, bit);
    }
  }
  const float total = v[0] + __shfl_xor_sync(0xffffffffu, v[0], 1);
#pragma unroll
  for (int r = 0; r < kMaxRows; ++r) out[r] = __shfl_sync(0xffffffffu, total, 2 * r);
}

// The common case (class-agnostic heads staged in shared memory as [nrows][K] + nrows biases, one feature vector
// for all heads, K a multiple of 1024): per 1024 features a lane loads 8 float4 of the feature row (all in flight),
// then per head row 8 LDS.128 at immediate offsets from one row pointer and 32 FMAs.
__device__ __forceinline__ void head_rows_staged(const float* staged, const float* f, int nrows, int K, int lane,
                                                 float (&out)[kMaxRows]) {
  float acc[16];
#pragma unroll
  for (int r = 0; r < 16; ++r) acc[r] = 0.f;
  const int K4 = K >> 2;
  const float4* f4 = reinterpret_cast<const float4*>(f) + lane;
  const float4* w4 = reinterpret_cast<const float4*>(staged) + lane;
  for (int base = 0; base < K4; base += 256) {
    float4 fs[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) fs[j] = f4[base + 32 * j];
#pragma unroll
    for (int r = 0; r < kMaxRows; ++r) {
      if (r < nrows) {
        const float4* wp = w4 + r * K4 + base;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 a = wp[32 * j];
          acc[r] = fmaf(a.x, fs[j].x, acc[r]);
          acc[r] = fmaf(a.y, fs[j].y, acc[r]);
          acc[r] = fmaf(a.z, fs[j].z, acc[r]);
          acc[r] = fmaf(a.w, fs[j].w, acc[r]);
        }
      }
    }
  }
  warp_sum16(acc, lane, out);
  const float* bias = staged + (size_t)nrows * K;
#pragma unroll
  for (int r = 0; r < kMaxRows; ++r) out[r] += bias[r < nrows ? r : 0];
}

// One box, one warp.  STAGED: `staged` is the CTA's shared-memory copy of the head rows in the row order below.
template <bool STAGED>
__device__ __forceinline__ void decode_box(const DecodeArgs& a, const int box, const BoxIn& in, const int lane,
                                           const float* staged) {
  const ovm_decode_cfg_t& cfg = a.cfg;
  const int K = cfg.fc_dim;
  const int NC = cfg.num_classes;
  const bool per_class = cfg.dims_priors_enabled != 0;
  const int cls = in.valid ? in.cls : 0;   // out of range: row 0 (the record becomes NaN, decode_core.cuh)
  const int csel = per_class ? cls : 0;

  const float x1 = in.x1, y1 = in.y1, x2 = in.x2, y2 = in.y2;
  // roi_heads.py:414-419
  const float src_w = x2 - x1, src_h = y2 - y1;
  const float src_scale = sqrtf(src_h * src_h + src_w * src_w);

  const float* frow = a.feats + (long long)box * a.ld_feats;
  bool vec4 = (K & 3) == 0 && (a.ld_feats & 3) == 0 && ((reinterpret_cast<uintptr_t>(a.feats) & 15) == 0);
  bool shared = true;
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    vec4 = vec4 && !(cfg.feat_offset[i] & 3);   // a vector off a 16-byte boundary: scalar loads
    shared = shared && cfg.feat_offset[i] == cfg.feat_offset[0];
  }

  // depth bin: nearest priors_z_scales[class, bin] to the 2-D diagonal (roi_heads.py:436-444)
  int bin = 0;
  if (cfg.cluster_bins > 1) {
    // torch.argmin: first minimum wins, and a NaN counts as smaller than everything
    float best = fabsf(a.w.priors_z_scales[(long long)cls * cfg.cluster_bins] - src_scale);
    for (int b = 1; b < cfg.cluster_bins; ++b) {
      const float d = fabsf(a.w.priors_z_scales[(long long)cls * cfg.cluster_bins + b] - src_scale);
      if (best == best && (d < best || d != d)) {
        best = d;
        bin = b;
      }
    }
  }

  // ---- heads: only the rows [fg, box_classes] keeps (roi_heads.py:452-477), all in one pass.
  // row order: deltas 0-1, dims 2-4, z 5, uncertainty 6 (if any), pose after that
  const int pose_n = cfg.pose_type == OVM_POSE_6D ? 6 : (cfg.pose_type == OVM_POSE_QUATERNION ? 4 : 3);
  // box_z.view(n, bins, C, -1)[i, bin, class] / view(n, C, -1)[i, class] (cube_head.py:192-201)
  const int z_row = cfg.cluster_bins > 1 ? bin * NC + cls : csel;
  const float* wr[kMaxRows];
  const float* fr[kMaxRows];
  float bs[kMaxRows], hv[kMaxRows];
#pragma unroll
  for (int r = 0; r < kMaxRows; ++r) {
    wr[r] = a.w.w_deltas;
    fr[r] = frow;
    bs[r] = 0.f;
  }
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    wr[j] = a.w.w_deltas + (long long)(csel * 2 + j) * K;
    fr[j] = frow + cfg.feat_offset[0];
    bs[j] = a.w.b_deltas[csel * 2 + j];
  }
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    wr[2 + j] = a.w.w_dims + (long long)(csel * 3 + j) * K;
    fr[2 + j] = frow + cfg.feat_offset[1];
    bs[2 + j] = a.w.b_dims[csel * 3 + j];
  }
  wr[5] = a.w.w_z + (long long)z_row * K;
  fr[5] = frow + cfg.feat_offset[3];
  bs[5] = a.w.b_z[z_row];
  const int conf_rows = cfg.use_confidence ? 1 : 0;
  if (conf_rows) {
    wr[6] = a.w.w_uncert + (long long)csel * K;
    fr[6] = frow + cfg.feat_offset[4];
    bs[6] = a.w.b_uncert[csel];
  }
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    if (j < pose_n) {
      // 6 + conf_rows + j, written so that the register arrays keep compile-time indices
      const float* wp = a.w.w_pose + (long long)(csel * pose_n + j) * K;
      const float bp = a.w.b_pose[csel * pose_n + j];
      if (conf_rows) {
        wr[7 + j] = wp; fr[7 + j] = frow + cfg.feat_offset[2]; bs[7 + j] = bp;
      } else {
        wr[6 + j] = wp; fr[6 + j] = frow + cfg.feat_offset[2]; bs[6 + j] = bp;
      }
    }
  }
  const int nrows = 6 + conf_rows + pose_n;
  if (STAGED) {
#pragma unroll
    for (int r = 0; r < kMaxRows; ++r) wr[r] = staged + (r < nrows ? r : 0) * K;
  }
  head_stamp(a, 3);
  if (STAGED && shared && vec4 && !(K & 1023))
    head_rows_staged(staged, frow + cfg.feat_offset[0], nrows, K, lane, hv);
  else if (shared)
    head_rows<true, STAGED>(wr, fr, bs, nrows, K, lane, vec4, hv);
  else
    head_rows<false, STAGED>(wr, fr, bs, nrows, K, lane, vec4, hv);
  head_stamp(a, 4);
  decode_record(a, box, in, bin, hv, lane == 0);
  head_stamp(a, 5);
}

// Grid-stride over boxes, one warp per box.  STAGED (class-agnostic heads without depth bins: every box uses the
// same <= 13 rows): the CTA first copies those rows into shared memory, in decode_box's row order.
template <bool STAGED>
__global__ void __launch_bounds__(kWarpsPerBlock * 32) head_decode_kernel(const DecodeArgs a) {
  extern __shared__ float4 staged4[];
  float* staged = reinterpret_cast<float*>(staged4);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  pdl_trigger();
  head_stamp(a, 0);
  if (STAGED) {   // head weights are parameters: staged while the FC kernel before this one is still running
    const ovm_decode_cfg_t& cfg = a.cfg;
    const int K = cfg.fc_dim;
    const int pose_n = cfg.pose_type == OVM_POSE_6D ? 6 : (cfg.pose_type == OVM_POSE_QUATERNION ? 4 : 3);
    const int conf_rows = cfg.use_confidence ? 1 : 0;
    const int nrows = 6 + conf_rows + pose_n;
    for (int r = warp; r < nrows; r += kWarpsPerBlock) {
      const float* src;
      if (r < 2) src = a.w.w_deltas + (long long)r * K;
      else if (r < 5) src = a.w.w_dims + (long long)(r - 2) * K;
      else if (r == 5) src = a.w.w_z;
      else if (r == 6 && conf_rows) src = a.w.w_uncert;
      else src = a.w.w_pose + (long long)(r - 6 - conf_rows) * K;
      float* dst = staged + (long long)r * K;
      if (!(K & 3) && !(reinterpret_cast<uintptr_t>(src) & 15)) {
        for (int i = lane; i < (K >> 2); i += 32)
          reinterpret_cast<float4*>(dst)[i] = __ldg(reinterpret_cast<const float4*>(src) + i);
      } else {
        for (int i = lane; i < K; i += 32) dst[i] = __ldg(src + i);
      }
      if (lane == 0) {   // the row's bias, after the rows
        float b;
        if (r < 2) b = a.w.b_deltas[r];
        else if (r < 5) b = a.w.b_dims[r - 2];
        else if (r == 5) b = a.w.b_z[0];
        else if (r == 6 && conf_rows) b = a.w.b_uncert[0];
        else b = a.w.b_pose[r - 6 - conf_rows];
        staged[(long long)nrows * K + r] = b;
      }
    }
  }
  // boxes, classes, scores and cameras are inputs of the call (complete before its first kernel, see
  // roi_order_kernel): the first box of every warp is loaded while the FC kernel is still running
  int box = blockIdx.x * kWarpsPerBlock + warp;
  BoxIn in{};
  if (box < a.n) in = load_box_in(a, box);
  if (STAGED) __syncthreads();
  head_stamp(a, 1);
  pdl_wait();   // the FC features
  head_stamp(a, 2);
  for (bool first = true; box < a.n; box += gridDim.x * kWarpsPerBlock, first = false) {
    if (!first) in = load_box_in(a, box);
    decode_box<STAGED>(a, box, in, lane, staged);
  }
}

}  // namespace

void head_set_trace(unsigned long long* buf) { g_head_trace = buf; }
unsigned long long* head_get_trace() { return g_head_trace; }

int head_decode_launch(const ovm_decode_cfg_t* cfg, const ovm_head_weights_t* w, const float* feats, int64_t ld_feats,
                       const float* boxes, const int64_t* classes, const float* scores, const int32_t* batch_idx,
                       const float* cams, int32_t num_images, int32_t n, float* records, float* raw,
                       cudaStream_t stream, const ovm_peer_records_t* peers) {
  OVM_REQUIRE(cfg && w, "ovm_head_decode: cfg / weights are NULL");
  OVM_REQUIRE(cfg->z_type >= OVM_Z_DIRECT && cfg->z_type <= OVM_Z_CLUSTERS, "ovm_head_decode: z_type %d", cfg->z_type);
  OVM_REQUIRE(cfg->pose_type >= OVM_POSE_6D && cfg->pose_type <= OVM_POSE_EULER,
              "Cuboid pose type %d is not recognized", cfg->pose_type);
  OVM_REQUIRE(cfg->dims_priors_func == OVM_DIMS_EXP || cfg->dims_priors_func == OVM_DIMS_SIGMOID,
              "ovm_head_decode: dims_priors_func %d", cfg->dims_priors_func);
  OVM_REQUIRE(cfg->fc_dim > 0 && cfg->num_classes > 0 && cfg->cluster_bins >= 1, "ovm_head_decode: bad sizes");
  if (cfg->cluster_bins > 1 && !cfg->dims_priors_enabled) {
    // the reference itself fails here: assignments is [n, num_classes] and indexes a [n, bins, 1] tensor
    // (roi_heads.py:449)
    set_error("ovm_head_decode: CLUSTER_BINS > 1 needs DIMS_PRIORS_ENABLED (the reference raises at roi_heads.py:449)");
    return OVM_EUNSUPPORTED;
  }
  OVM_REQUIRE(cfg->cluster_bins == 1 || w->priors_z_scales, "ovm_head_decode: priors_z_scales needed for CLUSTER_BINS > 1");
  OVM_REQUIRE(cfg->z_type != OVM_Z_CLUSTERS || w->priors_z_stats, "ovm_head_decode: priors_z_stats needed for Z_TYPE clusters");
  OVM_REQUIRE(!cfg->dims_priors_enabled || w->priors_dims_per_cat, "ovm_head_decode: priors_dims_per_cat needed");
  OVM_REQUIRE(w->w_deltas && w->b_deltas && w->w_dims && w->b_dims && w->w_pose && w->b_pose && w->w_z && w->b_z,
              "ovm_head_decode: a head weight is NULL");
  OVM_REQUIRE(!cfg->use_confidence || (w->w_uncert && w->b_uncert), "ovm_head_decode: uncertainty head is NULL");
  for (int i = 0; i < 5; ++i)
    OVM_REQUIRE(cfg->feat_offset[i] >= 0 && cfg->feat_offset[i] + cfg->fc_dim <= ld_feats,
                "ovm_head_decode: feat_offset[%d]=%d + fc_dim exceeds ld_feats", i, cfg->feat_offset[i]);
  if (n == 0) return OVM_OK;
  OVM_REQUIRE(n > 0 && feats && boxes && classes && batch_idx && cams && records && num_images > 0,
              "ovm_head_decode: NULL input with n=%d", n);
  OVM_REQUIRE(!(reinterpret_cast<uintptr_t>(records) & 15) && (!raw || !(reinterpret_cast<uintptr_t>(raw) & 15)),
              "ovm_head_decode: records / raw must be 16-byte aligned");
  DecodeArgs a{};
  a.cfg = *cfg;
  a.w = *w;
  a.feats = feats;
  a.ld_feats = ld_feats;
  a.boxes = boxes;
  a.classes = reinterpret_cast<const long long*>(classes);
  a.scores = scores;
  a.batch_idx = batch_idx;
  a.cams = cams;
  a.num_images = num_images;
  a.n = n;
  a.records = records;
  a.raw = raw;
  a.trace = g_head_trace;
  a.error_flag = nullptr;
  a.num_peers = 0;
  if (peers) {
    OVM_REQUIRE(peers->num_peers >= 0 && peers->num_peers <= OVM_MAX_PEERS, "ovm_lift_forward_peers: num_peers %d", peers->num_peers);
    for (int p = 0; p < peers->num_peers; ++p) {
      OVM_REQUIRE(peers->records[p] && !(reinterpret_cast<uintptr_t>(peers->records[p]) & 15),
                  "ovm_lift_forward_peers: peer buffer %d is NULL or not 16-byte aligned", p);
      a.peer_records[p] = peers->records[p];
    }
    a.num_peers = peers->num_peers;
  }
  const int pose_n = cfg->pose_type == OVM_POSE_6D ? 6 : (cfg->pose_type == OVM_POSE_QUATERNION ? 4 : 3);
  const int stage_rows = 6 + (cfg->use_confidence ? 1 : 0) + pose_n;
  const size_t stage_bytes = ((size_t)stage_rows * cfg->fc_dim + 16) * sizeof(float);   // rows, then their biases
  const bool staged = !cfg->dims_priors_enabled && cfg->cluster_bins == 1 && stage_bytes <= 160 * 1024 &&
                      cfg->fc_dim % 4 == 0;   // K % 4: rows stay 16-byte aligned in shared memory
  int grid = (n + kWarpsPerBlock - 1) / kWarpsPerBlock;
  if (grid > kNumSMs) grid = kNumSMs;
  if (staged) {
    static SmemOptIn opt;
    int rc = opt.ensure(head_decode_kernel<true>, stage_bytes);
    if (rc != OVM_OK) return rc;
    OVM_CUDA(launch_pdl(head_decode_kernel<true>, dim3(grid), dim3(kWarpsPerBlock * 32), stage_bytes, stream, a));
  } else {
    OVM_CUDA(launch_pdl(head_decode_kernel<false>, dim3(grid), dim3(kWarpsPerBlock * 32), 0, stream, a));
  }
  OVM_LAUNCHED("head_decode_kernel");
  return OVM_OK;
}

}  // namespace ovm
