// Kernel 1: multi-level ROIAlignV2 (level assignment + 7x7 bilinear pooling), one launch for all levels.
//
// Replaces d2 ROIPooler.forward / assign_boxes_to_levels / ROIAlign(aligned=True) and the torchvision
// roi_align kernel behind them (reference call site cubercnn/modeling/roi_heads/roi_heads.py:166-171,366;
// per-sample rule restated from torchvision/ops/roi_align.py:113-190 + the C++ kernel's range test).
//
// Design (NCHW fp32 in, (c,ph,pw)-major fp32/bf16 out):
//   * one CTA per (box, group of 4 channel chunks); a chunk = 32 channels = one warp, lane == channel.
//   * per box the bilinear taps of all samples are folded ONCE into two separable weight tables in shared
//     memory: Wx[pixel column][pw] and Wy[ph][pixel row] (sum of hx/lx resp. hy/ly over the bin's sample grid).
//     out[c,ph,pw] = (sum_y sum_x Wy[ph][y] * Wx[x][pw] * f[c,y,x]) / count  -- identical to the reference's
//     sum over samples of w1..w4 * f up to fp32 re-association.
//   * each warp streams the rows a bin-row needs ("band") for its 32 channels from global memory into its
//     private shared-memory buffer with 16-byte cp.async (footprint columns aligned down to 4 floats), laid
//     out [channel][row][col] with a channel stride of 4*odd floats, so the compute phase reads 128-bit
//     vectors with lane == channel free of bank conflicts.  No block-wide barrier after the tables are built.
//   * results are staged through shared memory and written as one contiguous run of 32*49 values per chunk.
//   * geometry the tables cannot hold (footprint wider than 124 columns, a bin taller than 64 rows), NHWC
//     input and P != 7 take the direct path: one thread per output, sampling global memory like torchvision.
#include <cfloat>
#include <climits>
#include <cmath>

#include "common.cuh"

namespace ovm {

namespace {

constexpr int kP = 7;
constexpr int kWarps = 4;
constexpr int kThreads = kWarps * 32;
constexpr int kMaxW = 124;                 // staged columns (multiple of 4)
constexpr int kMaxRows = 64;               // rows one output bin-row may touch
constexpr int kSMax = 128;                 // per-channel floats in a band buffer (4*odd <= 128 after padding)
constexpr int kBufFloats = 32 * kSMax;     // 16 KB per warp
constexpr int kWxFloats = kMaxW * 8;
constexpr int kWyFloats = kP * kMaxRows;
constexpr int kMetaInts = 64;
constexpr size_t kSmemBytes = (size_t)(kWarps * kBufFloats + kWxFloats + kWyFloats) * 4 + kMetaInts * 4;

struct RoiLevel {
  const float* data;
  int H, W;
  float scale;
  int layout;
};

struct RoiArgs {
  RoiLevel lv[OVM_MAX_LEVELS];
  float thr[OVM_MAX_LEVELS];  // s >= thr[j] (j = 1..L-1)  <=>  reference level index >= j
  int L, N, C, P, sampling_ratio;
  float canon_size, roi_scale;
  const float* boxes;
  const int* batch_idx;
  int n;
  void* out;
  int out_bf16;
  long long out_ld;
  int* level_out;
  int splits, chunks_per_split;
};

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src, int src_bytes) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

// One sample coordinate along an axis, op for op as torchvision computes it (no FMA contraction):
//   c = start + p * bin + (i + .5f) * bin / grid ; out of [-1, size] -> invalid ; clamp ; low/high ; l/h.
struct Tap {
  int lo, hi;
  float l, h;
  bool valid;
};
__device__ __forceinline__ Tap make_tap(float start, float bin, int grid, int p, int i, int size) {
  Tap t;
  float c = __fadd_rn(__fadd_rn(start, __fmul_rn((float)p, bin)),
                      __fdiv_rn(__fmul_rn((float)i + .5f, bin), (float)grid));
  t.valid = !(c < -1.0f || c > (float)size);
  if (c <= 0.f) c = 0.f;
  int lo = (int)c;
  int hi;
  if (lo >= size - 1) {
    hi = lo = size - 1;
    c = (float)lo;
  } else {
    hi = lo + 1;
  }
  t.lo = lo;
  t.hi = hi;
  t.l = __fsub_rn(c, (float)lo);
  t.h = __fsub_rn(1.f, t.l);
  return t;
}

struct BoxGeom {
  int level;  // -1: no level (NaN size)
  int b;
  float start_w, start_h, bin_w, bin_h;
  int gw, gh;
  float count;
};

__device__ __forceinline__ BoxGeom box_geometry(const RoiArgs& a, int box) {
  BoxGeom g;
  float x1 = a.boxes[4 * box + 0], y1 = a.boxes[4 * box + 1];
  float x2 = a.boxes[4 * box + 2], y2 = a.boxes[4 * box + 3];
  if (a.roi_scale > 0.f) {
    // ROIHeads3D.scale_proposals (roi_heads.py:310-327); heights use x2-x1 exactly as :317 does.
    float cx = __fdiv_rn(__fadd_rn(x1, x2), 2.f), cy = __fdiv_rn(__fadd_rn(y1, y2), 2.f);
    float w = __fsub_rn(x2, x1);
    float hw = __fmul_rn(__fmul_rn(0.5f, w), a.roi_scale);
    x1 = __fsub_rn(cx, hw);
    x2 = __fadd_rn(cx, hw);
    y1 = __fsub_rn(cy, hw);
    y2 = __fadd_rn(cy, hw);
  }
  g.b = a.batch_idx[box];
  int level = 0;
  if (a.L > 1) {
    // d2 assign_boxes_to_levels: floor(canonical_level + log2(sqrt(area)/canonical_size + 1e-8)), clamped.
    // sqrt / div / add are IEEE-exact on both sides; the log2+floor is a monotone step function of s whose
    // break points the host located with the same fp32 arithmetic the reference uses (roi_align_launch).
    float area = __fmul_rn(__fsub_rn(x2, x1), __fsub_rn(y2, y1));
    float s = __fadd_rn(__fdiv_rn(__fsqrt_rn(area), a.canon_size), 1e-8f);
    if (s != s) {
      level = -1;
    } else {
      for (int j = 1; j < a.L; ++j) level += (s >= a.thr[j]) ? 1 : 0;
    }
  }
  g.level = level;
  if (level < 0) return g;
  const float scale = a.lv[level].scale;
  g.start_w = __fsub_rn(__fmul_rn(x1, scale), 0.5f);
  g.start_h = __fsub_rn(__fmul_rn(y1, scale), 0.5f);
  float end_w = __fsub_rn(__fmul_rn(x2, scale), 0.5f);
  float end_h = __fsub_rn(__fmul_rn(y2, scale), 0.5f);
  float roi_w = __fsub_rn(end_w, g.start_w), roi_h = __fsub_rn(end_h, g.start_h);
  g.bin_w = __fdiv_rn(roi_w, (float)a.P);
  g.bin_h = __fdiv_rn(roi_h, (float)a.P);
  g.gw = a.sampling_ratio > 0 ? a.sampling_ratio : (int)ceilf(__fdiv_rn(roi_w, (float)a.P));
  g.gh = a.sampling_ratio > 0 ? a.sampling_ratio : (int)ceilf(__fdiv_rn(roi_h, (float)a.P));
  g.count = (float)max(g.gh * g.gw, 1);
  return g;
}

__device__ __forceinline__ void store_out(const RoiArgs& a, long long idx, float v) {
  if (a.out_bf16)
    reinterpret_cast<__nv_bfloat16*>(a.out)[idx] = __float2bfloat16_rn(v);
  else
    reinterpret_cast<float*>(a.out)[idx] = v;
}

// Direct path: thread per output scalar, global-memory sampling (any geometry, any P, NCHW or NHWC).
__device__ void direct_path(const RoiArgs& a, const BoxGeom& g, int box, int c_begin, int c_end) {
  const int PP = a.P * a.P;
  const int total = (c_end - c_begin) * PP;
  const long long out_row = (long long)box * a.out_ld;
  if (g.level < 0 || g.gw <= 0 || g.gh <= 0) {
    for (int o = threadIdx.x; o < total; o += blockDim.x) store_out(a, out_row + (long long)c_begin * PP + o, 0.f);
    return;
  }
  const RoiLevel lv = a.lv[g.level];
  const int H = lv.H, W = lv.W, C = a.C;
  for (int o = threadIdx.x; o < total; o += blockDim.x) {
    const int c = c_begin + o / PP;
    const int r = o % PP;
    const int ph = r / a.P, pw = r % a.P;
    float sum = 0.f;
    for (int iy = 0; iy < g.gh; ++iy) {
      Tap ty = make_tap(g.start_h, g.bin_h, g.gh, ph, iy, H);
      for (int ix = 0; ix < g.gw; ++ix) {
        Tap tx = make_tap(g.start_w, g.bin_w, g.gw, pw, ix, W);
        if (!(ty.valid && tx.valid)) continue;
        float v1, v2, v3, v4;
        if (lv.layout == OVM_LAYOUT_NCHW) {
          const float* p = lv.data + ((long long)g.b * C + c) * H * W;
          v1 = p[ty.lo * W + tx.lo];
          v2 = p[ty.lo * W + tx.hi];
          v3 = p[ty.hi * W + tx.lo];
          v4 = p[ty.hi * W + tx.hi];
        } else {
          const float* p = lv.data + (long long)g.b * H * W * C + c;
          v1 = p[((long long)ty.lo * W + tx.lo) * C];
          v2 = p[((long long)ty.lo * W + tx.hi) * C];
          v3 = p[((long long)ty.hi * W + tx.lo) * C];
          v4 = p[((long long)ty.hi * W + tx.hi) * C];
        }
        sum += ty.h * tx.h * v1 + ty.h * tx.l * v2 + ty.l * tx.h * v3 + ty.l * tx.l * v4;
      }
    }
    store_out(a, out_row + (long long)c_begin * PP + o, sum / g.count);
  }
}

// meta layout (ints): [0..6] x bin lo, [7..13] x bin hi, [14..20] y bin lo, [21..27] y bin hi (lo > hi: no valid sample)
template <bool ALIGNED16>
__global__ void __launch_bounds__(kThreads) roi_align_kernel(const RoiArgs a) {
  extern __shared__ float4 smem_f4[];
  float* smem = reinterpret_cast<float*>(smem_f4);
  float* bufs = smem;                                   // kWarps * kBufFloats
  float* Wx = bufs + kWarps * kBufFloats;               // [col][8]
  float* Wy = Wx + kWxFloats;                           // [ph][kMaxRows]
  int* meta = reinterpret_cast<int*>(Wy + kWyFloats);

  const int box = blockIdx.x / a.splits;
  const int split = blockIdx.x % a.splits;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int chunk_begin = split * a.chunks_per_split;
  const int c_begin = chunk_begin * 32;
  const int c_end = min(a.C, (chunk_begin + a.chunks_per_split) * 32);

  const BoxGeom g = box_geometry(a, box);
  if (split == 0 && tid == 0 && a.level_out) a.level_out[box] = g.level;

  bool direct = (a.P != kP) || g.level < 0 || g.gw <= 0 || g.gh <= 0;
  if (!direct) direct = a.lv[g.level].layout != OVM_LAYOUT_NCHW || (ALIGNED16 && (a.lv[g.level].W & 3));
  if (direct) {  // block-uniform
    direct_path(a, g, box, c_begin, c_end);
    return;
  }
  const RoiLevel lv = a.lv[g.level];
  const int H = lv.H, W = lv.W;

  // ---- pass A: pixel range of every bin (threads 0..6: x bins, 32..38: y bins)
  if (lane < kP && warp < 2) {
    const bool is_x = warp == 0;
    const float start = is_x ? g.start_w : g.start_h, bin = is_x ? g.bin_w : g.bin_h;
    const int grid = is_x ? g.gw : g.gh, size = is_x ? W : H;
    int mn = INT_MAX, mx = -1;
    for (int i = 0; i < grid; ++i) {
      Tap t = make_tap(start, bin, grid, lane, i, size);
      if (t.valid) {
        mn = min(mn, t.lo);
        mx = max(mx, t.hi);
      }
    }
    meta[(is_x ? 0 : 14) + lane] = mn;
    meta[(is_x ? 7 : 21) + lane] = mx;
  }
  __syncthreads();
  int x_lo = INT_MAX, x_hi = -1, ny_max = 0;
#pragma unroll
  for (int p = 0; p < kP; ++p) {
    x_lo = min(x_lo, meta[p]);
    x_hi = max(x_hi, meta[7 + p]);
    ny_max = max(ny_max, meta[21 + p] - meta[14 + p] + 1);
  }
  const long long out_row = (long long)box * a.out_ld;
  if (x_hi < 0 || ny_max <= 0) {  // every sample out of range: the reference's sum is empty
    for (int o = tid; o < (c_end - c_begin) * kP * kP; o += kThreads)
      store_out(a, out_row + (long long)c_begin * kP * kP + o, 0.f);
    return;
  }
  const int x0 = ALIGNED16 ? (x_lo & ~3) : x_lo;
  const int fwp = (x_hi - x0 + 1 + 3) & ~3;
  if (fwp > kMaxW || ny_max > kMaxRows) {
    direct_path(a, g, box, c_begin, c_end);
    return;
  }
  // ---- pass B: fold the taps into Wx / Wy
  for (int i = tid; i < fwp * 8; i += kThreads) Wx[i] = 0.f;
  for (int i = tid; i < kWyFloats; i += kThreads) Wy[i] = 0.f;
  __syncthreads();
  if (lane < kP && warp < 2) {
    const bool is_x = warp == 0;
    const float start = is_x ? g.start_w : g.start_h, bin = is_x ? g.bin_w : g.bin_h;
    const int grid = is_x ? g.gw : g.gh, size = is_x ? W : H;
    const int y_first = meta[14 + lane];
    for (int i = 0; i < grid; ++i) {
      Tap t = make_tap(start, bin, grid, lane, i, size);
      if (!t.valid) continue;
      if (is_x) {
        Wx[(t.lo - x0) * 8 + lane] += t.h;
        Wx[(t.hi - x0) * 8 + lane] += t.l;
      } else {
        Wy[lane * kMaxRows + (t.lo - y_first)] += t.h;
        Wy[lane * kMaxRows + (t.hi - y_first)] += t.l;
      }
    }
  }
  __syncthreads();

  // ---- per-warp streaming over channel chunks
  const int nx4 = fwp >> 2;
  int R = min(ny_max, kSMax / fwp);          // rows per staged group (>= 1 because fwp <= kMaxW < kSMax)
  int S4 = (R * fwp) >> 2;
  if (!(S4 & 1)) {                           // channel stride must be 4*odd floats
    if ((S4 + 1) * 4 <= kSMax) S4 += 1;
    else { R -= 1; S4 = (R * fwp) >> 2; if (!(S4 & 1)) S4 += 1; }
  }
  if (R < 1) {  // only when fwp*1 is 4*even and no room to pad: cannot happen with kMaxW=124 < 128, kept for safety
    direct_path(a, g, box, c_begin, c_end);
    return;
  }
  const int S = S4 * 4;
  float* buf = bufs + warp * kBufFloats;
  const uint32_t buf_s = smem_u32(buf);
  const float inv_div = g.count;
  const long long HW = (long long)H * W;

  for (int chunk = chunk_begin + warp; chunk < chunk_begin + a.chunks_per_split && chunk * 32 < a.C; chunk += kWarps) {
    const int c0 = chunk * 32;
    const int nch = min(32, a.C - c0);
    const float* img = lv.data + ((long long)g.b * a.C + c0) * HW;
    float acc[kP][kP];
#pragma unroll
    for (int i = 0; i < kP; ++i)
#pragma unroll
      for (int j = 0; j < kP; ++j) acc[i][j] = 0.f;

#pragma unroll
    for (int ph = 0; ph < kP; ++ph) {
      const int ys = meta[14 + ph];
      const int ny = meta[21 + ph] - ys + 1;   // <= 0: bin-row has no valid sample
      for (int r0 = 0; r0 < ny; r0 += R) {
        const int rows = min(R, ny - r0);
        __syncwarp();  // everyone finished reading the previous band
        const int items = rows * nx4;
        for (int it = lane; it < items; it += 32) {
          const int r = it / nx4, x4 = it - r * nx4;
          const float* src = img + (long long)(ys + r0 + r) * W + x0 + 4 * x4;
          const uint32_t dst = buf_s + (uint32_t)(r * fwp + 4 * x4) * 4u;
          if (ALIGNED16) {
            for (int cc = 0; cc < nch; ++cc) cp_async16(dst + (uint32_t)(cc * S) * 4u, src + cc * HW);
          } else {
            const int xb = x0 + 4 * x4;
            for (int cc = 0; cc < nch; ++cc) {
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const bool in = xb + e < W;
                cp_async4(dst + (uint32_t)(cc * S + e) * 4u, in ? (src + cc * HW + e) : img, in ? 4 : 0);
              }
            }
          }
        }
        cp_async_wait_all();
        __syncwarp();
        // compute: v[4 cols] = sum_rows Wy * f ; acc[ph][pw] += Wx[col][pw] * v
        const float* wy = Wy + ph * kMaxRows + r0;
        const float* mine = buf + lane * S;
        for (int x4 = 0; x4 < nx4; ++x4) {
          float v0 = 0.f, v1 = 0.f, v2 = 0.f, v3 = 0.f;
          for (int r = 0; r < rows; ++r) {
            const float4 f = *reinterpret_cast<const float4*>(mine + r * fwp + 4 * x4);
            const float w = wy[r];
            v0 = fmaf(w, f.x, v0);
            v1 = fmaf(w, f.y, v1);
            v2 = fmaf(w, f.z, v2);
            v3 = fmaf(w, f.w, v3);
          }
          const float vv[4] = {v0, v1, v2, v3};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float4 wa = *reinterpret_cast<const float4*>(Wx + (4 * x4 + e) * 8);
            const float4 wb = *reinterpret_cast<const float4*>(Wx + (4 * x4 + e) * 8 + 4);
            acc[ph][0] = fmaf(wa.x, vv[e], acc[ph][0]);
            acc[ph][1] = fmaf(wa.y, vv[e], acc[ph][1]);
            acc[ph][2] = fmaf(wa.z, vv[e], acc[ph][2]);
            acc[ph][3] = fmaf(wa.w, vv[e], acc[ph][3]);
            acc[ph][4] = fmaf(wb.x, vv[e], acc[ph][4]);
            acc[ph][5] = fmaf(wb.y, vv[e], acc[ph][5]);
            acc[ph][6] = fmaf(wb.z, vv[e], acc[ph][6]);
          }
        }
      }
    }
    // ---- stage the 32 x 49 results and write them as one contiguous run
    __syncwarp();
#pragma unroll
    for (int i = 0; i < kP; ++i)
#pragma unroll
      for (int j = 0; j < kP; ++j) buf[lane * (kP * kP) + i * kP + j] = acc[i][j] / inv_div;
    __syncwarp();
    const int nvals = nch * kP * kP;
    const long long o0 = out_row + (long long)c0 * kP * kP;
    if (a.out_bf16) {
      __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(a.out) + o0;  // o0 is even: c0 % 32 == 0, out_ld % 8 == 0
      for (int i = lane; i < (nvals >> 1); i += 32) {
        const float2 v = *reinterpret_cast<const float2*>(buf + 2 * i);
        *reinterpret_cast<__nv_bfloat162*>(out + 2 * i) = __floats2bfloat162_rn(v.x, v.y);
      }
      if ((nvals & 1) && lane == 0) out[nvals - 1] = __float2bfloat16_rn(buf[nvals - 1]);
    } else {
      float* out = reinterpret_cast<float*>(a.out) + o0;
      for (int i = lane; i < nvals; i += 32) out[i] = buf[i];
    }
  }
}

// Smallest fp32 s with floor(canonical_level + log2(s)) >= k, evaluated the way the reference's torch CPU
// path does (fp32 log2 correctly rounded around powers of two, then an fp32 add): checked in
// tests/test_levels.py against torch for +-64 ulp around every break point.
float level_threshold(int k, int canonical_level) {
  auto level_of = [&](float s) { return floorf((float)canonical_level + (float)log2((double)s)); };
  float s = ldexpf(1.f, k - canonical_level);
  if (level_of(s) >= (float)k) {
    for (int i = 0; i < 64; ++i) {
      float prev = nextafterf(s, 0.f);
      if (level_of(prev) >= (float)k) s = prev; else break;
    }
  } else {
    for (int i = 0; i < 64 && level_of(s) < (float)k; ++i) s = nextafterf(s, INFINITY);
  }
  return s;
}

}  // namespace

int roi_align_launch(const ovm_pooler_t* p, const float* boxes, const int32_t* batch_idx, int32_t n, void* out,
                     int32_t out_dtype, int64_t out_ld, int32_t* level_out, cudaStream_t stream) {
  OVM_REQUIRE(p != nullptr, "ovm_roi_align: pooler is NULL");
  OVM_REQUIRE(p->num_levels >= 1 && p->num_levels <= OVM_MAX_LEVELS, "ovm_roi_align: num_levels %d not in [1,%d]",
              p->num_levels, OVM_MAX_LEVELS);
  OVM_REQUIRE(p->channels >= 1 && p->num_images >= 1, "ovm_roi_align: bad channels/num_images");
  OVM_REQUIRE(p->pooler_resolution >= 1 && p->pooler_resolution <= 32, "ovm_roi_align: pooler_resolution %d",
              p->pooler_resolution);
  OVM_REQUIRE(out_dtype == OVM_F32 || out_dtype == OVM_BF16, "ovm_roi_align: out_dtype %d", out_dtype);
  const int PP = p->pooler_resolution * p->pooler_resolution;
  OVM_REQUIRE(out_ld >= (int64_t)p->channels * PP, "ovm_roi_align: out_ld %lld < C*P*P", (long long)out_ld);
  OVM_REQUIRE(out_dtype == OVM_F32 || (out_ld % 2 == 0), "ovm_roi_align: bf16 out_ld must be even");
  OVM_REQUIRE(p->canonical_box_size > 0, "ovm_roi_align: canonical_box_size must be > 0");
  if (n == 0) return OVM_OK;
  OVM_REQUIRE(n > 0 && boxes && batch_idx && out, "ovm_roi_align: NULL input with n=%d", n);

  RoiArgs a{};
  a.L = p->num_levels;
  // d2 ROIPooler.__init__: min_level = int(-log2(scales[0])), max_level = int(-log2(scales[-1])).
  const int min_level = (int)(-std::log2(p->levels[0].spatial_scale));
  const int max_level = (int)(-std::log2(p->levels[a.L - 1].spatial_scale));
  OVM_REQUIRE(a.L == max_level - min_level + 1, "[ROIPooler] Sizes of input featuremaps do not form a pyramid!");
  OVM_REQUIRE(0 <= min_level && min_level <= max_level, "[ROIPooler] bad level range %d..%d", min_level, max_level);
  bool aligned16 = true;
  for (int l = 0; l < a.L; ++l) {
    const ovm_level_t& s = p->levels[l];
    OVM_REQUIRE(s.data && s.height > 0 && s.width > 0, "ovm_roi_align: level %d is empty", l);
    OVM_REQUIRE(s.layout == OVM_LAYOUT_NCHW || s.layout == OVM_LAYOUT_NHWC, "ovm_roi_align: level %d layout", l);
    a.lv[l] = RoiLevel{s.data, s.height, s.width, (float)s.spatial_scale, s.layout};
    a.thr[l] = l == 0 ? 0.f : level_threshold(min_level + l, p->canonical_level);
    if ((reinterpret_cast<uintptr_t>(s.data) & 15) || (s.width & 3)) aligned16 = false;
  }
  a.N = p->num_images;
  a.C = p->channels;
  a.P = p->pooler_resolution;
  a.sampling_ratio = p->sampling_ratio;
  a.canon_size = (float)p->canonical_box_size;
  a.roi_scale = p->scale_roi_boxes;
  a.boxes = boxes;
  a.batch_idx = batch_idx;
  a.n = n;
  a.out = out;
  a.out_bf16 = out_dtype == OVM_BF16;
  a.out_ld = out_ld;
  a.level_out = level_out;
  const int chunks = (a.C + 31) / 32;
  a.chunks_per_split = chunks < kWarps ? chunks : kWarps;
  a.splits = (chunks + a.chunks_per_split - 1) / a.chunks_per_split;

  static bool attr_done = false;
  if (!attr_done) {
    OVM_CUDA(cudaFuncSetAttribute(roi_align_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
    OVM_CUDA(cudaFuncSetAttribute(roi_align_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
    attr_done = true;
  }
  const dim3 grid((unsigned)((long long)n * a.splits));
  if (aligned16)
    roi_align_kernel<true><<<grid, kThreads, kSmemBytes, stream>>>(a);
  else
    roi_align_kernel<false><<<grid, kThreads, kSmemBytes, stream>>>(a);
  OVM_LAUNCHED("roi_align_kernel");
  return OVM_OK;
}

}  // namespace ovm
