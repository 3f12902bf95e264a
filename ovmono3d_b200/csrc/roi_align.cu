// Kernel 1: multi-level ROIAlignV2 (level assignment + 7x7 bilinear pooling), one launch for all levels.
//
// Replaces d2 ROIPooler.forward / assign_boxes_to_levels / ROIAlign(aligned=True) and the torchvision
// roi_align kernel behind them (reference call site cubercnn/modeling/roi_heads/roi_heads.py:166-171,366;
// per-sample rule restated from torchvision/ops/roi_align.py:113-190 + the C++ kernel's range test).
//
// Design (NCHW fp32 in, (c,ph,pw)-major fp32/bf16 out):
//   * one CTA per (box, range of 32-channel chunks); 8 warps: warps 0..6 each own one output bin-row ph
//     (lane == channel), warp 7 is the TMA producer.
//   * per box the bilinear taps of all samples are folded ONCE into two separable weight tables in shared
//     memory: Wx[pixel column][pw] and Wy[ph][pixel row] (sum of hx/lx resp. hy/ly over the bin's sample grid).
//     out[c,ph,pw] = (sum_y sum_x Wy[ph][y] * Wx[x][pw] * f[c,y,x]) / count  -- identical to the reference's
//     sum over samples of w1..w4 * f up to fp32 re-association.
//   * the box's pixel footprint is brought from the NCHW map into shared memory by TMA
//     (cp.async.bulk.tensor.3d over [x, y, n*C+c]): one instruction per footprint row moves a
//     {Wt columns, 1 row, 32 channels} box, Wt in {8, 16, 32} floats with the matching 32/64/128-byte swizzle,
//     so that a lane reading 128-bit pieces of ITS channel's row is bank-conflict free.  Footprints wider than
//     Wt are walked in column tiles; a pipeline stage = (channel chunk, column tile), kStages deep, guarded by
//     full/empty mbarriers.  The start column is the footprint's first column rounded down to 4 floats (the
//     innermost TMA coordinate must be 16-byte aligned) and every footprint row is fetched exactly once per chunk.
//   * the seven warps read the same stage (each the rows of its own bin-row), keep 7 accumulators per lane,
//     and after the last column tile of a chunk the 32 x 49 results are staged through shared memory and
//     written as one contiguous run.
//   * geometry the stage cannot hold (rows x Wt floats per channel > kStageFloats / 32, footprints wider than
//     kMaxW), maps TMA cannot address (NHWC, width not a multiple of 4, unaligned base) and P != 7 take the
//     direct path: one thread per output, sampling global memory like torchvision.
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <tuple>

#include "common.cuh"

namespace ovm {

namespace {

constexpr int kP = 7;
constexpr int kThreads = 256;              // 7 consumer warps + 1 producer warp
constexpr int kConsumers = kP * 32;
// (stage limit of the default kernel: 24 KB = rows x Wt <= 192 floats per channel; 48 KB ring + tables + output
//  staging = 68 KB per CTA, three CTAs per SM)
constexpr int kMaxW = 128;                 // footprint columns the Wx table holds
constexpr int kMaxRows = 64;               // rows one output bin-row may touch
constexpr int kWxFloats = kMaxW * 8;
constexpr int kWyFloats = kP * kMaxRows;
constexpr int kOutFloats = 32 * kP * kP;   // one chunk of results
constexpr int kMetaInts = 72;              // 32 spare + the work item's header (kHdrInts)
constexpr int kMaxSlots = 8;               // ring slots (a slot is as large as the box's stage, not kStageFloats)
constexpr size_t roi_smem_bytes(int ring_floats, bool cl = false) {
  return 1024 + (size_t)ring_floats * 4 + (size_t)(2 * kOutFloats + (cl ? 2 : 1) * kWxFloats + kWyFloats) * 4 + kMetaInts * 4 +
         2 * kMaxSlots * 8;
}
constexpr int kWidths = 3;                 // tensor-map tile widths: Wt = 8, 16, 32
constexpr int kHeights = 5;                // tensor-map tile heights: 1, 2, 4, 8, 16 footprint rows per TMA instruction
constexpr int kMapsPerLevel = kWidths * kHeights;

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
  const int* order;             // boxes in descending cost (roi_order_kernel), or NULL: natural order
  int splits, chunks_per_split;
  int tma_ok[OVM_MAX_LEVELS];   // level has tensor maps (NCHW, 16-byte aligned base, width % 4 == 0)
  int min_slots;                // ring slots a stage geometry should leave (OVM_ROI_MINSLOTS)
  float split_factor;           // roi_order_kernel: an item may cost this many average boxes before the box is cut
  int xalign_mask;              // footprint start column is rounded down to this mask (~3 or ~7)
  int evict_first;              // feature loads carry the L2 evict_first policy (OVM_ROI_EVICT)
  int ring_floats, stage_floats;   // shared-memory ring / largest stage of the pooling kernel that will run (plan_box needs them)
  const int* box_hdr;              // [n][kHdrInts] per-box headers written by roi_plan_kernel (with `order`)
  int plan_boxes;                  // boxes (= planning warps) per CTA of roi_plan_kernel, <= kPlanBoxes
  int cl;                          // 0, or channels per lane (2 / 4) when every level is channels-last (NHWC) and TMA-addressable
  int chunk_ch;                    // channels per chunk: 32 (NCHW: lane = channel) or 32 cl (channels-last)
  unsigned long long* trace;    // optional [grid, 16] %globaltimer stamps (ovm_roi_set_trace), else NULL
};

// Everything a pooling CTA needs to know about its box before it can issue its first TMA load, computed ONCE per box by
// roi_plan_kernel (a warp per box, all SMs) instead of by every CTA that works on a piece of the box: round 1's pooling
// CTAs spent 3 of their 4.6 us prologue on this (sqrt / divisions of the level assignment, 28 sample taps for the bin
// ranges, the stage-geometry search), 2408 times per C2 batch, with two or three threads active.
constexpr int kHdrInts = 40;      // 160 bytes = 10 x int4 per work item
enum {
  H_BOX = 0, H_CHUNKS, H_LEVEL, H_IMAGE, H_FLAGS, H_WSEL, H_G, H_RMAX, H_X0, H_FW, H_YLO, H_ROWS, H_GW, H_GH, H_NYMAX, H_PAD0,
  H_START_W, H_START_H, H_BIN_W, H_BIN_H, H_COUNT, H_COST, H_PAD1, H_PAD2, H_META   // H_META .. H_META + 13: 28 x int16
};
enum { HF_DIRECT = 1, HF_EMPTY = 2 };
constexpr int kMetaNone = 0x7fff;   // int16 encoding of "no valid sample" for a bin's low end (high end: -1)

thread_local unsigned long long* g_roi_trace = nullptr;

// Development knobs (environment, read once): used by tools/roi_bench.py to compare schedules on the device.
struct RoiTune {
  int split_pct;    // OVM_ROI_SPLIT   cost of a work item, in % of the average box, above which a box is cut along the channels
  int split_cl;     // OVM_ROI_SPLIT_CL  the same for channels-last pyramids (cheaper prologue, at most 2-4 pieces per box: 300)
  int promo;        // OVM_ROI_PROMO   L2 promotion of the tensor maps: 0 none, 1 64 B, 2 128 B, 3 256 B
  int promo_cl;     // OVM_ROI_PROMO_CL  same for channels-last maps (a pixel's chunk is one 256-byte run)
  int cl;           // OVM_ROI_CL      channels per lane of the channels-last kernel: 4 (128-channel chunks), 2 (64), or
                    //                 0: channels-last pyramids take the direct path (no TMA kernel)
  int align8;       // OVM_ROI_ALIGN8  footprint starts on a 32-byte sector (8 floats) instead of 16 bytes
  int stages;       // OVM_ROI_STAGES  2: 48 KB ring (3 CTAs / SM), 3: 36 KB ring and 64 registers (4 CTAs / SM)
  int min_slots;    // OVM_ROI_MINSLOTS  prefer stage geometries (narrower tiles, row groups) that leave this many ring slots
  int evict;        // OVM_ROI_EVICT   1: the pyramid streams through the L2 as evict_first (to keep the pooled output there
                    //                 for fc1) -- measured 96 -> 126 us: boxes do reuse each other's footprint through the L2
};
const RoiTune& roi_tune() {
  static RoiTune t = [] {
    auto env = [](const char* k, int d) { const char* v = getenv(k); return v ? atoi(v) : d; };
    RoiTune r;
    r.split_pct = env("OVM_ROI_SPLIT", 400);
    r.split_cl = env("OVM_ROI_SPLIT_CL", env("OVM_ROI_SPLIT", 300));
    r.promo = env("OVM_ROI_PROMO", 2);
    r.promo_cl = env("OVM_ROI_PROMO_CL", 3);
    r.cl = env("OVM_ROI_CL", 4);
    if (r.cl != 0 && r.cl != 2 && r.cl != 4) r.cl = 4;
    r.align8 = env("OVM_ROI_ALIGN8", 0);
    r.stages = env("OVM_ROI_STAGES", 2);
    r.min_slots = env("OVM_ROI_MINSLOTS", 1);
    r.evict = env("OVM_ROI_EVICT", 0);
    return r;
  }();
  return t;
}

__device__ __forceinline__ void trace_stamp(const RoiArgs& a, int slot) {
  if (a.trace) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    a.trace[(size_t)blockIdx.x * 16 + slot] = t;
  }
}

struct RoiMaps {
  CUtensorMap m[OVM_MAX_LEVELS * kMapsPerLevel];   // [level][log2(Wt) - 3][log2(rows)]
};

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}"
        : "=r"(done)
        : "r"(bar), "r"(parity), "r"(0x989680u)   // suspend-time hint: sleep in the barrier unit, not in a spin loop
        : "memory");
  } while (!done);
}
// packed fp32 pairs (FFMA2 on sm_100): two independent fused multiply-adds per instruction
__device__ __forceinline__ uint64_t pack2(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ float lo2(uint64_t v) { return __uint_as_float((uint32_t)v); }
__device__ __forceinline__ float hi2(uint64_t v) { return __uint_as_float((uint32_t)(v >> 32)); }
__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int x, int y, int c) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst),
      "l"(map), "r"(bar), "r"(x), "r"(y), "r"(c)
      : "memory");
}
__device__ __forceinline__ void tma_load_3d_hint(uint32_t dst, const CUtensorMap* map, uint32_t bar, int x, int y, int c,
                                                 uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4, %5}], "
      "[%2], %6;" ::"r"(dst),
      "l"(map), "r"(bar), "r"(x), "r"(y), "r"(c), "l"(policy)
      : "memory");
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

__device__ __forceinline__ BoxGeom box_geometry_xy(const RoiArgs& a, float x1, float y1, float x2, float y2, int b) {
  BoxGeom g;
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
  g.b = b;
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
__device__ void direct_path(const RoiArgs& a, const BoxGeom& g, int box, int c_begin, int c_end, int tid, int nthreads) {
  const int PP = a.P * a.P;
  const int total = (c_end - c_begin) * PP;
  const long long out_row = (long long)box * a.out_ld;
  if (g.level < 0 || g.gw <= 0 || g.gh <= 0) {
    for (int o = tid; o < total; o += nthreads) store_out(a, out_row + (long long)c_begin * PP + o, 0.f);
    return;
  }
  const RoiLevel lv = a.lv[g.level];
  const int H = lv.H, W = lv.W, C = a.C;
  for (int o = tid; o < total; o += nthreads) {
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

// One (chunk, column tile) of one bin-row for one channel (lane): the vertical weighted sum over the bin-row's NY
// pixel rows (NY = 1..4 unrolled with the row weights in registers; NY = 0: any height, weights from shared memory),
// then the 7-wide horizontal one.  `slab0` is the lane's channel row in the first pixel row of the bin-row; 16-byte
// unit j of it sits at unit j ^ sw (TMA swizzle).
struct PoolAcc {
  uint64_t a01, a23, a45;
  float a6;
};
template <int NY>
__device__ __forceinline__ void consume_tile(const float* __restrict__ slab0, int slab_floats, int sw, int ncol4,
                                             const float (&wyr)[4], const float* __restrict__ wy, int ny,
                                             const float* __restrict__ wx, PoolAcc& acc) {
#pragma unroll 2
  for (int x4 = 0; x4 < ncol4; ++x4) {
    const float* p = slab0 + 4 * (x4 ^ sw);
    uint64_t v01 = 0, v23 = 0;
    if (NY > 0) {
#pragma unroll
      for (int r = 0; r < NY; ++r) {
        const float4 f = *reinterpret_cast<const float4*>(p + r * slab_floats);
        const uint64_t w2 = pack2(wyr[r], wyr[r]);
        v01 = ffma2(w2, pack2(f.x, f.y), v01);
        v23 = ffma2(w2, pack2(f.z, f.w), v23);
      }
    } else {
      for (int r = 0; r < ny; ++r) {
        const float4 f = *reinterpret_cast<const float4*>(p + r * slab_floats);
        const uint64_t w2 = pack2(wy[r], wy[r]);
        v01 = ffma2(w2, pack2(f.x, f.y), v01);
        v23 = ffma2(w2, pack2(f.z, f.w), v23);
      }
    }
    const float vv[4] = {lo2(v01), hi2(v01), lo2(v23), hi2(v23)};
    const float* w = wx + 32 * x4;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float4 wa = *reinterpret_cast<const float4*>(w + e * 8);
      const float4 wb = *reinterpret_cast<const float4*>(w + e * 8 + 4);
      const uint64_t v2 = pack2(vv[e], vv[e]);
      acc.a01 = ffma2(pack2(wa.x, wa.y), v2, acc.a01);
      acc.a23 = ffma2(pack2(wa.z, wa.w), v2, acc.a23);
      acc.a45 = ffma2(pack2(wb.x, wb.y), v2, acc.a45);
      acc.a6 = fmaf(wb.z, vv[e], acc.a6);
    }
  }
}

// Channels-last tiles ([row][column][64 channels], no swizzle): lane owns channels 2 lane, 2 lane + 1 of the chunk, so
// every FFMA2 works on the lane's channel PAIR -- one 64-bit shared-memory load per (row, column), one FFMA2 per row for
// the vertical pass and one per bin for the horizontal one (the column weights are stored duplicated, (w, w)).
template <int NY>
__device__ __forceinline__ void consume_tile_cl(const float* __restrict__ slab0, int slab_floats, int ncol,
                                                const float (&wyr)[4], const float* __restrict__ wy, int ny,
                                                const float2* __restrict__ wx2, uint64_t (&acc)[kP]) {
#pragma unroll 2
  for (int x = 0; x < ncol; ++x) {
    const float* p = slab0 + x * 64;
    uint64_t v = 0;
    if (NY > 0) {
#pragma unroll
      for (int r = 0; r < NY; ++r)
        v = ffma2(pack2(wyr[r], wyr[r]), *reinterpret_cast<const uint64_t*>(p + r * slab_floats), v);
    } else {
      for (int r = 0; r < ny; ++r) v = ffma2(pack2(wy[r], wy[r]), *reinterpret_cast<const uint64_t*>(p + r * slab_floats), v);
    }
    const ulonglong2* w = reinterpret_cast<const ulonglong2*>(wx2 + x * 8);
    const ulonglong2 w01 = w[0], w23 = w[1], w45 = w[2], w67 = w[3];
    acc[0] = ffma2(w01.x, v, acc[0]);
    acc[1] = ffma2(w01.y, v, acc[1]);
    acc[2] = ffma2(w23.x, v, acc[2]);
    acc[3] = ffma2(w23.y, v, acc[3]);
    acc[4] = ffma2(w45.x, v, acc[4]);
    acc[5] = ffma2(w45.y, v, acc[5]);
    acc[6] = ffma2(w67.x, v, acc[6]);
  }
}

// Four channels per lane ([row][column][128 channels] tiles): one 128-bit load per (row, column), the column weights are
// loaded once for two channel pairs -- 3 ny + 18 instructions per column for ny x 512 bytes (2 ny + 11 for 256 above).
template <int NY>
__device__ __forceinline__ void consume_tile_cl4(const float* __restrict__ slab0, int slab_floats, int ncol,
                                                 const float (&wyr)[4], const float* __restrict__ wy, int ny,
                                                 const float2* __restrict__ wx2, uint64_t (&acc)[kP], uint64_t (&bcc)[kP]) {
#pragma unroll 2
  for (int x = 0; x < ncol; ++x) {
    const float* p = slab0 + x * 128;
    uint64_t va = 0, vb = 0;
    if (NY > 0) {
#pragma unroll
      for (int r = 0; r < NY; ++r) {
        const ulonglong2 f = *reinterpret_cast<const ulonglong2*>(p + r * slab_floats);
        const uint64_t w2 = pack2(wyr[r], wyr[r]);
        va = ffma2(w2, f.x, va);
        vb = ffma2(w2, f.y, vb);
      }
    } else {
      for (int r = 0; r < ny; ++r) {
        const ulonglong2 f = *reinterpret_cast<const ulonglong2*>(p + r * slab_floats);
        const uint64_t w2 = pack2(wy[r], wy[r]);
        va = ffma2(w2, f.x, va);
        vb = ffma2(w2, f.y, vb);
      }
    }
    const ulonglong2* w = reinterpret_cast<const ulonglong2*>(wx2 + x * 8);
    const ulonglong2 w01 = w[0], w23 = w[1], w45 = w[2], w67 = w[3];
    acc[0] = ffma2(w01.x, va, acc[0]); bcc[0] = ffma2(w01.x, vb, bcc[0]);
    acc[1] = ffma2(w01.y, va, acc[1]); bcc[1] = ffma2(w01.y, vb, bcc[1]);
    acc[2] = ffma2(w23.x, va, acc[2]); bcc[2] = ffma2(w23.x, vb, bcc[2]);
    acc[3] = ffma2(w23.y, va, acc[3]); bcc[3] = ffma2(w23.y, vb, bcc[3]);
    acc[4] = ffma2(w45.x, va, acc[4]); bcc[4] = ffma2(w45.x, vb, bcc[4]);
    acc[5] = ffma2(w45.y, va, acc[5]); bcc[5] = ffma2(w45.y, vb, bcc[5]);
    acc[6] = ffma2(w67.x, va, acc[6]); bcc[6] = ffma2(w67.x, vb, bcc[6]);
  }
}

// One warp plans one box: level, sampling geometry, the pixel range of every bin, the stage geometry (tile width, row
// groups) of the TMA pipeline, and the footprint area the work-item planner balances by -> the 40-int header `h`
// (warp-private).  Lanes 0..27 each take one end sample of one bin; everything else is computed redundantly.
__device__ void plan_box_warp(const RoiArgs& a, float x1, float y1, float x2, float y2, int image, int box, int lane, int* h) {
  const BoxGeom g = box_geometry_xy(a, x1, y1, x2, y2, image);
  short* meta16 = reinterpret_cast<short*>(h + H_META);
  bool direct = (a.P != kP) || g.level < 0 || g.gw <= 0 || g.gh <= 0;
  if (!direct) direct = !a.tma_ok[g.level];
  int flags = 0, wsel = -1, G = kP, rmax = 0, x0 = 0, fw = 0, y_lo = 0, rows_total = 0, ny_max = 0;
  float cost = 1.f;
  if (g.level >= 0) {
    const RoiLevel lv = a.lv[g.level];
    // pixel rows x columns of the footprint, clipped to the map
    const float w = fminf(fmaxf(g.bin_w * (float)a.P, 0.f), (float)lv.W) + 2.f;
    const float hgt = fminf(fmaxf(g.bin_h * (float)a.P, 0.f), (float)lv.H) + 2.f;
    const float area = w * hgt;
    cost = (area == area) ? area : 1.f;
  }
  if (!direct) {   // warp-uniform
    const RoiLevel lv = a.lv[g.level];
    const int H = lv.H, W = lv.W;
    // ---- pixel range of every bin.  The sample coordinate is monotone in the sample index (every rounding step of
    // make_tap is), so a bin's range is [lo of its first valid sample, hi of its last valid one]; only a bin whose end
    // sample is off the map scans.
    if (lane < 4 * kP) {
      const int k = lane / kP, p = lane - k * kP;
      const bool is_x = !(k & 1), from_end = k >= 2;
      const float start = is_x ? g.start_w : g.start_h, bin = is_x ? g.bin_w : g.bin_h;
      const int grid = is_x ? g.gw : g.gh, size = is_x ? W : H;
      int v = from_end ? -1 : kMetaNone;
      if (bin >= 0.f) {
        for (int i = 0; i < grid; ++i) {
          const Tap t = make_tap(start, bin, grid, p, from_end ? grid - 1 - i : i, size);
          if (t.valid) {
            v = from_end ? t.hi : t.lo;
            break;
          }
        }
      } else {   // inverted box with a fixed sampling ratio: coordinates decrease, every sample is looked at
        for (int i = 0; i < grid; ++i) {
          const Tap t = make_tap(start, bin, grid, p, i, size);
          if (t.valid) v = from_end ? max(v, t.hi) : min(v, t.lo);
        }
      }
      meta16[(is_x ? 0 : 14) + (from_end ? 7 : 0) + p] = (short)min(v, kMetaNone);
    }
    __syncwarp();
    int x_lo = kMetaNone, x_hi = -1, yl = kMetaNone, y_hi = -1;
#pragma unroll
    for (int p = 0; p < kP; ++p) {
      x_lo = min(x_lo, (int)meta16[p]);
      x_hi = max(x_hi, (int)meta16[7 + p]);
      yl = min(yl, (int)meta16[14 + p]);
      y_hi = max(y_hi, (int)meta16[21 + p]);
      ny_max = max(ny_max, (int)meta16[21 + p] - (int)meta16[14 + p] + 1);
    }
    if (x_hi < 0 || y_hi < 0) {
      flags |= HF_EMPTY;   // every sample out of range: the reference's sum is empty
    } else {
      // ---- column tiling: the innermost TMA coordinate must be a multiple of 16 bytes (measured: anything else
      // faults), so the footprint starts at the 4-float boundary at or below its first column
      y_lo = yl;
      x0 = a.cl ? x_lo : (x_lo & a.xalign_mask);   // channels-last: the innermost TMA coordinate is the channel
      fw = x_hi - x0 + 1;
      rows_total = y_hi - y_lo + 1;
      // Row groups: a stage holds the pixel rows of G consecutive bin-rows (G = 7: the whole footprint; 4, 2, 1: tall
      // footprints are walked group by group).  Pick the largest G -- and for it the tile width with the least padding,
      // wider on ties -- whose stage fits the largest stage and leaves at least a.min_slots ring slots; if nothing gives
      // that depth, the largest G that fits at all.
      int best_pad = INT_MAX;
      auto try_cfg = [&](int gcand, int need_slots) {
        int rm = rows_total;
        if (gcand != kP) {
          rm = 0;
          for (int hh = 0; hh * gcand < kP; ++hh) {
            int lo = INT_MAX, hi = -1;
            for (int p = hh * gcand; p < min(kP, (hh + 1) * gcand); ++p) {
              if (meta16[21 + p] >= meta16[14 + p]) {
                lo = min(lo, (int)meta16[14 + p]);
                hi = max(hi, (int)meta16[21 + p]);
              }
            }
            rm = max(rm, hi < lo ? 1 : hi - lo + 1);
          }
        }
        int ws = -1, bp = INT_MAX;
#pragma unroll
        for (int k = kWidths - 1; k >= 0; --k) {
          const int wt = (a.cl == 4 ? 2 : a.cl == 2 ? 4 : 8) << k;   // columns per tile; channels-last tiles hold 64 / 128 channels per column
          const int pad = (fw + wt - 1) / wt * wt;
          const int stage = rm * (8 << k) * 32;
          if (stage <= a.stage_floats && stage * need_slots <= a.ring_floats && pad < bp) {
            bp = pad;
            ws = k;
          }
        }
        if (ws < 0) return false;
        wsel = ws; best_pad = bp; G = gcand; rmax = rm;
        return true;
      };
      const int gs[4] = {7, 4, 2, 1};
      bool ok = false;
#pragma unroll
      for (int i = 0; i < 4 && !ok; ++i) ok = try_cfg(gs[i], a.min_slots);
#pragma unroll
      for (int i = 0; i < 4 && !ok; ++i) ok = try_cfg(gs[i], 1);
      if (wsel < 0 || best_pad > kMaxW || ny_max > kMaxRows) direct = true;
    }
  }
  __syncwarp();
  if (lane == 0) {
    if (direct) flags = HF_DIRECT;
    h[H_BOX] = box; h[H_CHUNKS] = 0; h[H_LEVEL] = g.level; h[H_IMAGE] = image; h[H_FLAGS] = flags;
    h[H_WSEL] = wsel; h[H_G] = G; h[H_RMAX] = rmax; h[H_X0] = x0; h[H_FW] = fw; h[H_YLO] = y_lo; h[H_ROWS] = rows_total;
    h[H_GW] = g.level >= 0 ? g.gw : 0; h[H_GH] = g.level >= 0 ? g.gh : 0; h[H_NYMAX] = ny_max; h[H_PAD0] = 0;
    h[H_START_W] = __float_as_int(g.level >= 0 ? g.start_w : 0.f); h[H_START_H] = __float_as_int(g.level >= 0 ? g.start_h : 0.f);
    h[H_BIN_W] = __float_as_int(g.level >= 0 ? g.bin_w : 0.f); h[H_BIN_H] = __float_as_int(g.level >= 0 ? g.bin_h : 0.f);
    h[H_COUNT] = __float_as_int(g.level >= 0 ? g.count : 1.f); h[H_COST] = __float_as_int(cost);
    h[H_PAD1] = 0; h[H_PAD2] = 0;
  }
  __syncwarp();
}

__device__ void sort_items(const RoiArgs& a, int* __restrict__ ws, const int* hdr);
// planner timeline (tools/roi_trace.py --plan): trace rows 65000 + CTA, slots 0 entry, 1 after the wait, 2 boxes planned,
// 3 last CTA elected, 4 mean known, 5 histogram, 6 scan, 7 items written
__device__ __forceinline__ void plan_stamp(const RoiArgs& a, int slot) {
  if (a.trace && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    a.trace[(size_t)(65000 + blockIdx.x) * 16 + slot] = t;
  }
}

constexpr int kPlanBoxes = 32;  // most boxes (= planning warps) per CTA; the launch uses kPlanBoxesDefault (OVM_ROI_PLANBOXES)
// Measured at C2 (tools/roi_trace.py --plan: boxes planned / items written, us after entry): 32 per CTA (50 CTAs) 4.6 / 10.2,
// 16 (100 CTAs) 3.6 / 9.7, 11 (146 CTAs, one per SM) 3.3 / 9.5, 8 (200 CTAs: a second wave of 1024-thread CTAs) 2.8-6.7 / 12.3.
// Planning one box is a ~1000-instruction dependent chain (3 us even with 11 warps per SM); after it the last CTA needs
// ~6 us alone: election 1.3, costs + mean 1.8, histogram 1.1, scan 0.4, items 1.4.
constexpr int kPlanBoxesDefault = 16;
// First kernel of the lift: one warp per box, kPlanBoxes boxes per CTA, spread over the SMs.  Writes the per-box headers
// (hdr [n][kHdrInts]) and the level of every box; the last CTA to finish then sorts the work items (sort_items).
__global__ void __launch_bounds__(1024) roi_plan_kernel(const RoiArgs a, int* __restrict__ hdr, int* __restrict__ ws) {
  __shared__ __align__(16) int h[kPlanBoxes][kHdrInts];
  // Everything the caller's stream did before the call is complete and visible once this wait returns, and the trigger
  // comes after it -- so no later kernel of the chain can start before that point, and they may read the call's inputs
  // (boxes, classes, cameras, weights) ahead of their own wait.  It also covers the workspace, which the ROIAlign
  // kernel of the previous call may still be reading.
  plan_stamp(a, 0);
  pdl_wait();
  pdl_trigger();
  plan_stamp(a, 1);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int box = blockIdx.x * a.plan_boxes + warp;
  if (warp < a.plan_boxes && box < a.n) {
    const float x1 = a.boxes[4 * box + 0], y1 = a.boxes[4 * box + 1], x2 = a.boxes[4 * box + 2], y2 = a.boxes[4 * box + 3];
    const int image = a.batch_idx[box];
    plan_box_warp(a, x1, y1, x2, y2, image, box, lane, h[warp]);
    if (lane < kHdrInts / 4)
      reinterpret_cast<int4*>(hdr + (size_t)box * kHdrInts)[lane] = reinterpret_cast<const int4*>(h[warp])[lane];
    if (lane == 0 && a.level_out) a.level_out[box] = h[warp][H_LEVEL];
  }
  if (!ws) return;     // headers only
  // ---- the last CTA to finish turns the planned boxes into the sorted work items (threadfence-reduction pattern; the
  // ticket word ws[1] is zero between launches: the sorting CTA hands it back cleared)
  __shared__ int last;
  __threadfence();
  __syncthreads();
  plan_stamp(a, 2);
  if (threadIdx.x == 0) last = atomicAdd(ws + 1, 1) == (int)gridDim.x - 1;
  __syncthreads();
  if (!last) return;
  __threadfence();
  plan_stamp(a, 3);
  sort_items(a, ws, hdr);
  plan_stamp(a, 7);
}

// Work items for a grid of CTAs with very different run times.  One CTA turns the planned boxes into items
// (header, first chunk, chunks) so that no item is much more expensive than the average box -- a box whose footprint
// is k times the mean is cut into up to k pieces along the channel chunks (otherwise the largest box alone, walked
// by one CTA, sets the kernel time: measured 89 us for a 34 x 34 pixel footprint) -- and orders the items by
// cost, largest first, with a counting sort in shared memory.  ws layout: [0] item count, [1..7] unused, then
// 2 int32 per item: box, first chunk | chunks << 16.
// At most max(3, splits) * n items (see roi_align_launch).
__device__ void sort_items(const RoiArgs& a, int* __restrict__ ws, const int* hdr) {
  // one histogram per warp ([bucket][warp], rows padded to 33 words): boxes crowd into a few cost buckets, and a
  // single shared histogram serialised ~n atomics on the same word, twice
  __shared__ int hist[64 * 33];
  __shared__ int warp_tot[32];
  __shared__ float warp_sum[32];
  __shared__ float mean_area;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < 64 * 33; i += 1024) hist[i] = 0;
  __syncthreads();
  // (other CTAs wrote the headers: L2 loads)
  auto cost_of = [&](int box) { return __int_as_float(__ldcg(hdr + (size_t)box * kHdrInts + H_COST)); };
  constexpr int kKeep = 4;   // the costs of the first 4096 boxes stay in registers between the passes
  float areas[kKeep];
  float sum = 0.f;
#pragma unroll
  for (int k = 0; k < kKeep; ++k) {
    areas[k] = tid + k * 1024 < a.n ? cost_of(tid + k * 1024) : 0.f;
    sum += areas[k];
  }
  for (int box = tid + kKeep * 1024; box < a.n; box += 1024) sum += cost_of(box);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  if (lane == 0) warp_sum[warp] = sum;
  __syncthreads();
  if (warp == 0) {
    float t = warp_sum[lane];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if (lane == 0) mean_area = t / (float)a.n;
  }
  __syncthreads();
  plan_stamp(a, 4);
  const int chunks = (a.C + a.chunk_ch - 1) / a.chunk_ch;
  const float mean = mean_area * a.split_factor;
  // chunks per item: the largest power of two with area * cpi <= mean * chunks (never more than the launch's
  // chunks_per_split, which spreads a small batch over the SMs); pieces per box <= max(1, 2 area / mean)
  auto cpi_of = [&](float area) {
    int cpi = 1;
    while (cpi * 2 <= a.chunks_per_split && area * (float)(cpi * 2) <= mean * (float)chunks) cpi *= 2;
    if (cpi * 2 > a.chunks_per_split && area * (float)a.chunks_per_split <= mean * (float)chunks) cpi = a.chunks_per_split;
    if (chunks > 8 * cpi) cpi = (chunks + 7) / 8;   // at most 8 pieces per box
    return cpi;
  };
  auto slot_of = [&](float area, int cpi) {
    const float cost = area * (float)cpi / (mean_area * (float)chunks);   // in units of the average box
    return (63 - (int)fminf(cost * 8.f, 63.f)) * 33 + warp;
  };
#pragma unroll
  for (int k = 0; k < kKeep; ++k) {
    if (tid + k * 1024 < a.n) {
      const int cpi = cpi_of(areas[k]);
      atomicAdd(&hist[slot_of(areas[k], cpi)], (chunks + cpi - 1) / cpi);
    }
  }
  for (int box = tid + kKeep * 1024; box < a.n; box += 1024) {
    const float ar = cost_of(box);
    const int cpi = cpi_of(ar);
    atomicAdd(&hist[slot_of(ar, cpi)], (chunks + cpi - 1) / cpi);
  }
  __syncthreads();
  plan_stamp(a, 5);
  {   // exclusive scan of the 64 x 32 counters in (bucket, warp) order: two per thread
    const int e0 = 2 * tid, e1 = 2 * tid + 1;
    const int i0 = (e0 >> 5) * 33 + (e0 & 31), i1 = (e1 >> 5) * 33 + (e1 & 31);
    const int h0 = hist[i0], h1 = hist[i1];
    int incl = h0 + h1;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      const int own = warp_tot[lane];
      int t = own;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, t, o);
        if (lane >= o) t += u;
      }
      warp_tot[lane] = t - own;
      if (lane == 31) {
        ws[0] = t;
        ws[1] = 0;
      }
    }
    __syncthreads();
    const int before = warp_tot[warp] + incl - h0 - h1;
    hist[i0] = before;
    hist[i1] = before + h0;
  }
  __syncthreads();
  plan_stamp(a, 6);
  // an item = (box, first chunk | chunks << 16): the pooling CTA reads it, then the box's header (roi_plan_kernel).
  // (Copying the 160-byte header into every item made this one CTA move 0.6 MB: 28 us instead of 13 for the stage.)
  auto emit = [&](int box, float ar) {
    const int cpi = cpi_of(ar);
    const int pieces = (chunks + cpi - 1) / cpi;
    const int at = atomicAdd(&hist[slot_of(ar, cpi)], pieces);
    for (int j = 0; j < pieces; ++j)
      reinterpret_cast<int2*>(ws + 8)[at + j] = make_int2(box, (j * cpi) | (min(cpi, chunks - j * cpi) << 16));
  };
#pragma unroll
  for (int k = 0; k < kKeep; ++k)
    if (tid + k * 1024 < a.n) emit(tid + k * 1024, areas[k]);
  for (int box = tid + kKeep * 1024; box < a.n; box += 1024) emit(box, cost_of(box));
}

// One chunk's 32 x 49 staged results -> the box's output row (contiguous run), by the kConsumers pooling threads.
__device__ __forceinline__ void write_chunk(const RoiArgs& a, const float* ob, long long out_row, int chunk, int tid) {
  const int c0 = chunk * a.chunk_ch;
  const int nvals = min(a.chunk_ch, a.C - c0) * kP * kP;
  const long long o0 = out_row + (long long)c0 * kP * kP;
  if (a.out_bf16) {
    __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(a.out) + o0;  // o0 is even: c0 % 32 == 0, out_ld is even
    if (!(reinterpret_cast<uintptr_t>(out) & 15) && !(nvals & 7)) {
      for (int i = tid; i < (nvals >> 3); i += kConsumers) {   // 8 values -> one 16-byte store
        const float4 lo = *reinterpret_cast<const float4*>(ob + 8 * i);
        const float4 hi = *reinterpret_cast<const float4*>(ob + 8 * i + 4);
        __nv_bfloat162 b0 = __floats2bfloat162_rn(lo.x, lo.y), b1 = __floats2bfloat162_rn(lo.z, lo.w);
        __nv_bfloat162 b2 = __floats2bfloat162_rn(hi.x, hi.y), b3 = __floats2bfloat162_rn(hi.z, hi.w);
        uint4 pk;
        pk.x = *reinterpret_cast<uint32_t*>(&b0);
        pk.y = *reinterpret_cast<uint32_t*>(&b1);
        pk.z = *reinterpret_cast<uint32_t*>(&b2);
        pk.w = *reinterpret_cast<uint32_t*>(&b3);
        *reinterpret_cast<uint4*>(out + 8 * i) = pk;
      }
    } else {
      for (int i = tid; i < (nvals >> 1); i += kConsumers) {
        const float2 v = *reinterpret_cast<const float2*>(ob + 2 * i);
        *reinterpret_cast<__nv_bfloat162*>(out + 2 * i) = __floats2bfloat162_rn(v.x, v.y);
      }
      if ((nvals & 1) && tid == 0) out[nvals - 1] = __float2bfloat16_rn(ob[nvals - 1]);
    }
  } else {
    float* out = reinterpret_cast<float*>(a.out) + o0;
    if (!(reinterpret_cast<uintptr_t>(out) & 15) && !(nvals & 3)) {
      for (int i = tid; i < (nvals >> 2); i += kConsumers)
        *reinterpret_cast<float4*>(out + 4 * i) = *reinterpret_cast<const float4*>(ob + 4 * i);
    } else {
      for (int i = tid; i < nvals; i += kConsumers) out[i] = ob[i];
    }
  }
}

// meta layout (ints): [0..6] x bin lo, [7..13] x bin hi, [14..20] y bin lo, [21..27] y bin hi (lo > hi: no valid sample)
// kRingFloats: shared-memory ring of footprint stages; kStageFloats: largest stage (rows x Wt x 32 channels) the TMA
// path accepts.
// kCL: the pyramid is channels-last (NHWC): tiles are [row][column][64 channels] (tensor maps {C, W, N*H}, box {64, Wt, R},
// Wt = 4 / 8 / 16 columns, no swizzle), a chunk is 64 channels and lane owns a channel pair (consume_tile_cl).  A
// footprint pixel is then one contiguous 256-byte run: whole L2 lines, no partially used sectors, and the column range
// need not be rounded to 16 bytes.
template <int kRingFloats, int kStageFloats, int kMinBlocks, int kCL>   // kCL: 0 NCHW; 2 / 4: channels-last, channels per lane
__global__ void __launch_bounds__(kThreads, kMinBlocks) roi_align_kernel(const __grid_constant__ RoiMaps maps, const RoiArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;   // swizzled TMA slabs need 1024-byte alignment
  uint8_t* aligned = smem_raw + (base - raw);
  float* stages = reinterpret_cast<float*>(aligned);                 // kStages * kStageFloats
  float* outbuf = stages + kRingFloats;                   // 2 * kOutFloats
  constexpr int kWxF = kCL ? 2 * kWxFloats : kWxFloats;
  constexpr int kCh = kCL ? 32 * kCL : 32;                           // channels per chunk
  float* Wx = outbuf + 2 * kOutFloats;                               // [col][8] (channels-last: float2 (w, w) entries)
  float* Wy = Wx + kWxF;                                             // [ph][kMaxRows]
  int* meta = reinterpret_cast<int*>(Wy + kWyFloats);
  const uint32_t bars = smem_u32(meta + kMetaInts);                  // full[kStages], empty[kStages]

  int* hdr = meta + 32;                                              // kHdrInts ints, 16-byte aligned
  const short* meta16 = reinterpret_cast<const short*>(hdr + H_META);
  auto mt = [&](int i) { return (int)meta16[i]; };                   // bin ranges (see the layout comment above)
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  pdl_trigger();
  // shared-memory set-up that depends on nothing: the tap tables are cleared, the ring's barriers initialised
  for (int i = tid; i < (kWxF + kWyFloats) / 4; i += kThreads)
    reinterpret_cast<float4*>(Wx)[i] = make_float4(0.f, 0.f, 0.f, 0.f);   // Wy follows Wx
  if (tid == 0) {
    for (int s = 0; s < kMaxSlots; ++s) {
      mbar_init(bars + 8 * s, 1);                    // full: the producer's expect_tx arrive
      mbar_init(bars + 8 * (kMaxSlots + s), kP);     // empty: one arrive per consumer warp
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  pdl_wait();      // the work items come from roi_order_kernel
  if (tid == 0) trace_stamp(a, 0);
  if (a.order) {   // work items of roi_order_kernel; the grid is an upper bound of their number
    if ((int)blockIdx.x >= a.order[0]) return;
    if (tid < kHdrInts / 4) {
      const int2 item = reinterpret_cast<const int2*>(a.order + 8)[blockIdx.x];
      int4 v = reinterpret_cast<const int4*>(a.box_hdr + (size_t)kHdrInts * item.x)[tid];
      if (tid == 0) v.y = item.y;      // H_CHUNKS
      reinterpret_cast<int4*>(hdr)[tid] = v;
    }
  } else if (warp == 0) {   // no planner: one CTA per (box, split) in input order, its first warp plans the box
    const int bx = (int)(blockIdx.x / a.splits);
    plan_box_warp(a, a.boxes[4 * bx + 0], a.boxes[4 * bx + 1], a.boxes[4 * bx + 2], a.boxes[4 * bx + 3], a.batch_idx[bx], bx,
                  lane, hdr);
    if (lane == 0) {
      const int cb = (int)(blockIdx.x % a.splits) * a.chunks_per_split;
      hdr[H_CHUNKS] = cb | ((min((a.C + kCh - 1) / kCh, cb + a.chunks_per_split) - cb) << 16);
      if (cb == 0 && a.level_out) a.level_out[bx] = hdr[H_LEVEL];
    }
  }
  __syncthreads();
  if (tid == 0) trace_stamp(a, 1);
  const int box = hdr[H_BOX];
  const int chunk_begin = hdr[H_CHUNKS] & 0xffff, chunk_end = chunk_begin + (hdr[H_CHUNKS] >> 16);
  const int c_begin = chunk_begin * kCh;
  const int c_end = min(a.C, chunk_end * kCh);
  BoxGeom g;
  g.level = hdr[H_LEVEL];
  g.b = hdr[H_IMAGE];
  g.start_w = __int_as_float(hdr[H_START_W]); g.start_h = __int_as_float(hdr[H_START_H]);
  g.bin_w = __int_as_float(hdr[H_BIN_W]); g.bin_h = __int_as_float(hdr[H_BIN_H]);
  g.gw = hdr[H_GW]; g.gh = hdr[H_GH];
  g.count = __int_as_float(hdr[H_COUNT]);
  const int flags = hdr[H_FLAGS];
  const long long out_row = (long long)box * a.out_ld;
  if (flags & HF_DIRECT) {  // block-uniform: what TMA cannot address or a stage cannot hold
    direct_path(a, g, box, c_begin, c_end, tid, kThreads);
    return;
  }
  if (flags & HF_EMPTY) {   // every sample out of range: the reference's sum is empty
    for (int o = tid; o < (c_end - c_begin) * kP * kP; o += kThreads)
      store_out(a, out_row + (long long)c_begin * kP * kP + o, 0.f);
    return;
  }
  const RoiLevel lv = a.lv[g.level];
  const int H = lv.H, W = lv.W;
  const int wsel = hdr[H_WSEL], G = hdr[H_G], rmax = hdr[H_RMAX];
  const int x0 = hdr[H_X0], fw = hdr[H_FW], y_lo = hdr[H_YLO], rows_total = hdr[H_ROWS];
  auto group_span = [&](int Gs, int h, int& gy, int& rows) {   // pixel rows of group h (one dummy row if it has no sample)
    if (Gs == kP) {   // the common case: the whole footprint in one stage
      gy = y_lo;
      rows = rows_total;
      return;
    }
    int lo = INT_MAX, hi = -1;
    for (int p = h * Gs; p < min(kP, (h + 1) * Gs); ++p) {
      if (mt(21 + p) >= mt(14 + p)) {
        lo = min(lo, mt(14 + p));
        hi = max(hi, mt(21 + p));
      }
    }
    if (hi < lo) lo = hi = y_lo;
    gy = lo;
    rows = hi - lo + 1;
  };
  const int ngroups = (kP + G - 1) / G;
  const int Wt = (kCL == 4 ? 2 : kCL == 2 ? 4 : 8) << wsel;
  const int ntiles = ((fw + Wt - 1) / Wt * Wt) / Wt;
  const int slab_floats = 32 * (8 << wsel);      // one footprint row of one chunk: [32 channels][Wt] / [Wt][64 channels]
  // The ring is cut into slots of this box's stage size: small footprints get up to kMaxSlots stages in flight.
  const int slot_floats = rmax * slab_floats;
  const int nslots = min(kMaxSlots, kRingFloats / slot_floats);

  if (warp == kP) {
    // ===== TMA producer (one lane): starts as soon as the footprint is known, while the other warps build the tap
    // tables.  The maps view the level as {W, N*C, H}, so one instruction with box {Wt, 32, R} writes R footprint
    // rows of the chunk as R consecutive [32 channels][Wt] slabs; a stage is the binary decomposition of its row
    // count over R = 16, 8, 4, 2, 1 (each TMA instruction costs ~60 issue slots of operand marshalling).
    if (lane == 0) {
      const CUtensorMap* map = &maps.m[g.level * kMapsPerLevel + wsel * kHeights];
      int s = 0;
      uint32_t phase = 0;
      long long waited = 0;
      uint64_t policy = 0;
      if (a.evict_first) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
      for (int chunk = chunk_begin; chunk < chunk_end; ++chunk) {
        const int cc = kCL ? chunk * kCh : g.b * a.C + chunk * 32;   // channels-last: first channel; NCHW: first plane
        for (int grp = 0; grp < ngroups; ++grp) {
          int gy, rows;
          group_span(G, grp, gy, rows);
          for (int tile = 0; tile < ntiles; ++tile) {
            const uint32_t full = bars + 8 * s;
            const long long tw = a.trace ? clock64() : 0;
            mbar_wait(bars + 8 * (kMaxSlots + s), phase ^ 1u);
            if (a.trace) waited += clock64() - tw;
            mbar_expect_tx(full, (uint32_t)(rows * slab_floats * 4));
            const uint32_t dst = base + (uint32_t)(s * slot_floats) * 4u;
            const int xc = x0 + tile * Wt;
            int r = 0;
#pragma unroll
            for (int k = kHeights - 1; k >= 0; --k) {
              while (rows - r >= (1 << k)) {
                if (kCL)
                  tma_load_3d(dst + (uint32_t)(r * slab_floats) * 4u, map + k, full, cc, xc, g.b * H + gy + r);
                else if (a.evict_first)
                  tma_load_3d_hint(dst + (uint32_t)(r * slab_floats) * 4u, map + k, full, xc, cc, gy + r, policy);
                else
                  tma_load_3d(dst + (uint32_t)(r * slab_floats) * 4u, map + k, full, xc, cc, gy + r);
                r += 1 << k;
              }
            }
            if (++s == nslots) {
              s = 0;
              phase ^= 1u;
            }
          }
        }
      }
      if (a.trace) a.trace[(size_t)blockIdx.x * 16 + 14] = (unsigned long long)waited;   // clocks waiting for a free stage
    }
    return;
  }

  // ---- pass B (the seven consumer warps): fold the taps into Wx / Wy (zeroed before the first barrier, above)
  if (lane < kP && warp < 2) {
    const bool is_x = warp == 0;
    const float start = is_x ? g.start_w : g.start_h, bin = is_x ? g.bin_w : g.bin_h;
    const int grid = is_x ? g.gw : g.gh, size = is_x ? W : H;
    const int y_first = mt(14 + lane);
    for (int i = 0; i < grid; ++i) {
      Tap t = make_tap(start, bin, grid, lane, i, size);
      if (!t.valid) continue;
      if (is_x && kCL) {
        float2* w2 = reinterpret_cast<float2*>(Wx);
        float2& a0 = w2[(t.lo - x0) * 8 + lane];
        a0 = make_float2(a0.x + t.h, a0.y + t.h);
        float2& a1 = w2[(t.hi - x0) * 8 + lane];
        a1 = make_float2(a1.x + t.l, a1.y + t.l);
      } else if (is_x) {
        Wx[(t.lo - x0) * 8 + lane] += t.h;
        Wx[(t.hi - x0) * 8 + lane] += t.l;
      } else {
        Wy[lane * kMaxRows + (t.lo - y_first)] += t.h;
        Wy[lane * kMaxRows + (t.hi - y_first)] += t.l;
      }
    }
  }
  asm volatile("bar.sync 1, %0;" ::"n"(kConsumers) : "memory");
  if (tid == 0) {
    trace_stamp(a, 2);
    if (a.trace) {
      unsigned smid;
      asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
      a.trace[(size_t)blockIdx.x * 16 + 15] = ((unsigned long long)(rows_total) << 48) | ((unsigned long long)fw << 32) |
                                                ((unsigned long long)wsel << 24) | smid;
    }
  }

  // ===== consumers: warp == ph, lane == channel
  const int ph = warp;
  const int my_grp = ph / G;
  int my_gy, my_rows;
  group_span(G, my_grp, my_gy, my_rows);
  const int ys = mt(14 + ph);
  const int ny = mt(21 + ph) - ys + 1;          // <= 0: bin-row has no valid sample
  const float* wy = Wy + ph * kMaxRows;
  const float inv_count = 1.f / g.count;
  // 16-byte unit j of lane c's row sits at unit j ^ sw (the TMA swizzle of the slab, see the header comment)
  const int sw = wsel == 2 ? (lane & 7) : (wsel == 1 ? ((lane >> 1) & 3) : ((lane >> 2) & 1));
  const int nx4_tile = Wt >> 2;
  const int row_off = (ys - my_gy) * slab_floats + (kCL ? kCL * lane : lane * Wt);
  float wyr[4];   // row weights of a bin-row of up to 4 pixel rows (the table is zero past the bin-row's rows)
#pragma unroll
  for (int r = 0; r < 4; ++r) wyr[r] = wy[r];
  int s = 0;
  uint32_t phase = 0;
  bool first = true;
  long long waited = 0;
  for (int chunk = chunk_begin; chunk < chunk_end; ++chunk) {
    PoolAcc acc{0, 0, 0, 0.f};
    uint64_t acc2[kP] = {0, 0, 0, 0, 0, 0, 0};   // channels-last: bin pw of the lane's (first) channel pair
    uint64_t acc3[kP] = {0, 0, 0, 0, 0, 0, 0};   // kCL == 4: ... of its second channel pair
    for (int grp = 0; grp < ngroups; ++grp)
    for (int tile = 0; tile < ntiles; ++tile) {
      const long long tw = a.trace ? clock64() : 0;
      mbar_wait(bars + 8 * s, phase);
      if (a.trace) waited += clock64() - tw;
      if (tid == 0 && first) trace_stamp(a, 3);
      first = false;
      if (ny > 0 && grp == my_grp) {
        const float* slab0 = stages + s * slot_floats + row_off;
        const int ncol4 = min(nx4_tile, (fw - tile * Wt + 3) >> 2);   // columns past the footprint carry zero weight
        const float* wxt = Wx + tile * Wt * (kCL ? 16 : 8);
        if (kCL) {
          const int ncol = min(Wt, fw - tile * Wt);
          const float2* wx2 = reinterpret_cast<const float2*>(wxt);
          if (kCL == 4) {
            switch (ny) {
              case 1: consume_tile_cl4<1>(slab0, slab_floats, ncol, wyr, wy, ny, wx2, acc2, acc3); break;
              case 2: consume_tile_cl4<2>(slab0, slab_floats, ncol, wyr, wy, ny, wx2, acc2, acc3); break;
              case 3: consume_tile_cl4<3>(slab0, slab_floats, ncol, wyr, wy, ny, wx2, acc2, acc3); break;
              case 4: consume_tile_cl4<4>(slab0, slab_floats, ncol, wyr, wy, ny, wx2, acc2, acc3); break;
              default: consume_tile_cl4<0>(slab0, slab_floats, ncol, wyr, wy, ny, wx2, acc2, acc3); break;
            }
          } else
          switch (ny) {
            case 1: consume_tile_cl<1>(slab0, slab_floats, ncol, wyr, wy, ny, wx2, acc2); break;
            case 2: consume_tile_cl<2>(slab0, slab_floats, ncol, wyr, wy, ny, wx2, acc2); break;
            case 3: consume_tile_cl<3>(slab0, slab_floats, ncol, wyr, wy, ny, wx2, acc2); break;
            case 4: consume_tile_cl<4>(slab0, slab_floats, ncol, wyr, wy, ny, wx2, acc2); break;
            default: consume_tile_cl<0>(slab0, slab_floats, ncol, wyr, wy, ny, wx2, acc2); break;
          }
        } else
        switch (ny) {   // warp-uniform; a bin-row is rarely taller than 2-3 pixels
          case 1: consume_tile<1>(slab0, slab_floats, sw, ncol4, wyr, wy, ny, wxt, acc); break;
          case 2: consume_tile<2>(slab0, slab_floats, sw, ncol4, wyr, wy, ny, wxt, acc); break;
          case 3: consume_tile<3>(slab0, slab_floats, sw, ncol4, wyr, wy, ny, wxt, acc); break;
          case 4: consume_tile<4>(slab0, slab_floats, sw, ncol4, wyr, wy, ny, wxt, acc); break;
          default: consume_tile<0>(slab0, slab_floats, sw, ncol4, wyr, wy, ny, wxt, acc); break;
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(bars + 8 * (kMaxSlots + s));
      if (++s == nslots) {
        s = 0;
        phase ^= 1u;
      }
    }
    // ---- stage the chunk's 32 x 49 results and write them as one contiguous run
    // (channels-last: one buffer of 64 x 49 -- everybody must have written the previous chunk out before it is refilled)
    float* ob = kCL ? outbuf : outbuf + ((chunk - chunk_begin) & 1) * kOutFloats;
    if (kCL == 4) {   // 128 channels per chunk: staged as bf16 (the product's output type), or in two fp32 halves
      const int c0 = chunk * 128;
      if (a.out_bf16) {
        if (chunk > chunk_begin) asm volatile("bar.sync 2, %0;" ::"n"(kConsumers) : "memory");
        __nv_bfloat16* o16 = reinterpret_cast<__nv_bfloat16*>(ob) + 4 * lane * (kP * kP) + ph * kP;
#pragma unroll
        for (int p = 0; p < kP; ++p) {
          o16[p] = __float2bfloat16_rn(lo2(acc2[p]) * inv_count);
          o16[kP * kP + p] = __float2bfloat16_rn(hi2(acc2[p]) * inv_count);
          o16[2 * kP * kP + p] = __float2bfloat16_rn(lo2(acc3[p]) * inv_count);
          o16[3 * kP * kP + p] = __float2bfloat16_rn(hi2(acc3[p]) * inv_count);
        }
        asm volatile("bar.sync 1, %0;" ::"n"(kConsumers) : "memory");
        const int nvals = min(128, a.C - c0) * kP * kP;
        __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(a.out) + out_row + (long long)c0 * kP * kP;
        const __nv_bfloat16* src = reinterpret_cast<const __nv_bfloat16*>(ob);
        if (!(reinterpret_cast<uintptr_t>(out) & 15) && !(nvals & 7)) {
          for (int i = tid; i < (nvals >> 3); i += kConsumers)
            reinterpret_cast<uint4*>(out)[i] = reinterpret_cast<const uint4*>(src)[i];
        } else {
          for (int i = tid; i < nvals; i += kConsumers) out[i] = src[i];
        }
      } else {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          if (chunk > chunk_begin || h) asm volatile("bar.sync 2, %0;" ::"n"(kConsumers) : "memory");
          float* o = ob + 2 * lane * (kP * kP) + ph * kP;
#pragma unroll
          for (int p = 0; p < kP; ++p) {
            const uint64_t v = h ? acc3[p] : acc2[p];
            o[p] = lo2(v) * inv_count;
            o[kP * kP + p] = hi2(v) * inv_count;
          }
          asm volatile("bar.sync 1, %0;" ::"n"(kConsumers) : "memory");
          float* out = reinterpret_cast<float*>(a.out) + out_row;
          for (int i = tid; i < 64 * kP * kP; i += kConsumers) {   // staged channel lc -> channel 4 (lc / 2) + 2 h + (lc & 1)
            const int lc = i / (kP * kP), r = i - lc * (kP * kP);
            const int c = c0 + (lc >> 1) * 4 + 2 * h + (lc & 1);
            if (c < a.C) out[(long long)c * (kP * kP) + r] = ob[i];
          }
        }
      }
      continue;
    }
    if (kCL) {
      if (chunk > chunk_begin) asm volatile("bar.sync 2, %0;" ::"n"(kConsumers) : "memory");
      float* o = ob + 2 * lane * (kP * kP) + ph * kP;
#pragma unroll
      for (int p = 0; p < kP; ++p) {
        o[p] = lo2(acc2[p]) * inv_count;
        o[kP * kP + p] = hi2(acc2[p]) * inv_count;
      }
    } else {
      float* o = ob + lane * (kP * kP) + ph * kP;
      o[0] = lo2(acc.a01) * inv_count; o[1] = hi2(acc.a01) * inv_count;
      o[2] = lo2(acc.a23) * inv_count; o[3] = hi2(acc.a23) * inv_count;
      o[4] = lo2(acc.a45) * inv_count; o[5] = hi2(acc.a45) * inv_count;
      o[6] = acc.a6 * inv_count;
    }
    asm volatile("bar.sync 1, %0;" ::"n"(kConsumers) : "memory");
    if (tid == 0 && chunk - chunk_begin < 8) trace_stamp(a, 4 + chunk - chunk_begin);
    write_chunk(a, ob, out_row, chunk, tid);
  }
  if (tid == 0) {
    trace_stamp(a, 12);
    if (a.trace) a.trace[(size_t)blockIdx.x * 16 + 13] = (unsigned long long)waited;   // warp 0: clocks waiting for data
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

// Tensor maps of one pyramid level: fp32 [N*C, H, W] seen as 3-D {W, N*C, H} (the row dimension outermost, so
// that a box {Wt, 32, R} lands in shared memory as R consecutive [32 channels][Wt] slabs), Wt = 8 / 16 / 32 with
// the 32 / 64 / 128-byte swizzle, R = 1, 2, 4, 8, 16.  Out-of-range elements (columns past W, rows past H, channels
// past N*C) read as zero.
// Channels-last (cl): fp32 [N, H, W, C] seen as {C, W, N*H}, box {64, Wt, R} with Wt = 4 / 8 / 16 columns, no swizzle
// (a lane reads its channel pair, 32 lanes x 8 bytes are consecutive), 256-byte L2 promotion (a pixel's 64 channels).
// NC is then N (images) and C is passed separately.
int encode_level_maps(const float* data, int H, int W, long long NC, int cl_channels, int cl_vec, CUtensorMap* out) {
  EncodeTiledFn enc = tensormap_encode_fn();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled not available from the driver");
    return OVM_ECUDA;
  }
  static const CUtensorMapSwizzle swz[kWidths] = {CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_SWIZZLE_64B,
                                                 CU_TENSOR_MAP_SWIZZLE_128B};
  static const CUtensorMapL2promotion promo[4] = {CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_64B,
                                                  CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B};
  for (int k = 0; k < kWidths; ++k) {
    for (int j = 0; j < kHeights; ++j) {
      cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)NC, (cuuint64_t)H};
      cuuint64_t strides[2] = {(cuuint64_t)H * W * 4, (cuuint64_t)W * 4};
      cuuint32_t box[3] = {(cuuint32_t)(8 << k), 32, (cuuint32_t)(1 << j)};
      cuuint32_t estr[3] = {1, 1, 1};
      CUtensorMapSwizzle sw = swz[k];
      CUtensorMapL2promotion pr = promo[roi_tune().promo & 3];
      if (cl_channels) {
        const cuuint64_t C = (cuuint64_t)cl_channels;
        dims[0] = C; dims[1] = (cuuint64_t)W; dims[2] = (cuuint64_t)NC * H;
        strides[0] = C * 4; strides[1] = (cuuint64_t)W * C * 4;
        box[0] = (cuuint32_t)(32 * cl_vec); box[1] = (cuuint32_t)((cl_vec == 4 ? 2 : 4) << k);
        sw = CU_TENSOR_MAP_SWIZZLE_NONE;
        pr = promo[roi_tune().promo_cl & 3];
      }
      CUresult r = enc(&out[k * kHeights + j], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(data), dims,
                       strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw, pr, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled(feature map %dx%d, %lld planes, box %dx32x%d) failed with CUresult %d", H, W,
                  NC, 8 << k, 1 << j, (int)r);
        return OVM_ECUDA;
      }
    }
  }
  return OVM_OK;
}

// The maps depend only on (pointer, H, W, N*C): keep the last few so that a steady-state caller re-encodes nothing.
struct MapKey {
  const void* p;
  int H, W;
  long long NC;
  int cl_channels, cl_vec;
  bool operator<(const MapKey& o) const {
    return std::tie(p, H, W, NC, cl_channels, cl_vec) < std::tie(o.p, o.H, o.W, o.NC, o.cl_channels, o.cl_vec);
  }
};
struct MapVal {
  CUtensorMap m[kMapsPerLevel];
};
int level_maps(const float* data, int H, int W, long long NC, int cl_channels, int cl_vec, CUtensorMap* out) {
  static std::mutex mu;
  static std::map<MapKey, MapVal> cache;
  std::lock_guard<std::mutex> lock(mu);
  const MapKey key{data, H, W, NC, cl_channels, cl_vec};
  auto it = cache.find(key);
  if (it == cache.end()) {
    MapVal v;
    int rc = encode_level_maps(data, H, W, NC, cl_channels, cl_vec, v.m);
    if (rc != OVM_OK) return rc;
    if (cache.size() >= 256) cache.clear();
    it = cache.emplace(key, v).first;
  }
  memcpy(out, it->second.m, sizeof(it->second.m));
  return OVM_OK;
}

}  // namespace

void roi_set_trace(unsigned long long* buf) { g_roi_trace = buf; }

int roi_align_launch(const ovm_pooler_t* p, const float* boxes, const int32_t* batch_idx, int32_t n, void* out,
                     int32_t out_dtype, int64_t out_ld, int32_t* level_out, cudaStream_t stream, int* stage_idx) {
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
  RoiMaps maps;
  memset(&maps, 0, sizeof(maps));
  const RoiTune& tune = roi_tune();
  // channels-last fast path: every level NHWC, 16-byte aligned, C % 4 == 0 (the tensor maps' strides), P == 7
  bool all_cl = tune.cl && p->pooler_resolution == kP && !(p->channels & 3);
  for (int l = 0; l < a.L && all_cl; ++l)
    all_cl = p->levels[l].layout == OVM_LAYOUT_NHWC && p->levels[l].data && !(reinterpret_cast<uintptr_t>(p->levels[l].data) & 15);
  const int cl_vec = !all_cl ? 0 : (tune.cl == 4 && p->channels >= 128 ? 4 : 2);   // narrow pyramids: 64-channel chunks
  a.cl = cl_vec;
  a.chunk_ch = all_cl ? 32 * cl_vec : 32;
  for (int l = 0; l < a.L; ++l) {
    const ovm_level_t& s = p->levels[l];
    OVM_REQUIRE(s.data && s.height > 0 && s.width > 0, "ovm_roi_align: level %d is empty", l);
    OVM_REQUIRE(s.layout == OVM_LAYOUT_NCHW || s.layout == OVM_LAYOUT_NHWC, "ovm_roi_align: level %d layout", l);
    a.lv[l] = RoiLevel{s.data, s.height, s.width, (float)s.spatial_scale, s.layout};
    a.thr[l] = l == 0 ? 0.f : level_threshold(min_level + l, p->canonical_level);
    a.tma_ok[l] = all_cl || (s.layout == OVM_LAYOUT_NCHW && !(reinterpret_cast<uintptr_t>(s.data) & 15) && !(s.width & 3) &&
                             p->pooler_resolution == kP);
    if (a.tma_ok[l]) {
      int rc = all_cl ? level_maps(s.data, s.height, s.width, p->num_images, p->channels, cl_vec, &maps.m[l * kMapsPerLevel])
                      : level_maps(s.data, s.height, s.width, (long long)p->num_images * p->channels, 0, 0,
                                   &maps.m[l * kMapsPerLevel]);
      if (rc != OVM_OK) return rc;
    }
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
  a.order = p->order_ws;
  a.trace = g_roi_trace;
  // enough CTAs for two waves of two CTAs per SM: split the channel chunks of a box over several CTAs when n is small
  const int chunks = (a.C + a.chunk_ch - 1) / a.chunk_ch;
  int splits = (4 * kNumSMs + n - 1) / n;
  splits = splits < 1 ? 1 : (splits > chunks ? chunks : splits);
  a.chunks_per_split = (chunks + splits - 1) / splits;
  a.splits = (chunks + a.chunks_per_split - 1) / a.chunks_per_split;

  const int split_pct = all_cl ? tune.split_cl : tune.split_pct;
  a.split_factor = 0.01f * (float)(split_pct < 100 ? 100 : split_pct);   // < 1 would break the item bound
  // with work items: a box yields at most max(splits, 2 area / (split_factor mean area)) of them
  const long long nitems = (long long)n * a.splits + (a.order ? (long long)(2.0 * n / a.split_factor) + 8 : 0);
  // workspace: [0] item count, [1..7] unused, nitems 8-byte work items, then the n per-box headers
  const long long ws_items = (8 + 2 * nitems + 3) / 4 * 4;       // count + 8-byte items, rounded to 16 bytes
  const long long ws_need = ws_items + (long long)kHdrInts * n;
  OVM_REQUIRE(!a.order || (ws_need <= (long long)p->order_ws_len && !(reinterpret_cast<uintptr_t>(a.order) & 15)),
              "ovm_roi_align: order_ws holds %d int32, %lld needed (OVM_ORDER_WS_LEN), 16-byte aligned", p->order_ws_len,
              ws_need);
  a.xalign_mask = tune.align8 ? ~7 : ~3;
  a.evict_first = tune.evict;
  {
    static const int pb = [] { const char* v = getenv("OVM_ROI_PLANBOXES"); int x = v ? atoi(v) : 0; return x >= 1 && x <= kPlanBoxes ? x : 0; }();
    a.plan_boxes = pb ? pb : kPlanBoxesDefault;
  }
  static SmemOptIn optA, optB, optC, optD;
  auto kC = roi_align_kernel<12288, 6144, 0, 2>;   // channels-last pyramid, 64-channel chunks
  auto kD = roi_align_kernel<12288, 6144, 3, 4>;   // channels-last pyramid, 128-channel chunks
  auto kA = roi_align_kernel<12288, 6144, 0, 0>;   // 48 KB ring, stages up to 24 KB: 69 KB of shared memory, 3 CTAs / SM
  auto kB = roi_align_kernel<9216, 4608, 4, 0>;    // 36 KB ring, stages up to 18 KB, 64 registers: 4 CTAs / SM -- measured 125 vs 104 us (OVM_ROI_STAGES=3, tools/roi_bench.py)
  {
    const bool small = tune.stages == 3 && !all_cl;
    int rc = cl_vec == 4 ? optD.ensure(kD, roi_smem_bytes(12288, true))
           : all_cl ? optC.ensure(kC, roi_smem_bytes(12288, true))
                    : (small ? optB.ensure(kB, roi_smem_bytes(9216)) : optA.ensure(kA, roi_smem_bytes(12288)));
    if (rc != OVM_OK) return rc;
    a.ring_floats = small ? 9216 : 12288;
    a.stage_floats = small ? 4608 : 6144;
  }
  if (a.order) {
    int* box_hdr = p->order_ws + ws_items;
    a.box_hdr = box_hdr;
    OVM_CUDA(launch_pdl(roi_plan_kernel, dim3((unsigned)((n + a.plan_boxes - 1) / a.plan_boxes)), dim3(1024), 0, stream, a, box_hdr, p->order_ws));
    OVM_LAUNCHED("roi_plan_kernel");
  }
  if (stage_idx) {
    int rc = stage_mark(*stage_idx, stream);
    if (rc != OVM_OK) return rc;
  }
  const dim3 grid((unsigned)nitems);
  if (all_cl) {
    if (cl_vec == 4)
      OVM_CUDA(launch_pdl(kD, grid, dim3(kThreads), roi_smem_bytes(12288, true), stream, maps, a));
    else
      OVM_CUDA(launch_pdl(kC, grid, dim3(kThreads), roi_smem_bytes(12288, true), stream, maps, a));
    OVM_LAUNCHED("roi_align_kernel");
    return OVM_OK;
  }
  if (tune.stages == 3)
    OVM_CUDA(launch_pdl(kB, grid, dim3(kThreads), roi_smem_bytes(9216), stream, maps, a));
  else
    OVM_CUDA(launch_pdl(kA, grid, dim3(kThreads), roi_smem_bytes(12288), stream, maps, a));
  OVM_LAUNCHED("roi_align_kernel");
  return OVM_OK;
}

}  // namespace ovm

// ===== ovm_roi_align_static: the same launch with the planner's scratch in a file-scope device array =====================
// For callers that cannot allocate device memory (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc):
// no counterpart in the reference.  The array is part of the library's
// image: every GPU that loads the library holds OVM_ROI_STATIC_WS_BYTES of it until the process exits.
namespace ovm {
namespace {

constexpr long long kStaticWsInts = OVM_ORDER_WS_LEN(OVM_ROI_STATIC_MAX_BOXES);
static_assert(kStaticWsInts * 4 == OVM_ROI_STATIC_WS_BYTES, "OVM_ROI_STATIC_WS_BYTES must match OVM_ORDER_WS_LEN(max boxes)");
static __device__ __align__(256) int g_roi_static_ws[kStaticWsInts];

// In host code `g_roi_static_ws` is a host-side stand-in for the symbol, not a device address: the address a kernel can
// use comes from cudaGetSymbolAddress, asked once per device with that device current.
int roi_static_ws(int** out) {
  constexpr int kMaxDevices = 64;
  static std::mutex mu;
  static int* ptrs[kMaxDevices] = {};
  int dev = -1;
  OVM_CUDA(cudaGetDevice(&dev));
  OVM_REQUIRE(dev >= 0 && dev < kMaxDevices, "ovm_roi_align_static: device index %d not in [0,%d)", dev, kMaxDevices);
  std::lock_guard<std::mutex> lock(mu);
  if (!ptrs[dev]) {
    void* p = nullptr;
    OVM_CUDA(cudaGetSymbolAddress(&p, g_roi_static_ws));
    OVM_REQUIRE(p != nullptr && !(reinterpret_cast<uintptr_t>(p) & 255), "ovm_roi_align_static: bad scratch address %p", p);
    ptrs[dev] = static_cast<int*>(p);
  }
  *out = ptrs[dev];
  return OVM_OK;
}

}  // namespace
}  // namespace ovm

#pragma GCC visibility push(default)
extern "C" int ovm_roi_align_static(const ovm_pooler_t* pooler, const float* boxes, const int32_t* batch_idx, int32_t n,
                                    void* out, int32_t out_dtype, int64_t out_ld, int32_t* level_out, void* stream) {
  using namespace ovm;
  OVM_REQUIRE(pooler != nullptr, "ovm_roi_align_static: pooler is NULL");
  OVM_REQUIRE(n >= 0 && n <= OVM_ROI_STATIC_MAX_BOXES,
              "ovm_roi_align_static: n = %d boxes, the built-in scratch holds at most %d (OVM_ROI_STATIC_MAX_BOXES); "
              "use ovm_roi_align with a caller-allocated order_ws for larger batches", n, OVM_ROI_STATIC_MAX_BOXES);
  if (n == 0) return OVM_OK;
  int* ws = nullptr;
  int rc = roi_static_ws(&ws);
  if (rc != OVM_OK) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // The array keeps the previous call's contents.  Words 0..7 (item count, ticket counter) must start at zero: a call
  // hands the ticket back cleared, but one that was aborted would not have.  Items and headers are rewritten by the
  // planner before the pooling kernel reads them.
  OVM_CUDA(cudaMemsetAsync(ws, 0, 8 * sizeof(int), st));
  ovm_pooler_t p = *pooler;      // the caller's order_ws (if any) is ignored
  p.order_ws = ws;
  p.order_ws_len = (int32_t)kStaticWsInts;
  return roi_align_launch(&p, boxes, batch_idx, n, out, out_dtype, out_ld, level_out, st, nullptr);
}
#pragma GCC visibility pop
