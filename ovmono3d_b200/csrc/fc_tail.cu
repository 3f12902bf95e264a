// fc_tail_kernel: the last FC layer of CubeHead, its five linear heads and the whole cube decode in ONE launch.
//
//   h2 = ReLU(h1 . W2^T + b2)                reference cubercnn/modeling/roi_heads/cube_head.py:64-74,157 (last Linear)
//   heads = h2 . Wh^T + bh  (<= 13 rows)     cube_head.py:158-164
//   record = decode(heads, box, camera)      roi_heads.py:378-549,798-843 (decode_core.cuh)
//
// h2 never leaves the SM: the unfused sequence wrote it as fp32 [n, 1024] (6.5 MB at C2) from one kernel and read it
// back in the next, with a launch and a serial dependency in between (34 us = 18 % of the C2 step in round 1).
//
// Decomposition: a cluster of CS = FC_DIM / 128 CTAs owns 128 boxes.  CTA r of the cluster computes the 128 x 128
// block h2[:, 128 r : 128 r + 128] with tcgen05.mma (operands by TMA, 128-byte swizzle, fp32 accumulator in TMEM),
// applies bias + ReLU on the way out of TMEM (thread = box row, tcgen05.ld), and multiplies the block straight into
// its slice of the head rows (staged in shared memory, broadcast reads): 13 partial sums per box and CTA.  The
// partials are exchanged through distributed shared memory -- box row i belongs to CTA i / (128 / CS), every CTA
// stores its 13 partials of that row into the owner's buffer with st.shared::cluster -- one cluster barrier, then
// each CTA sums the 2 CS contributions of its 128 / CS boxes in a fixed order (deterministic) and decodes them, one
// thread per box.
//
// Precision: operands are bf16, fp16 (both kind::f16) or TF32 (kind::tf32: fp32 in shared memory, pre-rounded to 10
// mantissa bits).  fp16 is the default of the lift: it carries TF32's 10 mantissa bits in bf16's two bytes -- post-ReLU
// activations and Xavier-scale weights sit well inside its range (the producer saturates at +-65504 instead of making
// infinities) -- so fc2 gets TF32's accuracy (profiles/r02_bf16_attribution.json: the two are indistinguishable, and
// remove the fc2 half of the bf16 head's pose / corner error) at bf16's operand traffic: 4.6 us of main loop at C2
// instead of TF32's 9.2 (the layer is bound by operand bytes per SM, not by the tensor pipe).
#include <cuda.h>

#include <cstdlib>

#include "common.cuh"
#include "decode_core.cuh"

namespace ovm {

namespace {

constexpr int kTM = 128;               // boxes per cluster (UMMA M)
constexpr int kTN = 128;               // fc2 output columns per CTA (UMMA N)
constexpr int kRowBytes = 128;         // K bytes per operand row and stage = one 128-byte swizzle row
constexpr int kTailStages = 5;
constexpr int kTailABytes = kTM * kRowBytes;
constexpr int kTailBBytes = kTN * kRowBytes;
constexpr int kTailStageBytes = kTailABytes + kTailBBytes;
constexpr int kTailEpiWarps = 8;       // two per TMEM lane quadrant, each takes 64 of the CTA's 128 columns
constexpr int kTailThreads = (2 + kTailEpiWarps) * 32;
constexpr int kHwStride = 16;          // head rows per column in shared memory (13 used)
constexpr int kMaxCS = 8;
constexpr int kConvGroups = 4;         // column groups of the deferred finish of the layer before (TailPending)
constexpr int kPartFloats = 2 * kMaxCS * kMaxHeadRows * (kTM / kMaxCS);   // [2 CS contributors][13][128 / CS rows]
constexpr int kXchStride = 20;         // floats one box's two decoding threads exchange (M 9, has_m, X Y z u v score, 3 raw)
constexpr size_t kTailSmemBytes = 1024 + (size_t)kTailStages * kTailStageBytes + 128 /* barriers, tmem slot */ +
                                  (size_t)(kTN * kHwStride + kTN + 16 + kPartFloats + kTM * kXchStride) * 4;

struct TailArgs {
  DecodeArgs d;          // feats / ld_feats unused
  int K;                 // input features of the layer
  int N;                 // FC_DIM = CS * 128
  int cs;                // cluster size
  int kblocks;           // ceil(K * sizeof(element) / 128)
  int elems_per_row;     // K elements per 128-byte row: 32 (tf32) or 64 (bf16)
  const float* bias;     // [N]
  // the layer before only accumulated (TailPending): raw sums in acc [n, acc_ld]; this kernel writes a_out (the tensor
  // tmA describes, pitch a_ld elements) = round(relu(acc + bias_prev)) for the rows it loads and zeroes acc
  float* acc;
  long long acc_ld;
  const float* bias_prev;
  void* a_out;
  long long a_ld;
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
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
  } while (!done);
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
      "l"(map), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tcgen05_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// the same arrive on the barrier at this offset in every CTA of `mask`
__device__ __forceinline__ void tcgen05_commit_mc(uint32_t bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
               "h"(mask)
               : "memory");
}
// TMA load delivered to the same shared-memory offset (and counted on the barrier at the same offset) in every CTA of `mask`
__device__ __forceinline__ void tma_load_2d_mc(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(dst),
      "l"(map), "r"(bar), "r"(c0), "r"(c1), "h"(mask)
      : "memory");
}
__device__ __forceinline__ void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
enum { kFmtF16 = 0, kFmtBF16 = 1, kFmtTF32 = 2 };   // UMMA a_format / b_format codes of kind::f16 / kind::tf32
template <int FMT>
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  if (FMT == kFmtTF32) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  } else {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  }
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
// K-major operand, 128-byte swizzle, 1024 bytes between 8-row groups (see gemm_tcgen05.cu)
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t mapa_shared(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void st_cluster_f32(uint32_t addr, float v) {
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// Row r of the class-agnostic head stack in decode order (deltas 0-1, dims 2-4, z 5, uncertainty 6 if any, pose after)
__device__ __forceinline__ const float* agnostic_head_row(const ovm_head_weights_t& w, int r, int conf_rows, int K,
                                                          float* bias) {   // bias: NULL = not wanted (no load)
  const float* b;
  const float* row;
  if (r < 2) { b = w.b_deltas + r; row = w.w_deltas + (long long)r * K; }
  else if (r < 5) { b = w.b_dims + (r - 2); row = w.w_dims + (long long)(r - 2) * K; }
  else if (r == 5) { b = w.b_z; row = w.w_z; }
  else if (r == 6 && conf_rows) { b = w.b_uncert; row = w.w_uncert; }
  else { b = w.b_pose + (r - 6 - conf_rows); row = w.w_pose + (long long)(r - 6 - conf_rows) * K; }
  if (bias) *bias = *b;
  return row;
}

// ovm_head_set_trace: [grid, 8] %globaltimer stamps per CTA -- 0 entry, 1 set-up + staging done (cluster barrier passed),
// 2 the FC layer before has completed (programmatic-launch wait returned), 3 first operand stage landed, 4 accumulator
// complete, 5 epilogue (bias, ReLU, head partials) done, 6 partials exchanged (cluster barrier), 7 records written
__device__ __forceinline__ void tail_stamp(const TailArgs& g, int slot) {
  if (g.d.trace) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    g.d.trace[(size_t)blockIdx.x * 8 + slot] = t;
  }
}

// The finish of the layer before (TailPending) by the eight epilogue warps of one CTA; see the call site.  Kept out of line:
// inlined, its sixteen float4 in flight changed the register allocation of the whole kernel (109 -> 128) and cost the
// main loop and the decode 1.5 us.
template <int FMT>
__device__ __noinline__ void finish_prev_layer(float* acc, long long acc_ld, const float* bias_prev, void* a_out, long long a_ld,
                                               int K, int n, int cs, int rank, int m_tile, bool conv_grouped,
                                               uint32_t tfull_bar) {   // (scalars: a reference to the kernel's
                                                                       // parameter struct would copy it to local memory)
  auto conv_bar = [&](int c) { return c == 0 ? tfull_bar + 8 : tfull_bar + 16 + 8 * (uint32_t)c; };
    const int slice = kTM / cs;
    const int row0 = m_tile * kTM + rank * slice;
    auto finish4 = [&](float4 w, float4 bv, long long row, int c) {
      __stcg(reinterpret_cast<float4*>(acc + row * acc_ld + c), make_float4(0.f, 0.f, 0.f, 0.f));
      w.x += bv.x; w.y += bv.y; w.z += bv.z; w.w += bv.w;
      w.x = w.x < 0.f ? 0.f : w.x; w.y = w.y < 0.f ? 0.f : w.y;     // torch.relu keeps NaN
      w.z = w.z < 0.f ? 0.f : w.z; w.w = w.w < 0.f ? 0.f : w.w;
      if (FMT == kFmtTF32)
        *reinterpret_cast<float4*>(reinterpret_cast<float*>(a_out) + row * a_ld + c) = round_tf32(w);
      else
        *reinterpret_cast<uint2*>(reinterpret_cast<uint16_t*>(a_out) + row * a_ld + c) = pack16(w, FMT == kFmtF16);
    };
    const int et = threadIdx.x - 64;
    if (conv_grouped) {
      // slice * K = 16384 floats = 4096 float4 = 16 per thread, 4 per thread and group; k4g float4 per row and group
      const int k4g = K / (4 * kConvGroups), sh = __ffs(k4g) - 1;
      // the thread's column within a group is the same for its four rows (256 is a multiple of k4g): one bias float4 per
      // group, loaded ahead of everything (inside the loop it was one more round trip per group, behind the fence)
      float4 bvg[kConvGroups];
#pragma unroll
      for (int grp = 0; grp < kConvGroups; ++grp)
        bvg[grp] = __ldg(reinterpret_cast<const float4*>(bias_prev + (grp * k4g + (et & (k4g - 1))) * 4));
      float4 x[16];
#pragma unroll
      for (int u = 0; u < 16; ++u) {
        const int grp = u >> 2, rem = (u & 3) * 256 + et;
        const int r = rem >> sh, c = (grp * k4g + (rem & (k4g - 1))) * 4;
        x[u] = row0 + r < n ? __ldcg(reinterpret_cast<const float4*>(acc + (long long)(row0 + r) * acc_ld + c))
                                : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int grp = 0; grp < kConvGroups; ++grp) {
#pragma unroll
        for (int v = 0; v < 4; ++v) {
          const int rem = v * 256 + et;
          const int r = rem >> sh, c = (grp * k4g + (rem & (k4g - 1))) * 4;
          if (row0 + r < n) finish4(x[grp * 4 + v], bvg[grp], row0 + r, c);
        }
        asm volatile("fence.proxy.async.global;" ::: "memory");
        mbar_arrive(conv_bar(grp));
      }
    } else {
      const int k4 = K >> 2;
      const int total = slice * k4;
      constexpr int kBatch = 8;
      for (int i0 = et; i0 < total; i0 += kTailEpiWarps * 32 * kBatch) {
        float4 x[kBatch];
#pragma unroll
        for (int u = 0; u < kBatch; ++u) {
          const int i = i0 + u * kTailEpiWarps * 32;
          const int r = i / k4, c = (i - r * k4) * 4;
          if (i < total && row0 + r < n) x[u] = __ldcg(reinterpret_cast<const float4*>(acc + (long long)(row0 + r) * acc_ld + c));
        }
#pragma unroll
        for (int u = 0; u < kBatch; ++u) {
          const int i = i0 + u * kTailEpiWarps * 32;
          const int r = i / k4, c = (i - r * k4) * 4;
          if (i < total && row0 + r < n) finish4(x[u], __ldg(reinterpret_cast<const float4*>(bias_prev + c)), row0 + r, c);
        }
      }
      asm volatile("fence.proxy.async.global;" ::: "memory");
      mbar_arrive(conv_bar(0));
    }
}

template <int FMT>
__global__ void __launch_bounds__(kTailThreads, 1)
fc_tail_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const TailArgs g) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* aligned = smem_raw + (base - raw);
  const uint32_t smem_a = base;
  const uint32_t smem_b = base + kTailStages * kTailABytes;
  const uint32_t bars = base + kTailStages * kTailStageBytes;
  const uint32_t full_bar = bars;                          // kTailStages
  const uint32_t empty_bar = bars + 8 * kTailStages;       // kTailStages
  const uint32_t tfull_bar = bars + 16 * kTailStages;      // 1
  const uint32_t tmem_slot = tfull_bar + 16;
  // conversion barriers (TailPending) in the free words of the 128-byte barrier block: +88, +104, +112, +120
  auto conv_bar = [&](int c) { return c == 0 ? tfull_bar + 8 : tfull_bar + 16 + 8 * (uint32_t)c; };
  static_assert(16 * kTailStages + 16 + 8 * kConvGroups <= 128, "barrier block is 128 bytes");
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(aligned + kTailStages * kTailStageBytes + 16 * kTailStages + 16);
  float* hw_t = reinterpret_cast<float*>(aligned + kTailStages * kTailStageBytes + 128);   // [128 columns][16 rows]
  float* b2s = hw_t + kTN * kHwStride;                                                     // [128]
  float* hbias = b2s + kTN;                                                                // [16]
  float* part = hbias + 16;                                                                // [2 cs][13][rpc]
  float* xch = part + kPartFloats;                                                         // [rpc][kXchStride]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) tail_stamp(g, 0);
  const int cs = g.cs;
  const int rank = (int)cluster_ctarank();
  const int m_tile = blockIdx.x / cs;
  const int rpc = kTM / cs;                     // box rows this CTA decodes
  const ovm_decode_cfg_t& cfg = g.d.cfg;
  const int pose_n = cfg.pose_type == OVM_POSE_6D ? 6 : (cfg.pose_type == OVM_POSE_QUATERNION ? 4 : 3);
  const int conf_rows = cfg.use_confidence ? 1 : 0;
  const int nrows = 6 + conf_rows + pose_n;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kTailStages; ++s) {
      mbar_init(full_bar + 8 * s, 1);
      mbar_init(empty_bar + 8 * s, cs);      // a slot is free when every CTA of the cluster has consumed it (A is multicast)
    }
    mbar_init(tfull_bar, 1);
    for (int c = 0; c < kConvGroups; ++c) mbar_init(conv_bar(c), kTailEpiWarps * 32);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"((uint32_t)kTN) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  // the boxes this CTA decodes are inputs of the call (complete before the chain's first kernel): loaded ahead of the wait
  // Two threads per box: role 0 takes the pose / dims chain and writes the record, role 1 the centre chain (projected
  // centre, allocentric rotation, depth, back-projection, score).  On one thread the decode is a ~1100-instruction
  // dependent chain at ~10 cycles per instruction (5.9 us); the two chains are independent until the last matrix product.
  // The two roles of a box sit in DIFFERENT warps (role 1 starts at a warp boundary): with 16 boxes per CTA both used to
  // share one warp, whose divergent halves ran the two chains one after the other.
  const int e = threadIdx.x - 64;
  const int rs = rpc < 32 ? 32 : rpc;
  const int role = e >= rs ? 1 : 0, t = e - role * rs;
  const int my_row = m_tile * kTM + rank * rpc + t;
  const bool decoder = e >= 0 && t < rpc && e < rs + rpc && my_row < g.d.n;
  BoxIn in{};
  BoxCam cam{};
  if (decoder) in = load_box_in(g.d, my_row);   // two dependent global loads (image index -> camera row): issued first, in
                                                // flight under the staging loads below; consumed by box_camera after them
  // parameters of the model (not results of the kernels before): staged while fc1 is still running
  if (warp >= 2) {
    const int e = threadIdx.x - 64;             // 0 .. 255
    const int col0 = rank * kTN;
    // thread -> column j of head rows r0, r0 + 2, .. (consecutive threads read consecutive columns of one row).  All
    // eight loads are issued before the first store: as a load / store loop this was eight serial round trips (5 us of
    // the 6 us set-up, measured with a stamp ahead of the cluster barrier).
    static_assert(kTN * kHwStride == 8 * kTailEpiWarps * 32 && kTailEpiWarps * 32 == 2 * kTN, "8 head rows per thread");
    const int j = e & (kTN - 1), r0 = e >> 7;
    float v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int r = r0 + 2 * k;
      const int rr = r < nrows ? r : 0;
      const float* row = agnostic_head_row(g.d.w, rr, conf_rows, g.N, nullptr);
      v[k] = __ldg(row + col0 + j);
      if (r >= nrows) v[k] = 0.f;
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) hw_t[j * kHwStride + r0 + 2 * k] = v[k];
    if (e < kMaxHeadRows && e < nrows) {
      float b = 0.f;
      agnostic_head_row(g.d.w, e, conf_rows, g.N, &b);
      hbias[e] = b;
    }
    if (e < kTN) b2s[e] = __ldg(g.bias + col0 + e);
  }
  if (decoder) cam = box_camera(cfg, in);

  tcgen05_fence_before();
  cluster_sync_all();        // barriers initialised, TMEM allocated, staging visible; every CTA of the cluster has started
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  if (threadIdx.x == 0) tail_stamp(g, 1);
  pdl_trigger();
  pdl_wait();                // h1 (or its raw sums) comes from the FC layer before
  if (threadIdx.x == 0) tail_stamp(g, 2);

  // ===== finish of the layer before (TailPending), for exactly the rows this CTA's TMA loads (and multicasts): rows
  // [rank * slice, +slice) of the tile, all K columns.  Generic-proxy global stores -> fence.proxy.async -> mbarrier ->
  // TMA (async proxy) reads of the same addresses by the same CTA: no cluster-wide synchronisation is needed.  The
  // columns go in kConvGroups groups of whole K blocks, each with its own barrier, so that the producer starts on the
  // first group while the others are still being converted; every load is issued before the first result is used.
  const bool conv_grouped = g.acc && (kTM / cs) * g.K == 16 * 1024 && g.kblocks % kConvGroups == 0;
  if (g.acc && warp >= 2)
    finish_prev_layer<FMT>(g.acc, g.acc_ld, g.bias_prev, g.a_out, g.a_ld, g.K, g.d.n, cs, rank, m_tile, conv_grouped, tfull_bar);

  if (warp == 0) {
    if (lane == 0) {
      // ===== TMA producer.  The 128 x 128-byte A block of a stage is the same for the whole cluster: every CTA fetches
      // 128 / CS of its rows and multicasts them to all CS CTAs (each full barrier still sees 16 KB of A + 16 KB of B),
      // which takes the A traffic out of L2 from CS x to 1 x (106 -> 60 MB per launch at C2 with TF32 operands).
      int stage = 0;
      uint32_t phase = 0;
      const uint16_t mask = (uint16_t)((1u << cs) - 1u);
      const int slice_rows = kTM / cs;
      const int conv_every = conv_grouped ? g.kblocks / kConvGroups : g.kblocks;
      int conv_next = g.acc ? 0 : g.kblocks, conv_grp = 0;
      for (int kb = 0; kb < g.kblocks; ++kb) {
        if (kb == conv_next) {     // this CTA's rows of the next group of K blocks are written
          mbar_wait(conv_bar(conv_grp++), 0);
          conv_next += conv_every;
        }
        mbar_wait(empty_bar + 8 * stage, phase ^ 1);
        mbar_expect_tx(full_bar + 8 * stage, kTailStageBytes);
        const uint32_t a_dst = smem_a + stage * kTailABytes + rank * slice_rows * kRowBytes;
        if (cs > 1)
          tma_load_2d_mc(a_dst, &tmA, full_bar + 8 * stage, kb * g.elems_per_row, m_tile * kTM + rank * slice_rows, mask);
        else
          tma_load_2d(a_dst, &tmA, full_bar + 8 * stage, kb * g.elems_per_row, m_tile * kTM);
        tma_load_2d(smem_b + stage * kTailBBytes, &tmB, full_bar + 8 * stage, kb * g.elems_per_row, rank * kTN);
        if (++stage == kTailStages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ===== MMA issuer: D[128 x 128] += A[128 x 32 B] . B[128 x 32 B]^T, four per stage
      constexpr uint32_t fmt = (uint32_t)FMT;
      constexpr uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(kTN >> 3) << 17) | ((uint32_t)(kTM >> 4) << 24);
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = 0; kb < g.kblocks; ++kb) {
        mbar_wait(full_bar + 8 * stage, phase);
        tcgen05_fence_after();
        if (kb == 0) tail_stamp(g, 3);
        const uint64_t adesc = make_smem_desc(smem_a + stage * kTailABytes);
        const uint64_t bdesc = make_smem_desc(smem_b + stage * kTailBBytes);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma<FMT>(tmem_base, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (kb > 0 || k > 0) ? 1u : 0u);
        if (cs > 1) tcgen05_commit_mc(empty_bar + 8 * stage, (uint16_t)((1u << cs) - 1u));   // frees the slot in every CTA
        else tcgen05_commit(empty_bar + 8 * stage);
        if (++stage == kTailStages) {
          stage = 0;
          phase ^= 1;
        }
      }
      tcgen05_commit(tfull_bar);
    }
  } else {
    // ===== epilogue: thread = box row of the tile (TMEM lane), 64 columns per warp
    const int quad = warp & 3;
    const int half = (warp - 2) >> 2;
    const int row = quad * 32 + lane;
    float p[kMaxHeadRows];
#pragma unroll
    for (int k = 0; k < kMaxHeadRows; ++k) p[k] = 0.f;
    mbar_wait(tfull_bar, 0);
    tcgen05_fence_after();
    if (threadIdx.x == 64) tail_stamp(g, 4);
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
      uint32_t v[16];
      const int colb = half * 64 + c * 16;
      tmem_ld16(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)colb, v);
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        float x = __uint_as_float(v[i]) + b2s[colb + i];
        x = x < 0.f ? 0.f : x;                                  // torch.relu keeps NaN
        const float4* w4 = reinterpret_cast<const float4*>(hw_t + (colb + i) * kHwStride);
        const float4 w0 = w4[0], w1 = w4[1], w2 = w4[2], w3 = w4[3];
        p[0] = fmaf(x, w0.x, p[0]); p[1] = fmaf(x, w0.y, p[1]); p[2] = fmaf(x, w0.z, p[2]); p[3] = fmaf(x, w0.w, p[3]);
        p[4] = fmaf(x, w1.x, p[4]); p[5] = fmaf(x, w1.y, p[5]); p[6] = fmaf(x, w1.z, p[6]); p[7] = fmaf(x, w1.w, p[7]);
        p[8] = fmaf(x, w2.x, p[8]); p[9] = fmaf(x, w2.y, p[9]); p[10] = fmaf(x, w2.z, p[10]); p[11] = fmaf(x, w2.w, p[11]);
        p[12] = fmaf(x, w3.x, p[12]);
      }
    }
    tcgen05_fence_before();
    // partial sums of this row -> the CTA that decodes it
    const int dst = row / rpc, t = row - dst * rpc;
    const int contributor = rank * 2 + half;
    const uint32_t remote = mapa_shared(smem_u32(part), (uint32_t)dst);
#pragma unroll
    for (int k = 0; k < kMaxHeadRows; ++k)
      st_cluster_f32(remote + (uint32_t)(((contributor * kMaxHeadRows + k) * rpc + t) * 4), p[k]);
    if (threadIdx.x == 64) tail_stamp(g, 5);
  }
  __syncwarp();
  cluster_sync_all();        // every partial has landed (release / acquire at cluster scope)
  if (threadIdx.x == 64) tail_stamp(g, 6);
  if (warp == 2) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)kTN) : "memory");
  }
  float hv[kMaxHeadRows];
#pragma unroll
  for (int k = 0; k < kMaxHeadRows; ++k) hv[k] = 0.f;
  if (decoder) {
    // contributions outermost: 13 independent load + add chains per step (the other nesting ran one 16-long dependent
    // chain per row, 13 times in a row: 3.7 of the decode's 5.9 us).  Fixed order: deterministic.
    for (int c = 0; c < 2 * cs; ++c) {
      const float* pc = part + (size_t)c * kMaxHeadRows * rpc + t;
#pragma unroll
      for (int k = 0; k < kMaxHeadRows; ++k) hv[k] += pc[k * rpc];
    }
#pragma unroll
    for (int k = 0; k < kMaxHeadRows; ++k) hv[k] += hbias[k < nrows ? k : 0];
  }
  const HeadOut h = split_heads(cfg, hv);
  float R[9], dims[3];
  if (decoder && role == 0) decode_pose_dims(g.d, in, h.pose_raw, h.dims_raw, R, dims);
  if (decoder && role == 1) {
    const CentreOut o = decode_centre(g.d, in, cam, 0, h.deltas[0], h.deltas[1], h.z_raw, h.uncert);
    float* x = xch + t * kXchStride;
#pragma unroll
    for (int i = 0; i < 9; ++i) x[i] = o.M[i];
    x[9] = __int_as_float(o.has_m);
    x[10] = o.X; x[11] = o.Y; x[12] = o.z; x[13] = o.u_out; x[14] = o.v_out; x[15] = o.score;
    x[16] = h.deltas[0]; x[17] = h.deltas[1]; x[18] = h.z_raw; x[19] = h.uncert;
  }
  __syncthreads();
  if (decoder && role == 0) {
    const float* x = xch + t * kXchStride;
    if (__float_as_int(x[9])) {
      float M[9], T[9];
#pragma unroll
      for (int i = 0; i < 9; ++i) M[i] = x[i];
      matmul3(M, R, T);
#pragma unroll
      for (int i = 0; i < 9; ++i) R[i] = T[i];
    }
    const float raw13[13] = {x[16], x[17], x[18], h.dims_raw[0], h.dims_raw[1], h.dims_raw[2], x[19],
                             h.pose_raw[0], h.pose_raw[1], h.pose_raw[2], h.pose_raw[3], h.pose_raw[4], h.pose_raw[5]};
    write_record(g.d, my_row, in, R, dims, x[10], x[11], x[12], x[13], x[14], x[15], raw13, /*store_peers=*/false);
  }
  if (g.d.num_peers > 0) {
    // fused gather: the CTA's records (contiguous rows) go to every peer rank with all threads, 16 bytes each, instead
    // of 12 x peers dependent stores at the end of each decoding thread's chain
    __syncthreads();          // the decoding threads' record stores are visible to the block
    const int row0 = m_tile * kTM + rank * rpc;
    const int rows = min(rpc, g.d.n - row0);
    if (rows > 0) {
      const int per_peer = rows * (OVM_RECORD_WIDTH / 4);
      const float4* src = reinterpret_cast<const float4*>(g.d.records + (long long)row0 * OVM_RECORD_WIDTH);
      for (int i = threadIdx.x; i < per_peer * g.d.num_peers; i += kTailThreads) {
        const int p = i / per_peer, j = i - p * per_peer;
        reinterpret_cast<float4*>(g.d.peer_records[p] + (long long)row0 * OVM_RECORD_WIDTH)[j] = src[j];
      }
    }
  }
  if (threadIdx.x == 64) tail_stamp(g, 7);
}

// [rows, cols] row-major with leading dimension ld (elements) -> 2-D map, box {128 bytes of K, box_rows rows}, 128-byte
// swizzle, out-of-bounds elements read as zero (row and K tails).
int make_tail_map(CUtensorMap* map, const void* ptr, int64_t rows, int64_t cols, int64_t ld, int dtype, int box_rows) {
  EncodeTiledFn enc = tensormap_encode_fn();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled not available from the driver");
    return OVM_ECUDA;
  }
  const int esz = dtype == OVM_TF32 ? 4 : 2;
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * esz};
  cuuint32_t box[2] = {(cuuint32_t)(kRowBytes / esz), (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, dtype == OVM_TF32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : (dtype == OVM_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16), 2, const_cast<void*>(ptr), dims,
                   strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled (fc tail) failed with CUresult %d (rows=%lld cols=%lld ld=%lld)", (int)r, (long long)rows,
              (long long)cols, (long long)ld);
    return OVM_ECUDA;
  }
  return OVM_OK;
}

template <int FMT>
int launch_tail(const CUtensorMap& tmA, const CUtensorMap& tmB, const TailArgs& g, int grid, cudaStream_t stream) {
  static SmemOptIn opt;
  {
    int rc = opt.ensure(fc_tail_kernel<FMT>, kTailSmemBytes);
    if (rc != OVM_OK) return rc;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kTailThreads);
  cfg.dynamicSmemBytes = kTailSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  attr[1].id = cudaLaunchAttributeClusterDimension;
  attr[1].val.clusterDim.x = g.cs;
  attr[1].val.clusterDim.y = 1;
  attr[1].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  OVM_CUDA(cudaLaunchKernelEx(&cfg, fc_tail_kernel<FMT>, tmA, tmB, g));
  OVM_LAUNCHED("fc_tail_kernel");
  return OVM_OK;
}

}  // namespace

bool fc_tail_supported(const ovm_decode_cfg_t* cfg, int32_t N, int32_t K) {
  static const bool enabled = [] { const char* v = getenv("OVM_FUSE_TAIL"); return !v || atoi(v) != 0; }();
  if (!enabled || !cfg) return false;
  if (cfg->dims_priors_enabled || cfg->cluster_bins != 1) return false;   // head rows depend on the box's class / bin
  for (int i = 1; i < 5; ++i)
    if (cfg->feat_offset[i] != cfg->feat_offset[0]) return false;         // one FC stack per head (SHARED_FC = False)
  if (N != cfg->fc_dim || K <= 0) return false;
  return N == 128 || N == 256 || N == 512 || N == 1024;
}

int fc_tail_launch(const void* A, int64_t lda, int32_t a_dtype, const void* W, int64_t ldw, const float* bias, int32_t N,
                   int32_t K, const ovm_decode_cfg_t* cfg, const ovm_head_weights_t* w, const float* boxes,
                   const int64_t* classes, const float* scores, const int32_t* batch_idx, const float* cams,
                   int32_t num_images, int32_t n, float* records, float* raw, cudaStream_t stream,
                   const ovm_peer_records_t* peers, const TailPending* pending) {
  OVM_REQUIRE(cfg && w, "ovm_fc_tail: cfg / weights are NULL");
  OVM_REQUIRE(a_dtype == OVM_BF16 || a_dtype == OVM_TF32 || a_dtype == OVM_F16,
              "ovm_fc_tail: operand dtype %d (OVM_BF16, OVM_F16 or OVM_TF32)", a_dtype);
  OVM_REQUIRE(cfg->z_type >= OVM_Z_DIRECT && cfg->z_type <= OVM_Z_CLUSTERS, "ovm_fc_tail: z_type %d", cfg->z_type);
  OVM_REQUIRE(cfg->pose_type >= OVM_POSE_6D && cfg->pose_type <= OVM_POSE_EULER,
              "Cuboid pose type %d is not recognized", cfg->pose_type);
  if (!fc_tail_supported(cfg, N, K)) {
    set_error("ovm_fc_tail: needs class-agnostic heads (DIMS_PRIORS_ENABLED off, CLUSTER_BINS 1), SHARED_FC and "
              "FC_DIM in {128, 256, 512, 1024}; use ovm_gemm_bf16 + ovm_head_decode otherwise");
    return OVM_EUNSUPPORTED;
  }
  OVM_REQUIRE(cfg->z_type != OVM_Z_CLUSTERS || w->priors_z_stats, "ovm_fc_tail: priors_z_stats needed for Z_TYPE clusters");
  OVM_REQUIRE(w->w_deltas && w->b_deltas && w->w_dims && w->b_dims && w->w_pose && w->b_pose && w->w_z && w->b_z,
              "ovm_fc_tail: a head weight is NULL");
  OVM_REQUIRE(!cfg->use_confidence || (w->w_uncert && w->b_uncert), "ovm_fc_tail: uncertainty head is NULL");
  if (n == 0) return OVM_OK;
  const int esz = a_dtype == OVM_TF32 ? 4 : 2;
  OVM_REQUIRE(n > 0 && A && W && bias && boxes && classes && batch_idx && cams && records && num_images > 0,
              "ovm_fc_tail: NULL input with n=%d", n);
  OVM_REQUIRE(lda >= K && ldw >= K && (lda * esz) % 16 == 0 && (ldw * esz) % 16 == 0 &&
                  !(reinterpret_cast<uintptr_t>(A) & 15) && !(reinterpret_cast<uintptr_t>(W) & 15),
              "ovm_fc_tail: operands must be 16-byte aligned with 16-byte row pitches >= K");
  OVM_REQUIRE(!(reinterpret_cast<uintptr_t>(records) & 15) && (!raw || !(reinterpret_cast<uintptr_t>(raw) & 15)),
              "ovm_fc_tail: records / raw must be 16-byte aligned");
  TailArgs g{};
  g.d.cfg = *cfg;
  g.d.w = *w;
  g.d.boxes = boxes;
  g.d.classes = reinterpret_cast<const long long*>(classes);
  g.d.scores = scores;
  g.d.batch_idx = batch_idx;
  g.d.cams = cams;
  g.d.num_images = num_images;
  g.d.n = n;
  g.d.records = records;
  g.d.raw = raw;
  g.d.trace = head_get_trace();
  if (peers) {
    OVM_REQUIRE(peers->num_peers >= 0 && peers->num_peers <= OVM_MAX_PEERS, "ovm_lift_forward_peers: num_peers %d", peers->num_peers);
    for (int p = 0; p < peers->num_peers; ++p) {
      OVM_REQUIRE(peers->records[p] && !(reinterpret_cast<uintptr_t>(peers->records[p]) & 15),
                  "ovm_lift_forward_peers: peer buffer %d is NULL or not 16-byte aligned", p);
      g.d.peer_records[p] = peers->records[p];
    }
    g.d.num_peers = peers->num_peers;
  }
  g.K = K;
  g.N = N;
  g.cs = N / kTN;
  g.elems_per_row = kRowBytes / esz;
  g.kblocks = (K + g.elems_per_row - 1) / g.elems_per_row;
  g.bias = bias;
  if (pending) {
    OVM_REQUIRE(pending->acc && pending->bias && pending->ld >= K && pending->ld % 4 == 0 && K % 4 == 0 &&
                    !(reinterpret_cast<uintptr_t>(pending->acc) & 15) && !(reinterpret_cast<uintptr_t>(pending->bias) & 15),
                "ovm_fc_tail: pending accumulator must be [n, ld >= K] fp32, 16-byte aligned, K %% 4 == 0");
    g.acc = pending->acc;
    g.acc_ld = pending->ld;
    g.bias_prev = pending->bias;
    g.a_out = const_cast<void*>(A);
    g.a_ld = lda;
  }
  CUtensorMap tmA, tmB;
  int rc = make_tail_map(&tmA, A, n, K, lda, a_dtype, kTM / g.cs);   // each CTA multicasts its slice of the rows
  if (rc != OVM_OK) return rc;
  rc = make_tail_map(&tmB, W, N, K, ldw, a_dtype, kTN);
  if (rc != OVM_OK) return rc;
  const int grid = g.cs * ((n + kTM - 1) / kTM);
  if (a_dtype == OVM_TF32) return launch_tail<kFmtTF32>(tmA, tmB, g, grid, stream);
  if (a_dtype == OVM_F16) return launch_tail<kFmtF16>(tmA, tmB, g, grid, stream);
  return launch_tail<kFmtBF16>(tmA, tmB, g, grid, stream);
}

}  // namespace ovm
