// Kernel 2: the cube-head FC layers as tcgen05 GEMMs.
//
//   D[M,N] = act(A[M,K] . B[N,K]^T + bias)   A = activations (boxes x features), B = nn.Linear weight, both bf16
//   K-major; fp32 accumulation in TMEM.  Replaces the cuBLAS SGEMMs behind nn.Linear in CubeHead
//   (reference cubercnn/modeling/roi_heads/cube_head.py:64-74,157).
//
// One CTA = 6 warps: warp 0 TMA producer (one elected lane), warp 1 tcgen05.mma issuer (one elected lane),
// warps 2-5 epilogue (tcgen05.ld -> registers -> bias/ReLU -> global).  Operand tiles are 128 x 64 (A) and
// BN x 64 (B) bf16, brought in by TMA with the 128-byte swizzle into a kStages-deep ring guarded by
// full/empty mbarriers; tcgen05.commit releases a ring slot when the MMAs that read it have retired.
// The fp32 accumulator (128 lanes x BN columns) is double-buffered in TMEM so the epilogue of one
// work segment overlaps the MMAs of the next.
//
// Scheduling is a flat list of 64-wide K blocks: tile-major, K-minor.  CTA i of G owns blocks
// [i*T/G, (i+1)*T/G).  DIRECT mode launches G = #tiles, so every CTA owns exactly one whole tile and the
// epilogue applies bias/ReLU and stores.  STREAMK mode launches G = #SMs; a CTA's range may start and end
// inside a tile, and every segment is added to the zero-initialised fp32 output with red.global.add.v4.f32
// (bias/ReLU then run in bias_act_kernel).  fc1 of the lift (M = 1600, N = 1024, K = 12544) has only 52-104
// tiles for 148 SMs, which is why it needs the K split.
#include <cuda.h>
#include <cuda_fp16.h>

#include <cstdlib>
#include <cstring>

#include "common.cuh"

namespace ovm {

namespace {

constexpr int BM = 128;
constexpr int BK = 64;            // 64 bf16 = 128 bytes = one swizzle row
constexpr int UMMA_K = 16;
constexpr int kEpiWarps = 8;          // two per TMEM lane quadrant, each takes half of the tile's columns
constexpr int kGemmThreads = (2 + kEpiWarps) * 32;
constexpr int kChunk = 16;            // accumulator columns per tcgen05.ld
constexpr int kStgFloats = 32 * kChunk;   // swizzled 32 x 16 fp32 transpose tile per epilogue warp
enum { kModeF32 = 0, kModeBF16 = 1, kModeStreamK = 2 };

template <int BN>
struct GemmShape {
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBBytes = BN * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = BN == 256 ? 4 : (BN == 128 ? 6 : 8);
  static constexpr int kTmemCols = BN * 2 < 32 ? 32 : BN * 2;   // two accumulator stages
  static constexpr int kBarrierBytes = (2 * kStages + 4) * 8 + 16;
  static constexpr int kStgBytes = kEpiWarps * kStgFloats * 4;    // epilogue transpose tiles, one per warp
  static constexpr size_t kSmemBytes = 1024 + (size_t)kStages * kStageBytes + kBarrierBytes + kStgBytes;
};

struct GemmArgs {
  int M, N, K;
  int num_n_tiles, num_tiles, kb_per_tile;
  long long total_kb;
  int streamk;           // 1: red.add into fp32 out, no bias/relu
  int relu;
  int out_bf16;
  const float* bias;
  void* out;
  long long ldo;
  unsigned long long* trace;   // optional [grid, 16] globaltimer stamps (ovm_gemm_set_trace), else NULL
  // split-K with a fused finish: the CTA whose segment completes a tile (per-tile counter of K blocks) reads the
  // accumulated tile back, applies bias / ReLU, writes `fin_out` and hands accumulator and counter back zeroed
  int* tile_counters;          // NULL: raw accumulation only (a bias_act pass follows)
  void* fin_out;
  long long fin_ldo;
  int fin_bf16;
  int f16;               // 16-bit outputs (DIRECT epilogue / fused finish) are fp16 instead of bf16
  int round_tf32;        // fp32 outputs (DIRECT epilogue / fused finish) are rounded to TF32: the next layer is a kind::tf32 MMA
  int grid_finish;       // gemm_pair_kernel<false>: all CTAs finish 1/grid of the output after a grid-wide barrier
  // gemm_pair_kernel<true>: split-K partials meet in per-cluster slots (plain stores) instead of L2 atomics
  float* slots;          // [clusters][2 CTAs][128 rows][256 columns] fp32
  int* slot_flags;       // [clusters][2 CTAs][8 epilogue warps], zero between launches
};

thread_local unsigned long long* g_gemm_trace = nullptr;

__device__ __forceinline__ void trace_stamp(const GemmArgs& g, int slot) {
  if (g.trace) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    g.trace[(size_t)blockIdx.x * 16 + slot] = t;
  }
}

// ------------------------------------------------------------------ PTX wrappers
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
__device__ __forceinline__ void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
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
// (round_tf32, sat_f16, pack16: common.cuh -- shared with fc_tail.cu)
__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// Shared-memory matrix descriptor, K-major operand, 128-byte swizzle (cute::UMMA::SmemDescriptor, sm100):
// start>>4 [0,14) | LBO>>4 [16,30) (unused for swizzled K-major, 1) | SBO>>4 [32,46) = 1024 B between 8-row
// groups | version=1 [46,48) | layout SWIZZLE_128B=2 [61,64).
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// Instruction descriptor for kind::f16 (cute::UMMA::InstrDescriptor): D=f32 [4,6)=1, A=bf16 [7,10)=1,
// B=bf16 [10,13)=1, both K-major (bits 15,16 = 0), N>>3 at [17,23), M>>4 at [24,29).
template <int BN>
__device__ __forceinline__ constexpr uint32_t make_idesc() {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
}

// ------------------------------------------------------------------ work iteration shared by the three roles
struct Segment {
  int m_tile, n_tile, k0, k1;
};
struct WorkIter {
  long long cur, end;
  int kb_per_tile, num_n_tiles;
  __device__ WorkIter(const GemmArgs& g) {
    cur = g.total_kb * (long long)blockIdx.x / gridDim.x;
    end = g.total_kb * (long long)(blockIdx.x + 1) / gridDim.x;
    kb_per_tile = g.kb_per_tile;
    num_n_tiles = g.num_n_tiles;
  }
  __device__ bool next(Segment& s) {
    if (cur >= end) return false;
    const int tile = (int)(cur / kb_per_tile);
    s.k0 = (int)(cur - (long long)tile * kb_per_tile);
    const long long left = end - cur;
    s.k1 = (int)((long long)s.k0 + left < (long long)kb_per_tile ? (long long)s.k0 + left : (long long)kb_per_tile);
    s.m_tile = tile / num_n_tiles;
    s.n_tile = tile - s.m_tile * num_n_tiles;
    cur += s.k1 - s.k0;
    return true;
  }
};

template <int BN, int MODE>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmArgs g) {
  using Sh = GemmShape<BN>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;          // SWIZZLE_128B tiles need 1024-byte alignment
  const uint32_t smem_a = base;
  const uint32_t smem_b = base + Sh::kStages * Sh::kABytes;
  const uint32_t bars = base + Sh::kStages * Sh::kStageBytes;
  const uint32_t full_bar = bars;                         // kStages
  const uint32_t empty_bar = bars + 8 * Sh::kStages;      // kStages
  const uint32_t tfull_bar = bars + 16 * Sh::kStages;     // 2
  const uint32_t tempty_bar = tfull_bar + 16;             // 2
  const uint32_t tmem_slot = tempty_bar + 16;
  const uint32_t stage_buf = tmem_slot + 16;
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) trace_stamp(g, 0);

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < Sh::kStages; ++s) {
      mbar_init(full_bar + 8 * s, 1);
      mbar_init(empty_bar + 8 * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull_bar + 8 * s, 1);
      mbar_init(tempty_bar + 8 * s, kEpiWarps);   // one arrive per epilogue warp
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"((uint32_t)Sh::kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  if (threadIdx.x == 0) trace_stamp(g, 1);
  pdl_trigger();
  pdl_wait();   // A (and, split-K, the zeroed accumulator) come from the kernels before; set-up above overlapped them

  if (warp == 0) {
    if (lane == 0) {
      // ===== TMA producer
      WorkIter it(g);
      Segment s;
      int stage = 0;
      uint32_t phase = 0;
      while (it.next(s)) {
        for (int kb = s.k0; kb < s.k1; ++kb) {
          mbar_wait(empty_bar + 8 * stage, phase ^ 1);
          mbar_expect_tx(full_bar + 8 * stage, Sh::kStageBytes);
          tma_load_2d(smem_a + stage * Sh::kABytes, &tmA, full_bar + 8 * stage, kb * BK, s.m_tile * BM);
          tma_load_2d(smem_b + stage * Sh::kBBytes, &tmB, full_bar + 8 * stage, kb * BK, s.n_tile * BN);
          if (++stage == Sh::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
      trace_stamp(g, 2);   // last TMA issued
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ===== MMA issuer
      constexpr uint32_t idesc = make_idesc<BN>();
      WorkIter it(g);
      Segment s;
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      bool first = true;
      while (it.next(s)) {
        mbar_wait(tempty_bar + 8 * acc, acc_phase ^ 1);
        tcgen05_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)(acc * BN);
        for (int kb = s.k0; kb < s.k1; ++kb) {
          mbar_wait(full_bar + 8 * stage, phase);
          tcgen05_fence_after();
          if (first) {
            trace_stamp(g, 3);   // first operand tile landed
            first = false;
          }
          const uint64_t adesc = make_smem_desc(smem_a + stage * Sh::kABytes);
          const uint64_t bdesc = make_smem_desc(smem_b + stage * Sh::kBBytes);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            // advance 16 bf16 = 32 bytes along K inside the swizzle row: +2 in the (addr >> 4) field
            umma_bf16(tmem_d, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (kb > s.k0 || k > 0) ? 1u : 0u);
          }
          tcgen05_commit(empty_bar + 8 * stage);   // slot reusable once these MMAs retire
          if (++stage == Sh::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        tcgen05_commit(tfull_bar + 8 * acc);       // accumulator complete
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
      trace_stamp(g, 4);   // last MMA issued
    }
  } else {
    // ===== epilogue warps: TMEM lane quadrant = warp % 4, column half = (warp - 2) / 4
    // tcgen05.ld hands lane r the columns of accumulator row r.  Storing that directly scatters every warp store
    // over 32 output rows (measured: 17 us for one 128 x 256 fp32 tile).  Each 32 x 16 chunk is transposed
    // through a swizzled shared-memory tile instead (16-byte unit k of row r lives at unit k ^ ((r >> 1) & 3), so
    // both the row-wise 128-bit writes and the 8-rows-by-4-units 128-bit reads are bank-conflict free); after the
    // transpose lane (q, l4) owns columns 4*l4..4*l4+3 of rows q, q+8, q+16, q+24 and one warp instruction
    // writes eight 64-byte row segments.  Bias / ReLU are applied there.  MODE is a template parameter and the
    // full-chunk case has no per-element predicates: with one warp per scheduler the epilogue is bound by
    // instruction latency, not by bytes.
    const int quad = warp & 3;
    const int half = (warp - 2) >> 2;
    float* stg = reinterpret_cast<float*>(smem_raw + (stage_buf - raw)) + (warp - 2) * kStgFloats;
    const int q = lane >> 2, l4 = lane & 3;
    constexpr int kChunksPerWarp = BN / kChunk / 2;
    WorkIter it(g);
    Segment s;
    int acc = 0;
    uint32_t acc_phase = 0;
    while (it.next(s)) {
      mbar_wait(tfull_bar + 8 * acc, acc_phase);
      tcgen05_fence_after();
      if (threadIdx.x == 64) trace_stamp(g, 5);   // (last) accumulator ready
      const int row0 = s.m_tile * BM + quad * 32;
      const bool rows_full = row0 + 32 <= g.M;
#pragma unroll 1
      for (int ci = 0; ci < kChunksPerWarp; ++ci) {
        const int ch = half * kChunksPerWarp + ci;
        uint32_t v[kChunk];
        __syncwarp();   // tcgen05.ld is .sync.aligned; also: the previous chunk's readers are done with stg
        tmem_ld16(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * BN + ch * kChunk), v);
        const int col0 = s.n_tile * BN + ch * kChunk;
        if (threadIdx.x == 64 && ci < 8) trace_stamp(g, 8 + ci);   // chunk left TMEM
        if (row0 >= g.M || col0 >= g.N) continue;   // warp-uniform
        {
          float4* wr = reinterpret_cast<float4*>(stg + lane * kChunk);
          const int sw = (lane >> 1) & 3;
#pragma unroll
          for (int k = 0; k < 4; ++k)
            wr[k ^ sw] = make_float4(__uint_as_float(v[4 * k]), __uint_as_float(v[4 * k + 1]),
                                     __uint_as_float(v[4 * k + 2]), __uint_as_float(v[4 * k + 3]));
        }
        __syncwarp();
        const int c = col0 + 4 * l4;
        const bool full = rows_full && col0 + kChunk <= g.N;   // warp-uniform
        float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
        if (MODE != kModeStreamK && g.bias) {
          if (c + 3 < g.N) bv = __ldg(reinterpret_cast<const float4*>(g.bias + c));
          else {
            if (c < g.N) bv.x = __ldg(g.bias + c);
            if (c + 1 < g.N) bv.y = __ldg(g.bias + c + 1);
            if (c + 2 < g.N) bv.z = __ldg(g.bias + c + 2);
          }
        }
        const bool relu = MODE != kModeStreamK && g.relu;
#pragma unroll
        for (int rr = 0; rr < 32; rr += 8) {
          const int r = rr + q;
          float4 x = reinterpret_cast<const float4*>(stg + r * kChunk)[l4 ^ ((r >> 1) & 3)];
          if (MODE != kModeStreamK) {
            x.x += bv.x; x.y += bv.y; x.z += bv.z; x.w += bv.w;
            // torch.relu keeps NaN: compare-and-select, not fmaxf
            x.x = (relu && x.x < 0.f) ? 0.f : x.x;
            x.y = (relu && x.y < 0.f) ? 0.f : x.y;
            x.z = (relu && x.z < 0.f) ? 0.f : x.z;
            x.w = (relu && x.w < 0.f) ? 0.f : x.w;
            if (MODE == kModeF32 && g.round_tf32) x = round_tf32(x);
          }
          const long long off = (long long)(row0 + r) * g.ldo + c;
          if (full) {
            if (MODE == kModeStreamK) {
              red_add_v4(reinterpret_cast<float*>(g.out) + off, x.x, x.y, x.z, x.w);
            } else if (MODE == kModeF32) {
              *reinterpret_cast<float4*>(reinterpret_cast<float*>(g.out) + off) = x;
            } else {
              *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(g.out) + off) = pack16(x, g.f16);
            }
          } else if (row0 + r < g.M) {   // tile edge: element-wise
            const float xs[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              if (c + e >= g.N) continue;
              if (MODE == kModeStreamK) atomicAdd(reinterpret_cast<float*>(g.out) + off + e, xs[e]);
              else if (MODE == kModeF32) reinterpret_cast<float*>(g.out)[off + e] = xs[e];
              else if (g.f16) reinterpret_cast<__half*>(g.out)[off + e] = __float2half_rn(sat_f16(xs[e]));
              else reinterpret_cast<__nv_bfloat16*>(g.out)[off + e] = __float2bfloat16_rn(xs[e]);
            }
          }
        }
      }
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar + 8 * acc);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
      if (MODE == kModeStreamK && g.tile_counters) {
        // ---- fused finish (threadFenceReduction pattern): publish this segment, count its K blocks on the tile
        __threadfence();
        asm volatile("bar.sync 2, %0;" ::"n"(kEpiWarps * 32) : "memory");
        int* flag = reinterpret_cast<int*>(smem_raw + (tmem_slot - raw)) + 2;
        const int tile = s.m_tile * g.num_n_tiles + s.n_tile;
        if (threadIdx.x == 64) {
          const int done = atomicAdd(g.tile_counters + tile, s.k1 - s.k0) + (s.k1 - s.k0);
          *flag = done == g.kb_per_tile;
          if (done == g.kb_per_tile) g.tile_counters[tile] = 0;   // nobody else touches this tile any more
        }
        asm volatile("bar.sync 2, %0;" ::"n"(kEpiWarps * 32) : "memory");
        if (*flag) {
          __threadfence();
          float* accp = reinterpret_cast<float*>(g.out);
          const int colw = BN / 2;                      // this warp's half of the tile's columns
          const int cbase = s.n_tile * BN + half * colw;
          // 16 rows per batch with all loads in flight: the finish sits on the kernel's tail, its L2 round trips
          // must overlap (one row at a time cost 20 us)
          constexpr int kRows = 16;
          for (int c = cbase + 4 * lane; c < cbase + colw && c < g.N; c += 128) {
            float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
            if (g.bias) bv = __ldg(reinterpret_cast<const float4*>(g.bias + c));
            for (int rb = 0; rb < 32; rb += kRows) {
              float4 x[kRows];
#pragma unroll
              for (int r = 0; r < kRows; ++r) {
                const int row = row0 + rb + r;
                x[r] = row < g.M ? __ldcg(reinterpret_cast<const float4*>(accp + (long long)row * g.ldo + c))
                                 : make_float4(0.f, 0.f, 0.f, 0.f);      // L2: other SMs wrote it
              }
#pragma unroll
              for (int r = 0; r < kRows; ++r) {
                const int row = row0 + rb + r;
                if (row >= g.M) continue;
                __stcg(reinterpret_cast<float4*>(accp + (long long)row * g.ldo + c), make_float4(0.f, 0.f, 0.f, 0.f));
                float4 v = x[r];
                v.x += bv.x; v.y += bv.y; v.z += bv.z; v.w += bv.w;
                if (g.relu) {   // torch.relu keeps NaN
                  v.x = v.x < 0.f ? 0.f : v.x; v.y = v.y < 0.f ? 0.f : v.y;
                  v.z = v.z < 0.f ? 0.f : v.z; v.w = v.w < 0.f ? 0.f : v.w;
                }
                if (g.round_tf32) v = round_tf32(v);
                const long long off = (long long)row * g.fin_ldo + c;
                if (g.fin_bf16) {
                  *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(g.fin_out) + off) = pack16(v, g.f16);
                } else {
                  *reinterpret_cast<float4*>(reinterpret_cast<float*>(g.fin_out) + off) = v;
                }
              }
            }
          }
        }
      }
    }
    if (threadIdx.x == 64) trace_stamp(g, 6);     // epilogue done
  }

  tcgen05_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) trace_stamp(g, 7);
  if (warp == 2) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)Sh::kTmemCols) : "memory");
  }
}

// ------------------------------------------------------------------ split-K GEMM on CTA pairs (cta_group::2)
// fc1 of the lift is fed from L2: with 128 x 256 tiles per CTA the operand traffic of one launch is 494 MB (every
// A tile read by 4 column tiles, every W1 tile by 13 row tiles).  A CTA pair (cluster of 2, same TPC) works on a
// 256 x 256 tile instead: each CTA loads its own 128 rows of A and HALF of the B tile (128 of its 256 rows), the
// leader issues tcgen05.mma.cta_group::2 (M = 256: rows 0-127 accumulate in the leader's TMEM, 128-255 in the
// peer's; the B operand is read from both CTAs' shared memory) -- 32 KB instead of 48 KB per CTA and K block.
//   * full[s]   (leader's copy is used): the leader's producer arms it with the bytes of BOTH CTAs; the peer's TMA
//               loads (cta_group::2 form) complete on the leader's barrier.
//   * empty[s]  (one per CTA): tcgen05.commit multicast from the leader's MMA thread frees the slot in both CTAs.
//   * tfull[a]  (one per CTA): commit multicast, accumulator stage a is complete in both CTAs' TMEM.
//   * tempty[a] (leader's copy): 2 x kEpiWarps arrivals, the peer's epilogue warps arrive remotely.
// Work is the flat list of (pair tile, K block); cluster c of G owns a contiguous range (stream-K), partial tiles are
// added into the zeroed fp32 accumulator and finished by whichever CTA completes a 128-row tile (see above).
constexpr int kPairStages = 6;
constexpr int kPairBN = 256;
constexpr int kPairABytes = BM * BK * 2;                 // this CTA's 128 rows of A
constexpr int kPairBBytes = (kPairBN / 2) * BK * 2;      // this CTA's half of the B tile
constexpr int kPairStageBytes = kPairABytes + kPairBBytes;
constexpr size_t kPairSmemBytes = 1024 + (size_t)kPairStages * kPairStageBytes + (2 * kPairStages + 4) * 8 + 16 +
                                  kEpiWarps * kStgFloats * 4;

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
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_remote(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA load whose completion is signalled on a barrier of either CTA of the pair
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap* map, uint32_t bar_cluster, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
      "l"(map), "r"(bar_cluster), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tcgen05_commit_pair(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
               "h"((uint16_t)3)
               : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

struct PairSegment {
  int m_pair, n_tile, k0, k1;
};
struct PairIter {
  long long cur, end;
  int kb_per_tile, num_n_tiles;
  __device__ PairIter(const GemmArgs& g, int num_pair_tiles) {
    const long long total = (long long)num_pair_tiles * g.kb_per_tile;
    const int cluster = blockIdx.x >> 1, nclusters = gridDim.x >> 1;
    cur = total * cluster / nclusters;
    end = total * (cluster + 1) / nclusters;
    kb_per_tile = g.kb_per_tile;
    num_n_tiles = g.num_n_tiles;
  }
  __device__ bool next(PairSegment& s) {
    if (cur >= end) return false;
    const int tile = (int)(cur / kb_per_tile);
    s.k0 = (int)(cur - (long long)tile * kb_per_tile);
    const long long left = end - cur;
    s.k1 = (int)((long long)s.k0 + left < (long long)kb_per_tile ? (long long)s.k0 + left : (long long)kb_per_tile);
    s.m_pair = tile / num_n_tiles;
    s.n_tile = tile - s.m_pair * num_n_tiles;
    cur += s.k1 - s.k0;
    return true;
  }
};

__device__ __forceinline__ int ld_acquire_gpu(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(int* p, int v) {
  asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// SLOTS = false: partial tiles meet in the zeroed fp32 accumulator through red.global.add, the CTA that completes a tile
//   (per-tile counter) reads it back and finishes it (round 1).
// SLOTS = true: deterministic, no atomics.  In the flat (tile, K block) order a cluster's range starts inside at most
//   one tile: that first segment (k0 > 0) is a PARTIAL -- its raw fp32 accumulator is stored to the cluster's slot with
//   plain 16-byte stores and published with one release flag per epilogue warp.  Every later segment of the cluster
//   starts a tile (k0 == 0): the cluster OWNS that tile, keeps its own accumulator in tensor memory, and on the way out
//   adds the slots of the clusters whose ranges start inside the tile (cluster + 1, + 2, ...; the host guarantees at most
//   kMaxPartials of them) in a fixed order, then bias / ReLU / cast -> the layer's output.  The partials a cluster
//   produces are the first thing it does, so an owner never waits on anything that waits on it.
constexpr int kMaxPartials = 4;
template <bool SLOTS>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1)
gemm_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmArgs g,
                 const int num_pair_tiles) {
  constexpr int BN = kPairBN;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t smem_a = base;
  const uint32_t smem_b = base + kPairStages * kPairABytes;
  const uint32_t bars = base + kPairStages * kPairStageBytes;
  const uint32_t full_bar = bars;
  const uint32_t empty_bar = bars + 8 * kPairStages;
  const uint32_t tfull_bar = bars + 16 * kPairStages;
  const uint32_t tempty_bar = tfull_bar + 16;
  const uint32_t tmem_slot = tempty_bar + 16;
  const uint32_t stage_buf = tmem_slot + 16;
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  if (threadIdx.x == 0) trace_stamp(g, 0);

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kPairStages; ++s) {
      mbar_init(full_bar + 8 * s, 1);
      mbar_init(empty_bar + 8 * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull_bar + 8 * s, 1);
      mbar_init(tempty_bar + 8 * s, 2 * kEpiWarps);   // both CTAs' epilogue warps (used in the leader only)
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"((uint32_t)(2 * BN))
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  cluster_sync_all();        // both CTAs' barriers are initialised and both TMEM allocations are done
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  pdl_trigger();
  if (threadIdx.x == 0) trace_stamp(g, 1);
  pdl_wait();
  if (threadIdx.x == 0) trace_stamp(g, 3);   // the kernel before has completed

  if (warp == 0) {
    if (lane == 0) {
      // ===== TMA producer (both CTAs): own rows of A, own half of the B tile; completion on the LEADER's barrier
      PairIter it(g, num_pair_tiles);
      PairSegment s;
      int stage = 0;
      uint32_t phase = 0;
      while (it.next(s)) {
        const int a_row = s.m_pair * (2 * BM) + (int)rank * BM;
        const int b_row = s.n_tile * BN + (int)rank * (BN / 2);
        for (int kb = s.k0; kb < s.k1; ++kb) {
          mbar_wait(empty_bar + 8 * stage, phase ^ 1);
          const uint32_t full_leader = mapa_shared(full_bar + 8 * stage, 0);
          if (leader) mbar_expect_tx(full_bar + 8 * stage, 2 * kPairStageBytes);
          tma_load_2d_pair(smem_a + stage * kPairABytes, &tmA, full_leader, kb * BK, a_row);
          tma_load_2d_pair(smem_b + stage * kPairBBytes, &tmB, full_leader, kb * BK, b_row);
          if (++stage == kPairStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
      trace_stamp(g, 2);   // last TMA issued
    }
  } else if (warp == 1) {
    if (lane == 0 && leader) {
      // ===== MMA issuer (leader only): M = 256 over both CTAs, N = 256, K = 16 x 4 per stage
      constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)((2 * BM) >> 4) << 24);
      PairIter it(g, num_pair_tiles);
      PairSegment s;
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      while (it.next(s)) {
        mbar_wait(tempty_bar + 8 * acc, acc_phase ^ 1);
        tcgen05_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)(acc * BN);
        for (int kb = s.k0; kb < s.k1; ++kb) {
          mbar_wait(full_bar + 8 * stage, phase);
          tcgen05_fence_after();
          const uint64_t adesc = make_smem_desc(smem_a + stage * kPairABytes);
          const uint64_t bdesc = make_smem_desc(smem_b + stage * kPairBBytes);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k)
            umma_bf16_pair(tmem_d, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (kb > s.k0 || k > 0) ? 1u : 0u);
          tcgen05_commit_pair(empty_bar + 8 * stage);   // frees the slot in both CTAs once these MMAs retire
          if (++stage == kPairStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        tcgen05_commit_pair(tfull_bar + 8 * acc);        // accumulator complete in both CTAs
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
      trace_stamp(g, 4);   // last MMA issued
    }
  } else {
    // ===== epilogue warps (both CTAs): this CTA's 128 accumulator rows; same transpose as gemm_bf16_kernel
    const int quad = warp & 3;
    const int half = (warp - 2) >> 2;
    float* stg = reinterpret_cast<float*>(smem_raw + (stage_buf - raw)) + (warp - 2) * kStgFloats;
    const int q = lane >> 2, l4 = lane & 3;
    constexpr int kChunksPerWarp = BN / kChunk / 2;
    const uint32_t tempty_leader = mapa_shared(tempty_bar, 0);
    int* flag = reinterpret_cast<int*>(smem_raw + (tmem_slot - raw)) + 2;
    float* accp = reinterpret_cast<float*>(g.out);
    PairIter it(g, num_pair_tiles);
    PairSegment s;
    int acc = 0;
    uint32_t acc_phase = 0;
    while (it.next(s)) {
      mbar_wait(tfull_bar + 8 * acc, acc_phase);
      tcgen05_fence_after();
      if (threadIdx.x == 64) trace_stamp(g, 5);   // (last) accumulator ready
      const int m_tile = s.m_pair * 2 + (int)rank;
      const int row0 = m_tile * BM + quad * 32;
      const bool rows_full = row0 + 32 <= g.M;
      if (SLOTS) {
        const int cluster = blockIdx.x >> 1, nclusters = gridDim.x >> 1;
        const bool cta_valid = m_tile * BM < g.M;       // the odd last row tile of a pair may not exist (CTA-uniform)
        const int widx = warp - 2;
        if (s.k0 > 0) {
          // ---- partial: raw accumulator -> this cluster's slot
          float* slot = g.slots + ((size_t)cluster * 2 + rank) * (size_t)(BM * BN) + (size_t)(quad * 32) * BN;
          if (cta_valid) {
#pragma unroll 1
            for (int ci = 0; ci < kChunksPerWarp; ++ci) {
              const int ch = half * kChunksPerWarp + ci;
              uint32_t v[kChunk];
              __syncwarp();
              tmem_ld16(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * BN + ch * kChunk), v);
              float4* wr = reinterpret_cast<float4*>(stg + lane * kChunk);
              const int sw = (lane >> 1) & 3;
#pragma unroll
              for (int k = 0; k < 4; ++k)
                wr[k ^ sw] = make_float4(__uint_as_float(v[4 * k]), __uint_as_float(v[4 * k + 1]),
                                         __uint_as_float(v[4 * k + 2]), __uint_as_float(v[4 * k + 3]));
              __syncwarp();
#pragma unroll
              for (int rr = 0; rr < 32; rr += 8) {
                const int r = rr + q;
                const float4 x = reinterpret_cast<const float4*>(stg + r * kChunk)[l4 ^ ((r >> 1) & 3)];
                *reinterpret_cast<float4*>(slot + (size_t)r * BN + ch * kChunk + 4 * l4) = x;
              }
            }
          }
          tcgen05_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_remote(tempty_leader + 8 * acc);
          if (cta_valid) {
            __threadfence();      // every lane: its slot stores are visible GPU-wide before the flag
            __syncwarp();
            if (lane == 0) st_release_gpu(g.slot_flags + ((size_t)cluster * 2 + rank) * kEpiWarps + widx, 1);
          }
        } else {
          // ---- owner: own accumulator + the slots of the clusters that start inside this tile -> finished output
          const long long total = (long long)num_pair_tiles * g.kb_per_tile;
          const long long tile_end = (long long)(s.m_pair * g.num_n_tiles + s.n_tile + 1) * g.kb_per_tile;
          int P = 0;
          if (s.k1 < g.kb_per_tile)
            while (P < kMaxPartials && cluster + 1 + P < nclusters && total * (cluster + 1 + P) / nclusters < tile_end) ++P;
          const float* src[kMaxPartials];
#pragma unroll
          for (int p = 0; p < kMaxPartials; ++p)
            src[p] = g.slots + ((size_t)(cluster + 1 + (p < P ? p : 0)) * 2 + rank) * (size_t)(BM * BN) + (size_t)(quad * 32) * BN;
          if (cta_valid) {
#pragma unroll
            for (int p = 0; p < kMaxPartials; ++p)
              if (p < P) {
                const int* f = g.slot_flags + ((size_t)(cluster + 1 + p) * 2 + rank) * kEpiWarps + widx;
                while (ld_acquire_gpu(f) == 0) {
                }
              }
            if (threadIdx.x == 64) {
              trace_stamp(g, 8);                                   // every partial of the tile has been published
              if (g.trace) g.trace[(size_t)blockIdx.x * 16 + 11] = (unsigned long long)P;
            }
#pragma unroll 1
            for (int ci = 0; ci < kChunksPerWarp; ++ci) {
              if (threadIdx.x == 64 && ci == 1) trace_stamp(g, 9);   // first chunk merged and stored
              const int ch = half * kChunksPerWarp + ci;
              const int col0 = s.n_tile * BN + ch * kChunk;
              // the partials of this chunk first: their L2 round trips overlap the TMEM read and the transpose
              float4 pv[kMaxPartials][4];
#pragma unroll
              for (int p = 0; p < kMaxPartials; ++p)
#pragma unroll
                for (int rr = 0; rr < 4; ++rr)
                  pv[p][rr] = p < P ? __ldcg(reinterpret_cast<const float4*>(src[p] + (size_t)(rr * 8 + q) * BN + ch * kChunk + 4 * l4))
                                    : make_float4(0.f, 0.f, 0.f, 0.f);
              uint32_t v[kChunk];
              __syncwarp();
              tmem_ld16(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * BN + ch * kChunk), v);
              if (row0 >= g.M || col0 >= g.N) continue;   // warp-uniform
              float4* wr = reinterpret_cast<float4*>(stg + lane * kChunk);
              const int sw = (lane >> 1) & 3;
#pragma unroll
              for (int k = 0; k < 4; ++k)
                wr[k ^ sw] = make_float4(__uint_as_float(v[4 * k]), __uint_as_float(v[4 * k + 1]),
                                         __uint_as_float(v[4 * k + 2]), __uint_as_float(v[4 * k + 3]));
              __syncwarp();
              const int c = col0 + 4 * l4;
              float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
              if (g.bias && c + 3 < g.N) bv = __ldg(reinterpret_cast<const float4*>(g.bias + c));
#pragma unroll
              for (int rr = 0; rr < 4; ++rr) {
                const int r = rr * 8 + q;
                float4 x = reinterpret_cast<const float4*>(stg + r * kChunk)[l4 ^ ((r >> 1) & 3)];
#pragma unroll
                for (int p = 0; p < kMaxPartials; ++p) {   // fixed order: bit-reproducible
                  x.x += pv[p][rr].x; x.y += pv[p][rr].y; x.z += pv[p][rr].z; x.w += pv[p][rr].w;
                }
                x.x += bv.x; x.y += bv.y; x.z += bv.z; x.w += bv.w;
                if (g.relu) {   // torch.relu keeps NaN
                  x.x = x.x < 0.f ? 0.f : x.x; x.y = x.y < 0.f ? 0.f : x.y;
                  x.z = x.z < 0.f ? 0.f : x.z; x.w = x.w < 0.f ? 0.f : x.w;
                }
                if (g.round_tf32) x = round_tf32(x);
                if (row0 + r >= g.M || c + 3 >= g.N) continue;     // N % 4 == 0 (checked by the launcher)
                const long long off = (long long)(row0 + r) * g.fin_ldo + c;
                if (g.fin_bf16) {
                  *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(g.fin_out) + off) = pack16(x, g.f16);
                } else {
                  *reinterpret_cast<float4*>(reinterpret_cast<float*>(g.fin_out) + off) = x;
                }
              }
            }
            if (threadIdx.x == 64) trace_stamp(g, 10);              // last chunk merged and stored
            __syncwarp();
            if (lane == 0)   // consumed: hand the flags back cleared for the next launch
              for (int p = 0; p < P; ++p) g.slot_flags[((size_t)(cluster + 1 + p) * 2 + rank) * kEpiWarps + widx] = 0;
          }
          tcgen05_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_remote(tempty_leader + 8 * acc);
        }
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
        if (threadIdx.x == 64) trace_stamp(g, 7);
        continue;
      }
#pragma unroll 1
      for (int ci = 0; ci < kChunksPerWarp; ++ci) {
        const int ch = half * kChunksPerWarp + ci;
        uint32_t v[kChunk];
        __syncwarp();
        tmem_ld16(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * BN + ch * kChunk), v);
        const int col0 = s.n_tile * BN + ch * kChunk;
        if (row0 >= g.M || col0 >= g.N) continue;   // warp-uniform
        {
          float4* wr = reinterpret_cast<float4*>(stg + lane * kChunk);
          const int sw = (lane >> 1) & 3;
#pragma unroll
          for (int k = 0; k < 4; ++k)
            wr[k ^ sw] = make_float4(__uint_as_float(v[4 * k]), __uint_as_float(v[4 * k + 1]),
                                     __uint_as_float(v[4 * k + 2]), __uint_as_float(v[4 * k + 3]));
        }
        __syncwarp();
        const int c = col0 + 4 * l4;
        const bool full = rows_full && col0 + kChunk <= g.N;
#pragma unroll
        for (int rr = 0; rr < 32; rr += 8) {
          const int r = rr + q;
          const float4 x = reinterpret_cast<const float4*>(stg + r * kChunk)[l4 ^ ((r >> 1) & 3)];
          const long long off = (long long)(row0 + r) * g.ldo + c;
          if (full) {
            red_add_v4(accp + off, x.x, x.y, x.z, x.w);
          } else if (row0 + r < g.M) {
            const float xs[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
            for (int e = 0; e < 4; ++e)
              if (c + e < g.N) atomicAdd(accp + off + e, xs[e]);
          }
        }
      }
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_remote(tempty_leader + 8 * acc);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
      if (threadIdx.x == 64) trace_stamp(g, 7);   // (last) partial tile added to the accumulator
      if (g.tile_counters && !g.grid_finish && m_tile * BM < g.M) {   // CTA-uniform; the odd last row tile of a pair may not exist
        // ---- fused finish of this CTA's 128-row tile (see gemm_bf16_kernel)
        __threadfence();
        asm volatile("bar.sync 2, %0;" ::"n"(kEpiWarps * 32) : "memory");
        const int tile = m_tile * g.num_n_tiles + s.n_tile;
        if (threadIdx.x == 64) {
          const int done = atomicAdd(g.tile_counters + tile, s.k1 - s.k0) + (s.k1 - s.k0);
          *flag = done == g.kb_per_tile;
          if (done == g.kb_per_tile) g.tile_counters[tile] = 0;
        }
        asm volatile("bar.sync 2, %0;" ::"n"(kEpiWarps * 32) : "memory");
        if (*flag && row0 < g.M) {
          __threadfence();
          const int colw = BN / 2;
          const int cbase = s.n_tile * BN + half * colw;
          constexpr int kRows = 16;
          for (int c = cbase + 4 * lane; c < cbase + colw && c < g.N; c += 128) {
            float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
            if (g.bias) bv = __ldg(reinterpret_cast<const float4*>(g.bias + c));
            for (int rb = 0; rb < 32; rb += kRows) {
              float4 x[kRows];
#pragma unroll
              for (int r = 0; r < kRows; ++r) {
                const int row = row0 + rb + r;
                x[r] = row < g.M ? __ldcg(reinterpret_cast<const float4*>(accp + (long long)row * g.ldo + c))
                                 : make_float4(0.f, 0.f, 0.f, 0.f);
              }
#pragma unroll
              for (int r = 0; r < kRows; ++r) {
                const int row = row0 + rb + r;
                if (row >= g.M) continue;
                __stcg(reinterpret_cast<float4*>(accp + (long long)row * g.ldo + c), make_float4(0.f, 0.f, 0.f, 0.f));
                float4 w = x[r];
                w.x += bv.x; w.y += bv.y; w.z += bv.z; w.w += bv.w;
                if (g.relu) {
                  w.x = w.x < 0.f ? 0.f : w.x; w.y = w.y < 0.f ? 0.f : w.y;
                  w.z = w.z < 0.f ? 0.f : w.z; w.w = w.w < 0.f ? 0.f : w.w;
                }
                if (g.round_tf32) w = round_tf32(w);
                const long long off = (long long)row * g.fin_ldo + c;
                if (g.fin_bf16) {
                  *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(g.fin_out) + off) = pack16(w, g.f16);
                } else {
                  *reinterpret_cast<float4*>(reinterpret_cast<float*>(g.fin_out) + off) = w;
                }
              }
            }
          }
        }
      }
    }
    if (!SLOTS && g.tile_counters && g.grid_finish) {
      // ---- finish by everybody.  With per-tile counters the 52 CTAs that happen to complete a tile each read 128 KB
      // back behind the other CTAs' atomics (5-8 us at the very end of the kernel, 96 SMs idle).  Here every CTA waits
      // until ALL partials have been added (the grid is one wave: every CTA is resident, so the barrier cannot
      // deadlock) and then finishes 1/grid of the output: 44 KB per CTA, every load in flight at once.
      int* cnt = g.tile_counters;        // [0] arrivals, [1] departures; zero between launches
      __threadfence();
      asm volatile("bar.sync 2, %0;" ::"n"(kEpiWarps * 32) : "memory");
      if (threadIdx.x == 64) {
        atomicAdd(cnt, 1);
        while (ld_acquire_gpu(cnt) < (int)gridDim.x) {
        }
      }
      asm volatile("bar.sync 2, %0;" ::"n"(kEpiWarps * 32) : "memory");
      if (threadIdx.x == 64) trace_stamp(g, 8);     // every partial of every tile has been added
      const int n4 = g.N >> 2;
      const long long total4 = (long long)g.M * n4;
      const int et = threadIdx.x - 64;              // 0 .. 255
      constexpr int kInFlight = 4;
      for (long long i0 = (long long)blockIdx.x * (kEpiWarps * 32) + et; i0 < total4;
           i0 += (long long)gridDim.x * (kEpiWarps * 32) * kInFlight) {
        float4 x[kInFlight];
        long long idx[kInFlight];
#pragma unroll
        for (int u = 0; u < kInFlight; ++u) {
          idx[u] = i0 + (long long)u * gridDim.x * (kEpiWarps * 32);
          if (idx[u] < total4) {
            const int row = (int)(idx[u] / n4), c = (int)(idx[u] - (long long)row * n4) * 4;
            x[u] = __ldcg(reinterpret_cast<const float4*>(accp + (long long)row * g.ldo + c));
          }
        }
#pragma unroll
        for (int u = 0; u < kInFlight; ++u) {
          if (idx[u] >= total4) continue;
          const int row = (int)(idx[u] / n4), c = (int)(idx[u] - (long long)row * n4) * 4;
          __stcg(reinterpret_cast<float4*>(accp + (long long)row * g.ldo + c), make_float4(0.f, 0.f, 0.f, 0.f));
          float4 w = x[u];
          if (g.bias) {
            const float4 bv = __ldg(reinterpret_cast<const float4*>(g.bias + c));
            w.x += bv.x; w.y += bv.y; w.z += bv.z; w.w += bv.w;
          }
          if (g.relu) {   // torch.relu keeps NaN
            w.x = w.x < 0.f ? 0.f : w.x; w.y = w.y < 0.f ? 0.f : w.y;
            w.z = w.z < 0.f ? 0.f : w.z; w.w = w.w < 0.f ? 0.f : w.w;
          }
          if (g.round_tf32) w = round_tf32(w);
          const long long off = (long long)row * g.fin_ldo + c;
          if (g.fin_bf16) {
            *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(g.fin_out) + off) = pack16(w, g.f16);
          } else {
            *reinterpret_cast<float4*>(reinterpret_cast<float*>(g.fin_out) + off) = w;
          }
        }
      }
      asm volatile("bar.sync 2, %0;" ::"n"(kEpiWarps * 32) : "memory");
      if (threadIdx.x == 64 && atomicAdd(cnt + 1, 1) == (int)gridDim.x - 1) {   // the last one out clears the barrier
        cnt[0] = 0;
        cnt[1] = 0;
      }
    }
  }

  if (threadIdx.x == 64) trace_stamp(g, 6);   // epilogue (and finish) done
  tcgen05_fence_before();
  cluster_sync_all();        // the leader's MMAs write the peer's TMEM: neither CTA may release before both are done
  if (warp == 2) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)(2 * BN)) : "memory");
  }
}

// ------------------------------------------------------------------ small elementwise kernels
__global__ void bias_act_kernel(const float* in, long long ldi, const float* __restrict__ bias, int relu,
                                int M, int N, void* out, int out_bf16, long long ldo, int clear_in, int tf32, int f16) {
  // one thread per 4 consecutive columns (N % 4 == 0 is required by the launcher)
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int n4 = N >> 2;
  pdl_trigger();
  pdl_wait();
  if (idx >= (long long)M * n4) return;
  const int m = (int)(idx / n4), j = (int)(idx - (long long)m * n4) * 4;
  float4 v = *reinterpret_cast<const float4*>(in + (long long)m * ldi + j);
  // split-K accumulator: hand it back zeroed so that the next layer / call needs no memset launch
  if (clear_in) *reinterpret_cast<float4*>(const_cast<float*>(in) + (long long)m * ldi + j) = make_float4(0.f, 0.f, 0.f, 0.f);
  if (bias) {
    const float4 b = *reinterpret_cast<const float4*>(bias + j);
    v.x += b.x; v.y += b.y; v.z += b.z; v.w += b.w;
  }
  if (relu) {   // torch.relu keeps NaN
    v.x = v.x < 0.f ? 0.f : v.x; v.y = v.y < 0.f ? 0.f : v.y; v.z = v.z < 0.f ? 0.f : v.z; v.w = v.w < 0.f ? 0.f : v.w;
  }
  if (tf32) v = round_tf32(v);
  if (out_bf16) {
    *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(out) + (long long)m * ldo + j) = pack16(v, f16);
  } else {
    *reinterpret_cast<float4*>(reinterpret_cast<float*>(out) + (long long)m * ldo + j) = v;
  }
}

__global__ void cast_bf16_kernel(const float* __restrict__ in, long long ld_in, int rows, int cols,
                                 __nv_bfloat16* __restrict__ out, long long ld_out) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)rows * ld_out) return;
  const int r = (int)(idx / ld_out), c = (int)(idx - (long long)r * ld_out);
  out[idx] = __float2bfloat16_rn(c < cols ? in[(long long)r * ld_in + c] : 0.f);
}

// x (fp32) -> fp16 (saturated), columns >= cols zero-filled up to ld_out: the weight operand of a kind::f16 tail layer
__global__ void cast_f16_kernel(const float* __restrict__ in, long long ld_in, int rows, int cols,
                                __half* __restrict__ out, long long ld_out) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)rows * ld_out) return;
  const int r = (int)(idx / ld_out), c = (int)(idx - (long long)r * ld_out);
  out[idx] = __float2half_rn(c < cols ? sat_f16(in[(long long)r * ld_in + c]) : 0.f);
}

// W (fp32) -> fp32 rounded to TF32, columns >= cols zero-filled up to ld_out: the weight operand of a kind::tf32 layer
__global__ void round_tf32_kernel(const float* __restrict__ in, long long ld_in, int rows, int cols,
                                  float* __restrict__ out, long long ld_out) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)rows * ld_out) return;
  const int r = (int)(idx / ld_out), c = (int)(idx - (long long)r * ld_out);
  out[idx] = c < cols ? round_tf32(in[(long long)r * ld_in + c]) : 0.f;
}

// ------------------------------------------------------------------ host side
// bf16 [rows, cols] row-major with leading dimension ld -> 2-D map, box {64 cols, box_rows}, 128-byte swizzle,
// out-of-bounds elements read as zero (M, N and K tails).
int make_map(CUtensorMap* map, const void* ptr, int64_t rows, int64_t cols, int64_t ld, int box_rows) {
  EncodeTiledFn enc = tensormap_encode_fn();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled not available from the driver");
    return OVM_ECUDA;
  }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%lld cols=%lld ld=%lld)", (int)r, (long long)rows,
              (long long)cols, (long long)ld);
    return OVM_ECUDA;
  }
  return OVM_OK;
}

template <int BN, int MODE>
int launch_gemm_mode(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmArgs& g, int grid, cudaStream_t stream) {
  static SmemOptIn opt;
  {
    int rc = opt.ensure(gemm_bf16_kernel<BN, MODE>, GemmShape<BN>::kSmemBytes);
    if (rc != OVM_OK) return rc;
  }
  OVM_CUDA(launch_pdl(gemm_bf16_kernel<BN, MODE>, dim3(grid), dim3(kGemmThreads), GemmShape<BN>::kSmemBytes, stream, tmA,
                      tmB, g));
  OVM_LAUNCHED("gemm_bf16_kernel");
  return OVM_OK;
}

template <int BN>
int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmArgs& g, int grid, cudaStream_t stream) {
  if (g.streamk) return launch_gemm_mode<BN, kModeStreamK>(tmA, tmB, g, grid, stream);
  if (g.out_bf16) return launch_gemm_mode<BN, kModeBF16>(tmA, tmB, g, grid, stream);
  return launch_gemm_mode<BN, kModeF32>(tmA, tmB, g, grid, stream);
}

}  // namespace

int gemm_bf16_launch(const void* A, int64_t lda, const void* B, int64_t ldb, int32_t M, int32_t N, int32_t K,
                     const float* bias, int32_t relu, void* out, int32_t out_dtype, int64_t ldo, int32_t mode,
                     int32_t* mode_used, cudaStream_t stream, const GemmFinish* fin) {
  OVM_REQUIRE(M >= 0 && N > 0 && K > 0, "ovm_gemm_bf16: bad shape M=%d N=%d K=%d", M, N, K);
  if (M == 0) return OVM_OK;
  OVM_REQUIRE(A && B && out, "ovm_gemm_bf16: NULL pointer");
  OVM_REQUIRE(lda >= K && ldb >= K && lda % 8 == 0 && ldb % 8 == 0, "ovm_gemm_bf16: lda/ldb must be >= K and multiples of 8");
  OVM_REQUIRE(!(reinterpret_cast<uintptr_t>(A) & 15) && !(reinterpret_cast<uintptr_t>(B) & 15),
              "ovm_gemm_bf16: A and B must be 16-byte aligned");
  OVM_REQUIRE(out_dtype == OVM_F32 || out_dtype == OVM_BF16 || out_dtype == OVM_TF32 || out_dtype == OVM_F16,
              "ovm_gemm_bf16: out_dtype %d", out_dtype);
  OVM_REQUIRE(ldo >= N && ldo % 8 == 0 && !(reinterpret_cast<uintptr_t>(out) & 15),
              "ovm_gemm_bf16: ldo must be >= N, a multiple of 8, and out 16-byte aligned");
  OVM_REQUIRE(mode >= OVM_GEMM_AUTO && mode <= OVM_GEMM_STREAMK, "ovm_gemm_bf16: mode %d", mode);

  // Tile width: the widest of 256/128/64 that still yields enough tiles; small N uses 64 (B rows past N are
  // zero-filled by TMA and masked in the epilogue).
  int BN = N >= 256 ? 256 : (N >= 128 ? 128 : 64);
  GemmArgs g{};
  g.M = M; g.N = N; g.K = K;
  const int num_m_tiles = (M + BM - 1) / BM;

  g.num_n_tiles = (N + BN - 1) / BN;
  g.num_tiles = num_m_tiles * g.num_n_tiles;
  g.kb_per_tile = (K + BK - 1) / BK;
  g.total_kb = (long long)g.num_tiles * g.kb_per_tile;
  if (mode == OVM_GEMM_AUTO) {
    // K split pays when the tiles cannot fill the SMs (or leave a ragged second wave) and K is long enough
    // that the extra fp32 read-modify-write of the output is small next to the operand traffic.
    const bool ragged = g.num_tiles < kNumSMs || (g.num_tiles % kNumSMs) * 2 < kNumSMs;
    mode = (out_dtype == OVM_F32 && g.kb_per_tile >= 32 && g.num_tiles < 2 * kNumSMs && ragged && !bias && !relu)
               ? OVM_GEMM_STREAMK : OVM_GEMM_DIRECT;
  }
  if (mode == OVM_GEMM_DIRECT) {
    // a grid that does not fill the SMs is bound by the per-CTA critical path (prologue + K loop + epilogue of one
    // tile): narrower tiles shorten it as long as they still fit one wave (fc2: 52 tiles of 256 -> 104 of 128)
    while (BN > 64 && num_m_tiles * ((N + BN / 2 - 1) / (BN / 2)) <= kNumSMs) BN /= 2;
    g.num_n_tiles = (N + BN - 1) / BN;
    g.num_tiles = num_m_tiles * g.num_n_tiles;
    g.total_kb = (long long)g.num_tiles * g.kb_per_tile;
  }
  if (mode == OVM_GEMM_STREAMK) {
    OVM_REQUIRE(out_dtype == OVM_F32 && !bias && !relu,
                "ovm_gemm_bf16: STREAMK accumulates raw fp32 partial sums (no bias/relu, fp32 out)");
  }
  if (mode_used) *mode_used = mode;
  g.streamk = mode == OVM_GEMM_STREAMK;
  g.relu = relu;
  g.out_bf16 = out_dtype == OVM_BF16 || out_dtype == OVM_F16;     // 16-bit store path
  g.f16 = out_dtype == OVM_F16;
  g.round_tf32 = out_dtype == OVM_TF32;
  g.bias = bias;
  g.out = out;
  g.ldo = ldo;
  g.trace = g_gemm_trace;
  if (fin) {
    OVM_REQUIRE(g.streamk && fin->tile_counters && fin->out && fin->num_counters >= g.num_tiles && N % 4 == 0 &&
                    fin->ldo >= N && fin->ldo % 4 == 0 && !(reinterpret_cast<uintptr_t>(fin->out) & 15),
                "ovm_gemm_bf16: fused split-K finish needs STREAMK, %d tile counters and an aligned output", g.num_tiles);
    g.tile_counters = fin->defer ? nullptr : fin->tile_counters;
    g.bias = fin->bias;
    g.relu = fin->relu;
    g.fin_out = fin->out;
    g.fin_ldo = fin->ldo;
    g.fin_bf16 = fin->out_dtype == OVM_BF16 || fin->out_dtype == OVM_F16;
    g.f16 = fin->out_dtype == OVM_F16;
    g.round_tf32 = fin->out_dtype == OVM_TF32;
  }
  CUtensorMap tmA, tmB;
  int rc = make_map(&tmA, A, M, K, lda, BM);
  if (rc != OVM_OK) return rc;
  static const bool pair_ok = [] { const char* v = getenv("OVM_GEMM_PAIR"); return !v || atoi(v) != 0; }();
  if (fin && BN == kPairBN && pair_ok && num_m_tiles >= 2) {
    // CTA pairs on 256 x 256 tiles: a third less operand traffic out of L2 (see gemm_pair_kernel)
    rc = make_map(&tmB, B, N, K, ldb, kPairBN / 2);
    if (rc != OVM_OK) return rc;
    const int num_pair_tiles = ((num_m_tiles + 1) / 2) * g.num_n_tiles;
    const long long total = (long long)num_pair_tiles * g.kb_per_tile;
    const int clusters = (int)(total < kNumSMs / 2 ? total : kNumSMs / 2);
    // deterministic slot exchange when the workspace has the slots and no tile is shared by more than kMaxPartials + 1 clusters
    // Opt-in (OVM_GEMM_SLOTS=1): measured at C2, the slot exchange costs 56.4 us against 51.6 us for red.global.add --
    // the owners' merge reads (2.6 partial tiles of 128 KB per CTA, ~1 us per L2 round trip on lines other SMs have just
    // written, 12 loads in flight per lane) are slower than the L2's atomic units; prefetching the next chunk's partials
    // needs registers the epilogue does not have (spills: 69 us).  What it buys is a bit-reproducible layer.
    static const bool slots_ok = [] { const char* v = getenv("OVM_GEMM_SLOTS"); return v && atoi(v) != 0; }();
    const long long starts_per_tile = ((long long)g.kb_per_tile * clusters + total - 1) / total;   // cluster ranges starting in one tile, at most
    const bool use_slots = !fin->defer && slots_ok && fin->slots && fin->slot_flags && starts_per_tile <= kMaxPartials &&
                           fin->slots_len >= (long long)clusters * 2 * BM * kPairBN && fin->flags_len >= clusters * 2 * kEpiWarps &&
                           !(reinterpret_cast<uintptr_t>(fin->slots) & 15);
    // Opt-in (OVM_GEMM_GRIDFIN=1).  Measured at C2: the fc1 stage alone 53.4 -> 51.7 us, but the step 0.168 -> 0.172 ms:
    // with every fc1 CTA held to the end, the next kernel's CTAs (resident 13 us early on the SMs of fc1 CTAs that finish
    // first, staging their head weights) lose their head start.
    static const bool gridfin = [] { const char* v = getenv("OVM_GEMM_GRIDFIN"); return v && atoi(v) != 0; }();
    g.grid_finish = !fin->defer && gridfin && !use_slots && fin->num_counters >= 2 && 2 * clusters <= kNumSMs;
    if (use_slots) {
      g.slots = fin->slots;
      g.slot_flags = fin->slot_flags;
      static SmemOptIn opt;
      rc = opt.ensure(gemm_pair_kernel<true>, kPairSmemBytes);
      if (rc != OVM_OK) return rc;
      OVM_CUDA(launch_pdl(gemm_pair_kernel<true>, dim3(2 * clusters), dim3(kGemmThreads), kPairSmemBytes, stream, tmA, tmB, g,
                          num_pair_tiles));
    } else {
      static SmemOptIn opt;
      rc = opt.ensure(gemm_pair_kernel<false>, kPairSmemBytes);
      if (rc != OVM_OK) return rc;
      OVM_CUDA(launch_pdl(gemm_pair_kernel<false>, dim3(2 * clusters), dim3(kGemmThreads), kPairSmemBytes, stream, tmA, tmB, g,
                          num_pair_tiles));
    }
    OVM_LAUNCHED("gemm_pair_kernel");
    return OVM_OK;
  }
  rc = make_map(&tmB, B, N, K, ldb, BN);
  if (rc != OVM_OK) return rc;
  int grid = g.num_tiles;
  if (g.streamk) grid = (int)(g.total_kb < kNumSMs ? g.total_kb : kNumSMs);
  switch (BN) {
    case 256: return launch_gemm<256>(tmA, tmB, g, grid, stream);
    case 128: return launch_gemm<128>(tmA, tmB, g, grid, stream);
    default: return launch_gemm<64>(tmA, tmB, g, grid, stream);
  }
}

void gemm_set_trace(unsigned long long* buf) { g_gemm_trace = buf; }

int bias_act_launch(const float* in, int64_t ldi, const float* bias, int32_t relu, int32_t M, int32_t N, void* out,
                    int32_t out_dtype, int64_t ldo, cudaStream_t stream, bool clear_in) {
  if (M == 0) return OVM_OK;
  OVM_REQUIRE(in && out && M > 0 && N > 0, "ovm_bias_act: bad arguments");
  OVM_REQUIRE(N % 4 == 0 && ldi % 4 == 0 && ldo % 4 == 0, "ovm_bias_act: N, ldi, ldo must be multiples of 4");
  OVM_REQUIRE(out_dtype == OVM_F32 || out_dtype == OVM_BF16 || out_dtype == OVM_TF32 || out_dtype == OVM_F16,
              "ovm_bias_act: out_dtype %d", out_dtype);
  const long long total = (long long)M * (N / 4);
  OVM_CUDA(launch_pdl(bias_act_kernel, dim3((unsigned)((total + 255) / 256)), dim3(256), 0, stream, in, (long long)ldi, bias,
                      (int)relu, (int)M, (int)N, out, (int)(out_dtype == OVM_BF16 || out_dtype == OVM_F16), (long long)ldo,
                      clear_in ? 1 : 0, (int)(out_dtype == OVM_TF32), (int)(out_dtype == OVM_F16)));
  OVM_LAUNCHED("bias_act_kernel");
  return OVM_OK;
}

int cast_bf16_launch(const float* in, int64_t ld_in, int32_t rows, int32_t cols, void* out, int64_t ld_out,
                     cudaStream_t stream) {
  if (rows == 0) return OVM_OK;
  OVM_REQUIRE(in && out && rows > 0 && cols > 0 && ld_in >= cols && ld_out >= cols, "ovm_cast_bf16: bad arguments");
  const long long total = (long long)rows * ld_out;
  cast_bf16_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(in, ld_in, rows, cols,
                                                                       reinterpret_cast<__nv_bfloat16*>(out), ld_out);
  OVM_LAUNCHED("cast_bf16_kernel");
  return OVM_OK;
}

int cast_f16_launch(const float* in, int64_t ld_in, int32_t rows, int32_t cols, void* out, int64_t ld_out, cudaStream_t stream) {
  if (rows == 0) return OVM_OK;
  OVM_REQUIRE(in && out && rows > 0 && cols > 0 && ld_in >= cols && ld_out >= cols, "ovm_cast_f16: bad arguments");
  const long long total = (long long)rows * ld_out;
  cast_f16_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(in, ld_in, rows, cols, reinterpret_cast<__half*>(out), ld_out);
  OVM_LAUNCHED("cast_f16_kernel");
  return OVM_OK;
}

int round_tf32_launch(const float* in, int64_t ld_in, int32_t rows, int32_t cols, float* out, int64_t ld_out,
                      cudaStream_t stream) {
  if (rows == 0) return OVM_OK;
  OVM_REQUIRE(in && out && rows > 0 && cols > 0 && ld_in >= cols && ld_out >= cols, "ovm_round_tf32: bad arguments");
  const long long total = (long long)rows * ld_out;
  round_tf32_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(in, ld_in, rows, cols, out, ld_out);
  OVM_LAUNCHED("round_tf32_kernel");
  return OVM_OK;
}

}  // namespace ovm
