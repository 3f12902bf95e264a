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

#include <cstring>

#include "common.cuh"

namespace ovm {

namespace {

constexpr int BM = 128;
constexpr int BK = 64;            // 64 bf16 = 128 bytes = one swizzle row
constexpr int UMMA_K = 16;
constexpr int kGemmThreads = 192;

template <int BN>
struct GemmShape {
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBBytes = BN * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = BN == 256 ? 4 : (BN == 128 ? 6 : 8);
  static constexpr int kTmemCols = BN * 2 < 32 ? 32 : BN * 2;   // two accumulator stages
  static constexpr int kBarrierBytes = (2 * kStages + 4) * 8 + 16;
  static constexpr size_t kSmemBytes = 1024 + (size_t)kStages * kStageBytes + kBarrierBytes;
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
};

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

template <int BN>
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
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

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
      mbar_init(tempty_bar + 8 * s, 4);   // one arrive per epilogue warp
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
      while (it.next(s)) {
        mbar_wait(tempty_bar + 8 * acc, acc_phase ^ 1);
        tcgen05_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)(acc * BN);
        for (int kb = s.k0; kb < s.k1; ++kb) {
          mbar_wait(full_bar + 8 * stage, phase);
          tcgen05_fence_after();
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
    }
  } else {
    // ===== epilogue warps: TMEM lane quadrant = warp % 4
    const int quad = warp & 3;
    WorkIter it(g);
    Segment s;
    int acc = 0;
    uint32_t acc_phase = 0;
    while (it.next(s)) {
      mbar_wait(tfull_bar + 8 * acc, acc_phase);
      tcgen05_fence_after();
      const int row = s.m_tile * BM + quad * 32 + lane;
      const bool row_ok = row < g.M;
#pragma unroll 1
      for (int ch = 0; ch < BN / 32; ++ch) {
        uint32_t v[32];
        __syncwarp();   // tcgen05.ld is .sync.aligned: reconverge after the predicated stores below
        tmem_ld32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * BN + ch * 32), v);
        const int col0 = s.n_tile * BN + ch * 32;
        if (!row_ok || col0 >= g.N) continue;
        const bool full = col0 + 32 <= g.N;
        if (g.streamk) {
          float* dst = reinterpret_cast<float*>(g.out) + (long long)row * g.ldo + col0;
          if (full) {
#pragma unroll
            for (int j = 0; j < 8; ++j)
              red_add_v4(dst + 4 * j, __uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]),
                         __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (col0 + j < g.N) atomicAdd(dst + j, __uint_as_float(v[j]));
          }
        } else {
          float f[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            float x = __uint_as_float(v[j]);
            if (g.bias && (full || col0 + j < g.N)) x += __ldg(g.bias + col0 + j);
            f[j] = g.relu ? fmaxf(x, 0.f) : x;
          }
          if (g.out_bf16) {
            __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(g.out) + (long long)row * g.ldo + col0;
            if (full) {
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                uint4 pk;
                __nv_bfloat162 t0 = __floats2bfloat162_rn(f[8 * j + 0], f[8 * j + 1]);
                __nv_bfloat162 t1 = __floats2bfloat162_rn(f[8 * j + 2], f[8 * j + 3]);
                __nv_bfloat162 t2 = __floats2bfloat162_rn(f[8 * j + 4], f[8 * j + 5]);
                __nv_bfloat162 t3 = __floats2bfloat162_rn(f[8 * j + 6], f[8 * j + 7]);
                pk.x = *reinterpret_cast<uint32_t*>(&t0);
                pk.y = *reinterpret_cast<uint32_t*>(&t1);
                pk.z = *reinterpret_cast<uint32_t*>(&t2);
                pk.w = *reinterpret_cast<uint32_t*>(&t3);
                *reinterpret_cast<uint4*>(dst + 8 * j) = pk;
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (col0 + j < g.N) dst[j] = __float2bfloat16_rn(f[j]);
            }
          } else {
            float* dst = reinterpret_cast<float*>(g.out) + (long long)row * g.ldo + col0;
            if (full) {
#pragma unroll
              for (int j = 0; j < 8; ++j)
                *reinterpret_cast<float4*>(dst + 4 * j) = make_float4(f[4 * j], f[4 * j + 1], f[4 * j + 2], f[4 * j + 3]);
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (col0 + j < g.N) dst[j] = f[j];
            }
          }
        }
      }
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar + 8 * acc);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 2) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)Sh::kTmemCols) : "memory");
  }
}

// ------------------------------------------------------------------ small elementwise kernels
__global__ void bias_act_kernel(const float* __restrict__ in, long long ldi, const float* __restrict__ bias, int relu,
                                int M, int N, void* out, int out_bf16, long long ldo) {
  // one thread per 4 consecutive columns (N % 4 == 0 is required by the launcher)
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int n4 = N >> 2;
  if (idx >= (long long)M * n4) return;
  const int m = (int)(idx / n4), j = (int)(idx - (long long)m * n4) * 4;
  float4 v = *reinterpret_cast<const float4*>(in + (long long)m * ldi + j);
  if (bias) {
    const float4 b = *reinterpret_cast<const float4*>(bias + j);
    v.x += b.x; v.y += b.y; v.z += b.z; v.w += b.w;
  }
  if (relu) {
    v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f);
  }
  if (out_bf16) {
    __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
    uint2 pk;
    pk.x = *reinterpret_cast<uint32_t*>(&a);
    pk.y = *reinterpret_cast<uint32_t*>(&b);
    *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(out) + (long long)m * ldo + j) = pk;
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

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
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

// bf16 [rows, cols] row-major with leading dimension ld -> 2-D map, box {64 cols, box_rows}, 128-byte swizzle,
// out-of-bounds elements read as zero (M, N and K tails).
int make_map(CUtensorMap* map, const void* ptr, int64_t rows, int64_t cols, int64_t ld, int box_rows) {
  EncodeTiledFn enc = get_encode_fn();
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

template <int BN>
int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmArgs& g, int grid, cudaStream_t stream) {
  static bool attr_done = false;
  if (!attr_done) {
    OVM_CUDA(cudaFuncSetAttribute(gemm_bf16_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)GemmShape<BN>::kSmemBytes));
    attr_done = true;
  }
  gemm_bf16_kernel<BN><<<grid, kGemmThreads, GemmShape<BN>::kSmemBytes, stream>>>(tmA, tmB, g);
  OVM_LAUNCHED("gemm_bf16_kernel");
  return OVM_OK;
}

}  // namespace

int gemm_bf16_launch(const void* A, int64_t lda, const void* B, int64_t ldb, int32_t M, int32_t N, int32_t K,
                     const float* bias, int32_t relu, void* out, int32_t out_dtype, int64_t ldo, int32_t mode,
                     int32_t* mode_used, cudaStream_t stream) {
  OVM_REQUIRE(M >= 0 && N > 0 && K > 0, "ovm_gemm_bf16: bad shape M=%d N=%d K=%d", M, N, K);
  if (M == 0) return OVM_OK;
  OVM_REQUIRE(A && B && out, "ovm_gemm_bf16: NULL pointer");
  OVM_REQUIRE(lda >= K && ldb >= K && lda % 8 == 0 && ldb % 8 == 0, "ovm_gemm_bf16: lda/ldb must be >= K and multiples of 8");
  OVM_REQUIRE(!(reinterpret_cast<uintptr_t>(A) & 15) && !(reinterpret_cast<uintptr_t>(B) & 15),
              "ovm_gemm_bf16: A and B must be 16-byte aligned");
  OVM_REQUIRE(out_dtype == OVM_F32 || out_dtype == OVM_BF16, "ovm_gemm_bf16: out_dtype %d", out_dtype);
  OVM_REQUIRE(ldo >= N && ldo % 8 == 0 && !(reinterpret_cast<uintptr_t>(out) & 15),
              "ovm_gemm_bf16: ldo must be >= N, a multiple of 8, and out 16-byte aligned");
  OVM_REQUIRE(mode >= OVM_GEMM_AUTO && mode <= OVM_GEMM_STREAMK, "ovm_gemm_bf16: mode %d", mode);

  // Tile width: the widest of 256/128/64 that still yields enough tiles; small N uses 64 (B rows past N are
  // zero-filled by TMA and masked in the epilogue).
  const int BN = N >= 256 ? 256 : (N >= 128 ? 128 : 64);
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
  if (mode == OVM_GEMM_STREAMK) {
    OVM_REQUIRE(out_dtype == OVM_F32 && !bias && !relu,
                "ovm_gemm_bf16: STREAMK accumulates raw fp32 partial sums (no bias/relu, fp32 out)");
  }
  if (mode_used) *mode_used = mode;
  g.streamk = mode == OVM_GEMM_STREAMK;
  g.relu = relu;
  g.out_bf16 = out_dtype == OVM_BF16;
  g.bias = bias;
  g.out = out;
  g.ldo = ldo;
  CUtensorMap tmA, tmB;
  int rc = make_map(&tmA, A, M, K, lda, BM);
  if (rc != OVM_OK) return rc;
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

int bias_act_launch(const float* in, int64_t ldi, const float* bias, int32_t relu, int32_t M, int32_t N, void* out,
                    int32_t out_dtype, int64_t ldo, cudaStream_t stream) {
  if (M == 0) return OVM_OK;
  OVM_REQUIRE(in && out && M > 0 && N > 0, "ovm_bias_act: bad arguments");
  OVM_REQUIRE(N % 4 == 0 && ldi % 4 == 0 && ldo % 4 == 0, "ovm_bias_act: N, ldi, ldo must be multiples of 4");
  OVM_REQUIRE(out_dtype == OVM_F32 || out_dtype == OVM_BF16, "ovm_bias_act: out_dtype %d", out_dtype);
  const long long total = (long long)M * (N / 4);
  bias_act_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(in, ldi, bias, relu, M, N, out,
                                                                     out_dtype == OVM_BF16, ldo);
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

}  // namespace ovm
