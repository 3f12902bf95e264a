// Detection post-processing either side of the lift: segmented NMS and the GroundingDINO
// post-processing that produces the lift's prompt boxes.
//
// Reference: torchvision.ops.nms as called at cubercnn/modeling/roi_heads/roi_heads_gdino.py:272 and (through d2's
// batched_nms) cubercnn/modeling/roi_heads/fast_rcnn.py:105; get_grounding_output roi_heads_gdino.py:174-202 and
// grounding_dino_inference_detector :262-281.
//
// One CTA owns one image from start to finish: score keys are sorted in shared memory (bitonic, 64-bit keys =
// descending score | ascending row, which is the stable descending order torch's sort produces), the sorted boxes
// stay in shared memory, and the greedy pass walks them 32 positions per round (three barriers).  Nothing goes back to the
// host in between (torchvision's CUDA nms copies its n x n/64 mask to the host and scans it there).
#include <cooperative_groups.h>

#include <algorithm>
#include <cstdlib>

#include "common.cuh"

namespace ovm {
namespace {
namespace cg = cooperative_groups;

constexpr int kPostThreads = 1024;
constexpr int kPostWarps = kPostThreads / 32;
constexpr uint64_t kPadKey = ~0ull;

thread_local unsigned long long* g_post_trace = nullptr;   // ovm_post_set_trace

__device__ __forceinline__ void post_stamp(unsigned long long* trace, int slot) {
  if (trace && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    trace[slot] = t;
  }
}

// Cluster barrier in its two halves and a 32-bit store into another CTA's shared memory.  (cooperative_groups'
// cluster.sync() compiles to the same barrier plus MEMBAR.ALL.GPU and CCTL.IVALL; measured ~850 cycles per round.)
__device__ __forceinline__ void cluster_arrive_release() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
}
__device__ __forceinline__ void cluster_wait_acquire() {
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void st_shared_cluster_u32(const void* local_smem, int cta_rank, uint32_t v) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(local_smem)), "r"(cta_rank));
  asm volatile("st.shared::cluster.u32 [%0], %1;" ::"r"(remote), "r"(v) : "memory");
}

// fp32 -> 32-bit key whose unsigned ascending order is the descending order of the floats (-0 == +0)
__device__ __forceinline__ uint32_t score_key_desc(float s) {
  if (s == 0.0f) s = 0.0f;
  uint32_t u = __float_as_uint(s);
  u ^= (u >> 31) ? 0xFFFFFFFFu : 0x80000000u;
  return ~u;
}

__device__ __forceinline__ int next_pow2(int n) {
  int p = 32;
  while (p < n) p <<= 1;
  return p;
}

// ascending bitonic sort of npad (power of two) 64-bit keys in shared memory; all threads of the CTA take part
__device__ void block_sort(uint64_t* keys, int npad) {
  __syncthreads();
  for (int k = 2; k <= npad; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = threadIdx.x; t < (npad >> 1); t += blockDim.x) {
        const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
        const int l = i | j;
        const uint64_t a = keys[i], b = keys[l];
        const bool up = (i & k) == 0;
        if ((a > b) == up) {
          keys[i] = b;
          keys[l] = a;
        }
      }
      __syncthreads();
    }
  }
}

// torchvision's CPU kernel (ops/cpu/nms_kernel.cpp): areas and intersections in fp32, no contraction
__device__ __forceinline__ float box_area(const float4& b) {
  return __fmul_rn(__fsub_rn(b.z, b.x), __fsub_rn(b.w, b.y));
}

struct IouThr {
  double thr;     // the reference's threshold (compared in double, as torchvision's CPU kernel does)
  float thr_f;    // (float)thr, for the one-product decision far from the threshold
  bool fast;      // thr in a range where that decision is safe
};
__device__ __forceinline__ IouThr make_thr(double thr) {
  IouThr t;
  t.thr = thr;
  t.thr_f = (float)thr;
  t.fast = thr > 0.0 && thr < 1.0e6;
  return t;
}

__device__ __forceinline__ bool suppresses(const float4& a, float area_a, const float4& b, float area_b,
                                           const IouThr& t) {
  const float xx1 = fmaxf(a.x, b.x), yy1 = fmaxf(a.y, b.y);
  const float xx2 = fminf(a.z, b.z), yy2 = fminf(a.w, b.w);
  const float w = fmaxf(0.0f, __fsub_rn(xx2, xx1)), h = fmaxf(0.0f, __fsub_rn(yy2, yy1));
  const float inter = __fmul_rn(w, h);
  if (inter <= 0.0f && t.thr >= 0.0) return false;   // 0 / union is 0 (or NaN for two empty boxes): never above thr
  const float uni = __fsub_rn(__fadd_rn(area_a, area_b), inter);
  // far from the threshold the quotient's rounding cannot matter (|error| < 1e-6 relative): decide with one product;
  // only within 1e-4 of it, or for a degenerate union, do the reference's division and its comparison in double
  const float tu = t.thr_f * uni;
  if (t.fast && uni > 0.0f && tu > 1.0e-30f && tu < 3.0e38f) {
    if (inter > tu * 1.0001f) return true;
    if (inter < tu * 0.9999f) return false;
  }
  const float ovr = __fdiv_rn(inter, uni);
  return (double)ovr > t.thr;
}

// Greedy suppression over boxes already in decreasing score order, 32 sorted positions per round.
//   1. the 32 x 32 tests inside the round's block, one per thread (warp = candidate, lane = earlier candidate);
//   2. warp 0 resolves the block on the resulting bit rows (fixed point, see below): K = its kept candidates;
//   3. every thread tests its own later positions against the kept candidates only.
// Three barriers per 32 positions whatever the number of survivors (a barrier per kept box costs 0.5 us each).
// cls == nullptr: every pair competes.  alive[i] is 1 on entry for i < n.  blockDim.x must be kPostThreads.
//
// C > 1: the CTA is one of a cluster of C that all hold the same sorted boxes.  Steps 1 and 2 run in every CTA (same
// inputs, same K).  Step 3 is a per-thread chain of popc(K) dependent-latency tests (measured: splitting the
// *positions* over the cluster does not shorten it), so the round's kept candidates are split instead: CTA r tests
// every later position against the candidates c = r mod C only.  Each CTA's alive[] then holds a partial view; before
// the next round every CTA publishes its view of the next block (one 32-bit word per CTA, stored into every CTA's
// shared memory through DSMEM), one cluster barrier replaces the round's last CTA barrier, and step 2 ANDs the C
// words.  A block's final state is the K every CTA computes for it, so each CTA ends with the complete result.
template <int C>
__device__ void block_greedy(const float4* sb, const int32_t* cls, volatile uint8_t* alive, int n, double thr,
                             uint32_t* mrow /* [33] shared */, unsigned long long* trace = nullptr,
                             uint32_t* xbits = nullptr /* C > 1: [2][C] shared words, each CTA's view of the next block */) {
  static_assert(C == 1 || C == 8, "the candidate split below is written for clusters of 8");
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int rank = C > 1 ? (int)cg::this_cluster().block_rank() : 0;
  long long c1 = 0, c2 = 0, c3 = 0, c4 = 0, t0 = 0, t1 = 0, t2 = 0, t3c = 0;   // thread 0's cycles in the three steps (trace only)
  const IouThr T = make_thr(thr);
  __syncthreads();
  for (int i0 = 0; i0 < n; i0 += 32) {
    if (trace) t0 = clock64();
    {
      const int c = i0 + warp, e = i0 + lane;
      bool sup = false;
      if (lane < warp && c < n && (!cls || cls[c] == cls[e])) {
        const float4 be = sb[e], bc = sb[c];
        sup = suppresses(be, box_area(be), bc, box_area(bc), T);
      }
      const uint32_t m = __ballot_sync(0xFFFFFFFFu, sup);
      if (lane == 0) mrow[warp] = m;
    }
    if (C > 1 && i0 > 0) cluster_wait_acquire();   // the other CTAs' views of this block have landed (arrive: end of last round)
    __syncthreads();
    if (trace) t1 = clock64();
    if (warp == 0) {
      // Greedy resolution of the block as a fixed point instead of a 32-step dependent chain (which, with the other
      // 31 warps parked at the barrier and nothing to hide ALU latency, cost ~1000 cycles per round): a candidate is
      // kept once every earlier candidate that overlaps it is dead, dead once one of them is kept.  The lowest
      // undecided candidate is decided in every pass; typical depth is 2-4 passes of three ballots.
      const uint32_t m = mrow[lane];
      // C > 1: the block is alive where every CTA of the cluster still has it alive
      bool al;
      if (C > 1 && i0 > 0) {
        uint32_t merged = 0xFFFFFFFFu;
#pragma unroll
        for (int r = 0; r < C; ++r) merged &= xbits[((i0 >> 5) & 1) * C + r];
        al = i0 + lane < n && ((merged >> lane) & 1u) != 0u;
      } else {
        al = i0 + lane < n && alive[i0 + lane] != 0;
      }
      uint32_t dead = ~__ballot_sync(0xFFFFFFFFu, al);
      uint32_t K = 0;
      bool undecided = al;
      while (true) {
        const bool k_now = undecided && (m & ~dead) == 0u;
        const bool d_now = undecided && (m & K) != 0u;
        K |= __ballot_sync(0xFFFFFFFFu, k_now);
        dead |= __ballot_sync(0xFFFFFFFFu, d_now);
        undecided = undecided && !k_now && !d_now;
        if (__ballot_sync(0xFFFFFFFFu, undecided) == 0u) break;
      }
      if (lane == 0) mrow[32] = K;
      if (i0 + lane < n) alive[i0 + lane] = (K >> lane) & 1u;
    }
    __syncthreads();
    if (trace) t2 = clock64();
    const uint32_t K = mrow[32];
    auto test_position = [&](int j, uint32_t Kuse) {
      if (!alive[j]) return;
      const float4 bj = sb[j];
      const float aj = box_area(bj);
      const int cj = cls ? cls[j] : 0;
      uint32_t k = Kuse;
      const int nk = __popc(Kuse);
      bool dead = false;      // no early exit: the tests are independent, so their latencies overlap
#pragma unroll 4
      for (int t = 0; t < nk; ++t) {
        const int c = i0 + __ffs(k) - 1;
        k &= k - 1;
        const float4 bc = sb[c];
        const bool same = !cls || cls[c] == cj;
        // (a four-compare early-out for pairs that do not touch was measured: slower on clustered boxes, where
        // some lane of most warps overlaps and the warp then runs both paths)
        dead |= same && suppresses(bc, box_area(bc), bj, aj, T);
      }
      if (dead) alive[j] = 0;
    };
    if (C == 1) {
      if (K)
        for (int j = i0 + 32 + threadIdx.x; j < n; j += kPostThreads) test_position(j, K);
      __syncthreads();
    } else {
      const uint32_t Kmine = K & (0x01010101u << rank);          // C == 8: candidates rank, rank + 8, rank + 16, rank + 24
      if (Kmine)
        for (int j = i0 + 32 + threadIdx.x; j < n; j += kPostThreads) test_position(j, Kmine);
      if (warp == 0 && i0 + 32 < n) {
        // this warp owns the next block's positions (j = i0 + 32 + lane): publish this CTA's view of them, lane r
        // storing the word into CTA r's shared memory
        __syncwarp();
        const int j = i0 + 32 + lane;
        const uint32_t bits = __ballot_sync(0xFFFFFFFFu, j < n && alive[j]);
        // (two sets of words, alternating by round: a fast CTA may publish round r + 1 while a slow one still reads r)
        if (lane < C) st_shared_cluster_u32(xbits + (((i0 >> 5) + 1) & 1) * C + rank, lane, bits);
      }
      if (trace) t3c = clock64();
      // split barrier: arrive now, wait after the next round's pair tests (which need nothing from the other CTAs)
      cluster_arrive_release();
    }
    if (trace) {
      const long long t3 = clock64();
      c1 += t1 - t0;
      c2 += t2 - t1;
      c3 += t3 - t2;
      if (C > 1) c4 += t3 - t3c;      // of which: the cluster barrier
    }
  }
  if (C > 1 && n > 0) cluster_wait_acquire();   // pairs the last arrive: no CTA leaves while another may still store into it
  if (C > 1) __syncthreads();
  if (trace && threadIdx.x == 0) {     // slots 6, 7: cycles of steps 1+2 and of step 3, summed over the rounds
    trace[6] = (unsigned long long)(c1 << 32 | (c2 & 0xFFFFFFFFll));
    trace[7] = (unsigned long long)(c4 << 32 | (c3 & 0xFFFFFFFFll));
  }
}

// Ordered compaction of the surviving sorted positions: calls emit(rank, position) for each, returns the count.
template <typename Emit>
__device__ int block_compact(const volatile uint8_t* alive, int n, int* warp_tot /* [kPostWarps + 1] shared */,
                             Emit emit) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int base = 0;
  for (int start = 0; start < n; start += kPostThreads) {
    const int pos = start + threadIdx.x;
    const bool f = pos < n && alive[pos];
    const uint32_t m = __ballot_sync(0xFFFFFFFFu, f);
    if (lane == 0) warp_tot[warp] = __popc(m);
    __syncthreads();
    if (warp == 0) {
      int v = warp_tot[lane];
      int incl = v;
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= o) incl += y;
      }
      warp_tot[lane] = incl - v;
      if (lane == 31) warp_tot[kPostWarps] = incl;
    }
    __syncthreads();
    if (f) emit(base + warp_tot[warp] + __popc(m & ((1u << lane) - 1u)), pos);
    base += warp_tot[kPostWarps];
    __syncthreads();
  }
  return base;
}

struct PostSmem {
  uint64_t* keys;    // [npad]
  float4* sb;        // [npad] sorted boxes
  int32_t* scls;     // [npad] class of the sorted box
  uint8_t* alive;    // [npad]
  int* warp_tot;     // [kPostWarps + 1]
  float* red;        // [kPostWarps]
  unsigned char* tail;
};

__device__ __forceinline__ PostSmem carve(unsigned char* base, int npad) {
  PostSmem s;
  s.sb = reinterpret_cast<float4*>(base);
  s.keys = reinterpret_cast<uint64_t*>(base + (size_t)npad * 16);
  s.scls = reinterpret_cast<int32_t*>(base + (size_t)npad * 24);
  s.warp_tot = reinterpret_cast<int*>(base + (size_t)npad * 28);
  s.red = reinterpret_cast<float*>(base + (size_t)npad * 28 + 256);
  s.alive = reinterpret_cast<uint8_t*>(base + (size_t)npad * 28 + 512);
  s.tail = base + (size_t)npad * 29 + 512;
  return s;
}
size_t post_smem_bytes(int npad) { return (size_t)npad * 29 + 512; }

// ---------------------------------------------------------------------------------------------------------------
// ovm_nms

template <int C>
__global__ void __launch_bounds__(kPostThreads, 1)
nms_kernel(const float* __restrict__ boxes, const float* __restrict__ scores, const int64_t* __restrict__ classes,
           const int32_t* __restrict__ seg_offsets, int max_len, double thr, int mode, int32_t* __restrict__ keep,
           int32_t* __restrict__ keep_count) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int seg = blockIdx.x / C;       // a cluster of C CTAs per segment
  const int r0 = seg_offsets[seg];
  const int n = seg_offsets[seg + 1] - r0;
  if (n <= 0 || n > max_len) {
    if (threadIdx.x == 0) keep_count[seg] = n <= 0 ? 0 : -1;
    return;
  }
  const int npad = next_pow2(n);
  PostSmem s = carve(smem_raw, npad);
  const float4* gb = reinterpret_cast<const float4*>(boxes) + r0;

  // torchvision batched_nms' coordinate trick: offset = class * (max coordinate of the segment + 1)
  float off_unit = 0.0f;
  if (mode == OVM_NMS_COORD_TRICK) {
    float m = -INFINITY;
    for (int i = threadIdx.x; i < n; i += kPostThreads) {
      const float4 b = gb[i];
      m = fmaxf(fmaxf(fmaxf(m, b.x), b.y), fmaxf(b.z, b.w));
    }
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xFFFFFFFFu, m, o));
    if ((threadIdx.x & 31) == 0) s.red[threadIdx.x >> 5] = m;
    __syncthreads();
    m = s.red[0];
    for (int w = 1; w < kPostWarps; ++w) m = fmaxf(m, s.red[w]);
    off_unit = __fadd_rn(m, 1.0f);
  }
  for (int i = threadIdx.x; i < npad; i += kPostThreads) {
    s.keys[i] = i < n ? ((uint64_t)score_key_desc(scores[r0 + i]) << 32) | (uint32_t)i : kPadKey;
    s.alive[i] = i < n;
  }
  block_sort(s.keys, npad);
  for (int p = threadIdx.x; p < n; p += kPostThreads) {
    const int i = (int)(uint32_t)s.keys[p];
    float4 b = gb[i];
    int c = 0;
    if (mode != OVM_NMS_AGNOSTIC) {
      const int64_t cl = classes[r0 + i];
      c = (int)cl;
      if (mode == OVM_NMS_COORD_TRICK) {
        const float o = __fmul_rn((float)cl, off_unit);
        b.x = __fadd_rn(b.x, o);
        b.y = __fadd_rn(b.y, o);
        b.z = __fadd_rn(b.z, o);
        b.w = __fadd_rn(b.w, o);
      }
    }
    s.sb[p] = b;
    s.scls[p] = c;
  }
  if (C > 1) {
    // block_greedy's first round already stores into the peer CTAs' exchange words, which alias s.red: every CTA of the
    // cluster must have started (CUDA's rule for any distributed-shared-memory access) and be past its own reads of
    // s.red in the coordinate-trick reduction above before a faster peer may do that
    cluster_arrive_release();
    cluster_wait_acquire();
  }
  block_greedy<C>(s.sb, mode == OVM_NMS_PER_CLASS ? s.scls : nullptr, s.alive, n, thr,
                  reinterpret_cast<uint32_t*>(s.warp_tot), nullptr, reinterpret_cast<uint32_t*>(s.red));
  if (C > 1 && blockIdx.x % C != 0) return;     // every CTA has the result; the first of the cluster writes it
  int32_t* kout = keep + r0;
  const uint64_t* keys = s.keys;
  const int cnt = block_compact(s.alive, n, s.warp_tot, [&](int rank, int pos) { kout[rank] = (int32_t)(uint32_t)keys[pos]; });
  if (threadIdx.x == 0) keep_count[seg] = cnt;
}

// ---------------------------------------------------------------------------------------------------------------
// ovm_gdino_post

// Query scores (roi_heads_gdino.py:186-202, 296-317): one warp per query, spread over the whole GPU (one CTA doing all
// 900 queries took 78 of the post-processing's 144 us).  sigmoid of the query's token logits into shared memory,
// phrase sums over each phrase's token columns, max / first argmax.  Results go to q_score / q_cls [B, nq].
constexpr int kScoreWarps = 8;

__global__ void __launch_bounds__(kScoreWarps * 32)
gdino_score_kernel(const float* __restrict__ logits, const int32_t* __restrict__ phrase_ranges, int nq, int T, int P,
                   float* __restrict__ q_score, int64_t* __restrict__ q_cls) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* rows = reinterpret_cast<float*>(smem_raw);                           // [kScoreWarps][T]
  int2* pr = reinterpret_cast<int2*>(rows + (size_t)kScoreWarps * T);         // [P] phrase ranges, clamped to [0, T]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = blockIdx.y;
  for (int p = threadIdx.x; p < P; p += blockDim.x)
    pr[p] = make_int2(min(max(phrase_ranges[2 * p], 0), T), min(max(phrase_ranges[2 * p + 1], 0), T));
  __syncthreads();
  const int q = blockIdx.x * kScoreWarps + warp;
  if (q >= nq) return;
  // only the token columns some phrase covers are ever summed: [t_lo, t_hi) (GroundingDINO fills the columns past the
  // caption with -inf, whose sigmoid would take the slow paths of expf and the reciprocal for nothing)
  int t_lo = T, t_hi = 0;
  for (int p = lane; p < P; p += 32) {
    t_lo = min(t_lo, pr[p].x);
    t_hi = max(t_hi, pr[p].y);
  }
  for (int o = 16; o > 0; o >>= 1) {
    t_lo = min(t_lo, __shfl_xor_sync(0xFFFFFFFFu, t_lo, o));
    t_hi = max(t_hi, __shfl_xor_sync(0xFFFFFFFFu, t_hi, o));
  }
  float* row = rows + (size_t)warp * T;
  const float* src = logits + ((size_t)b * nq + q) * T;
  for (int t = t_lo + lane; t < t_hi; t += 32) row[t] = __frcp_rn(__fadd_rn(1.0f, expf(-src[t])));   // = 1 / (1 + e^-x)
  __syncwarp();
  float best = -INFINITY;
  int best_p = 0x7FFFFFFF;
  for (int p = lane; p < P; p += 32) {
    const int t0 = pr[p].x, t1 = pr[p].y;
    float sum = 0.0f;
    for (int t = t0; t < t1; ++t) sum = __fadd_rn(sum, row[t]);
    if (sum > best || (sum != sum && best == best)) {   // NaN wins like torch.max, first occurrence kept
      best = sum;
      best_p = p;
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    const float ob = __shfl_xor_sync(0xFFFFFFFFu, best, o);
    const int op = __shfl_xor_sync(0xFFFFFFFFu, best_p, o);
    const bool o_nan = ob != ob, m_nan = best != best;
    const bool take = (o_nan && (!m_nan || op < best_p)) || (!m_nan && !o_nan && (ob > best || (ob == best && op < best_p)));
    if (take) {
      best = ob;
      best_p = op;
    }
  }
  if (lane == 0) {
    q_score[(size_t)b * nq + q] = best;
    q_cls[(size_t)b * nq + q] = best_p;
  }
}

// The rest for one image per CTA.  out_scores / out_classes hold the query scores / classes of gdino_score_kernel on
// entry (indexed by query); they are copied to shared memory first and then overwritten with the result (indexed by
// keep rank).
template <int C>
__global__ void __launch_bounds__(kPostThreads, 1)
gdino_post_kernel(const float* __restrict__ boxes_cxcywh, int nq, float img_w, float img_h, float box_thr,
                  double nms_thr, float* __restrict__ out_boxes, float* out_scores, int64_t* out_classes,
                  int32_t* __restrict__ out_query, int32_t* __restrict__ out_count, unsigned long long* trace) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int b = blockIdx.x / C;         // a cluster of C CTAs per image
  trace = (trace && blockIdx.x % C == 0) ? trace + b * 8 : nullptr;
  const int npad = next_pow2(nq);
  PostSmem s = carve(smem_raw, npad);
  float* q_score = reinterpret_cast<float*>(s.tail);                 // [npad]
  int32_t* q_cls = reinterpret_cast<int32_t*>(s.tail + (size_t)npad * 4);   // [npad]
  post_stamp(trace, 0);
  for (int i = threadIdx.x; i < nq; i += kPostThreads) {
    q_score[i] = out_scores[(size_t)b * nq + i];
    q_cls[i] = (int32_t)out_classes[(size_t)b * nq + i];
  }
  if (C > 1) cg::this_cluster().sync();   // every CTA of the cluster has its copy before the first one overwrites the buffers
  __syncthreads();
  post_stamp(trace, 1);

  // 2. candidates = queries above the box threshold, in stable decreasing score order
  for (int i = threadIdx.x; i < npad; i += kPostThreads) {
    const bool ok = i < nq && q_score[i] > box_thr;
    s.keys[i] = ok ? ((uint64_t)score_key_desc(q_score[i]) << 32) | (uint32_t)i : kPadKey;
  }
  block_sort(s.keys, npad);
  post_stamp(trace, 2);
  // number of candidates = first pad key (keys are sorted)
  int* n_cand = s.warp_tot + kPostWarps;
  if (threadIdx.x == 0) *n_cand = npad;
  __syncthreads();
  for (int i = threadIdx.x; i < npad; i += kPostThreads)
    if (s.keys[i] == kPadKey && (i == 0 || s.keys[i - 1] != kPadKey)) *n_cand = i;
  __syncthreads();
  const int n = *n_cand;
  __syncthreads();

  // 3. boxes: cxcywh (normalised) * (w, h, w, h) -> xyxy (roi_heads_gdino.py:270-271, 283-287)
  const float4* gb = reinterpret_cast<const float4*>(boxes_cxcywh) + (size_t)b * nq;
  for (int p = threadIdx.x; p < npad; p += kPostThreads) {
    s.alive[p] = p < n;
    if (p < n) {
      const float4 c = gb[(uint32_t)s.keys[p]];
      const float cx = __fmul_rn(c.x, img_w), cy = __fmul_rn(c.y, img_h);
      const float hw = __fmul_rn(0.5f, __fmul_rn(c.z, img_w)), hh = __fmul_rn(0.5f, __fmul_rn(c.w, img_h));
      s.sb[p] = make_float4(__fsub_rn(cx, hw), __fsub_rn(cy, hh), __fadd_rn(cx, hw), __fadd_rn(cy, hh));
    }
  }
  post_stamp(trace, 3);
  block_greedy<C>(s.sb, nullptr, s.alive, n, nms_thr, reinterpret_cast<uint32_t*>(s.warp_tot), trace,
                  reinterpret_cast<uint32_t*>(s.red));
  post_stamp(trace, 4);
  if (C > 1 && blockIdx.x % C != 0) return;

  // 4. survivors in keep order
  float4* ob = reinterpret_cast<float4*>(out_boxes) + (size_t)b * nq;
  float* os = out_scores + (size_t)b * nq;
  int64_t* oc = out_classes + (size_t)b * nq;
  int32_t* oq = out_query ? out_query + (size_t)b * nq : nullptr;
  const uint64_t* keys = s.keys;
  const float4* sb = s.sb;
  const int cnt = block_compact(s.alive, n, s.warp_tot, [&](int rank, int pos) {
    const int q = (int)(uint32_t)keys[pos];
    ob[rank] = sb[pos];
    os[rank] = q_score[q];
    oc[rank] = (int64_t)q_cls[q];
    if (oq) oq[rank] = q;
  });
  if (threadIdx.x == 0) out_count[b] = cnt;
  post_stamp(trace, 5);
}

// ---------------------------------------------------------------------------------------------------------------
// ovm_box_inference: FastRCNNOutputLayers.predict_probs / predict_boxes + fast_rcnn_inference_single_image
// (reference cubercnn/modeling/roi_heads/fast_rcnn.py:57-143; d2 Box2BoxTransform.apply_deltas, Boxes.clip).
//   box_predict_kernel    warp per proposal: softmax, per-class boxes, clip, validity, per-image candidate stats
//   box_class_nms_kernel  CTA per (class, image): candidates above the score threshold, sorted, suppressed
//   box_topk_kernel       CTA per image: the per-class survivors merged in decreasing score order, first top-k

struct BoxCfg {
  int K, agnostic;
  float wx, wy, ww, wh, scale_clamp, score_thresh;
  double nms_thresh;
  int topk, coord_trick_max;
};

__device__ __forceinline__ bool finite_f(float x) { return fabsf(x) <= 3.402823466e38f; }

__global__ void __launch_bounds__(256)
box_predict_kernel(BoxCfg cfg, const float* __restrict__ logits, int64_t ld_logits, const float* __restrict__ deltas,
                   int64_t ld_deltas, const float* __restrict__ proposals, const int32_t* __restrict__ seg_offsets,
                   const int32_t* __restrict__ image_sizes, int N, int R, float* __restrict__ probs,
                   float* __restrict__ cand_boxes, int32_t* __restrict__ image_stats) {
  const int lane = threadIdx.x & 31;
  const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= R) return;
  const int K = cfg.K, Kb = cfg.agnostic ? 1 : K;
  // image of this proposal (segments are sorted; N is small)
  int img = 0;
  {
    int lo = 0, hi = N;           // largest i with seg_offsets[i] <= r
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (seg_offsets[mid] <= r) lo = mid; else hi = mid;
    }
    img = lo;
  }
  const float im_h = (float)image_sizes[2 * img], im_w = (float)image_sizes[2 * img + 1];
  // softmax over the K + 1 logits (F.softmax(scores, -1), d2 fast_rcnn.py predict_probs)
  const float* lg = logits + (size_t)r * ld_logits;
  float mx = -INFINITY;
  bool fin = true;
  for (int k = lane; k <= K; k += 32) {
    const float v = lg[k];
    fin &= finite_f(v);
    mx = fmaxf(mx, v);
  }
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xFFFFFFFFu, mx, o));
  float sum = 0.0f;
  for (int k = lane; k <= K; k += 32) sum += expf(lg[k] - mx);
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xFFFFFFFFu, sum, o);
  // boxes (Box2BoxTransform.apply_deltas, then Boxes.clip to the image)
  const float4 p = reinterpret_cast<const float4*>(proposals)[r];
  const float pw = __fsub_rn(p.z, p.x), ph = __fsub_rn(p.w, p.y);
  const float pcx = __fadd_rn(p.x, __fmul_rn(0.5f, pw)), pcy = __fadd_rn(p.y, __fmul_rn(0.5f, ph));
  const float* dl = deltas + (size_t)r * ld_deltas;
  float4* cb = reinterpret_cast<float4*>(cand_boxes) + (size_t)r * Kb;
  for (int k = lane; k < Kb; k += 32) {
    const float dx = __fdiv_rn(dl[4 * k], cfg.wx), dy = __fdiv_rn(dl[4 * k + 1], cfg.wy);
    const float dw = fminf(__fdiv_rn(dl[4 * k + 2], cfg.ww), cfg.scale_clamp);
    const float dh = fminf(__fdiv_rn(dl[4 * k + 3], cfg.wh), cfg.scale_clamp);
    const float cx = __fadd_rn(__fmul_rn(dx, pw), pcx), cy = __fadd_rn(__fmul_rn(dy, ph), pcy);
    const float w = __fmul_rn(expf(dw), pw), h = __fmul_rn(expf(dh), ph);
    float4 b = make_float4(__fsub_rn(cx, __fmul_rn(0.5f, w)), __fsub_rn(cy, __fmul_rn(0.5f, h)),
                           __fadd_rn(cx, __fmul_rn(0.5f, w)), __fadd_rn(cy, __fmul_rn(0.5f, h)));
    fin &= finite_f(b.x) && finite_f(b.y) && finite_f(b.z) && finite_f(b.w);   // isfinite on the unclipped boxes
    b.x = fminf(fmaxf(b.x, 0.0f), im_w);
    b.y = fminf(fmaxf(b.y, 0.0f), im_h);
    b.z = fminf(fmaxf(b.z, 0.0f), im_w);
    b.w = fminf(fmaxf(b.w, 0.0f), im_h);
    cb[k] = b;
  }
  __syncwarp();       // class-agnostic: every lane reads the one box lane 0 wrote
  fin = __all_sync(0xFFFFFFFFu, fin);
  // probabilities; a row with a non-finite number is dropped as a whole (fast_rcnn.py:75-78): NaN never passes `>`
  int cnt = 0;
  float cmax = 0.0f;
  float* pr = probs + (size_t)r * K;
  for (int k = lane; k < K; k += 32) {
    const float pk = fin ? __fdiv_rn(expf(lg[k] - mx), sum) : __int_as_float(0x7FC00000);
    pr[k] = pk;
    if (pk > cfg.score_thresh) {
      ++cnt;
      const float4 b = cb[cfg.agnostic ? 0 : k];     // own write (k == lane-strided) or, agnostic, lane 0's
      cmax = fmaxf(cmax, fmaxf(fmaxf(b.x, b.y), fmaxf(b.z, b.w)));
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    cnt += __shfl_xor_sync(0xFFFFFFFFu, cnt, o);
    cmax = fmaxf(cmax, __shfl_xor_sync(0xFFFFFFFFu, cmax, o));
  }
  if (lane == 0 && cnt) {
    atomicAdd(&image_stats[2 * img], cnt);
    atomicMax(&image_stats[2 * img + 1], __float_as_int(cmax));     // clipped coordinates are >= 0: int order = float order
  }
}

__global__ void __launch_bounds__(kPostThreads, 1)
box_class_nms_kernel(BoxCfg cfg, const float* __restrict__ probs, const float* __restrict__ cand_boxes,
                     const int32_t* __restrict__ seg_offsets, const int32_t* __restrict__ image_stats,
                     int max_per_image, int32_t* __restrict__ class_keep, int32_t* __restrict__ class_count) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int k = blockIdx.x, img = blockIdx.y, K = cfg.K;
  const int r0 = seg_offsets[img];
  const int len = min(seg_offsets[img + 1] - r0, max_per_image);
  int32_t* kout = class_keep + ((size_t)img * K + k) * max_per_image;
  if (len <= 0 || image_stats[2 * img] == 0) {
    if (threadIdx.x == 0) class_count[img * K + k] = 0;
    return;
  }
  const int npad = next_pow2(len);
  PostSmem s = carve(smem_raw, npad);
  int* n_cand = s.warp_tot + kPostWarps;
  for (int i = threadIdx.x; i < npad; i += kPostThreads) {
    float sc = 0.0f;
    const bool ok = i < len && (sc = probs[(size_t)(r0 + i) * K + k]) > cfg.score_thresh;
    s.keys[i] = ok ? ((uint64_t)score_key_desc(sc) << 32) | (uint32_t)i : kPadKey;
  }
  block_sort(s.keys, npad);
  if (threadIdx.x == 0) *n_cand = npad;
  __syncthreads();
  for (int i = threadIdx.x; i < npad; i += kPostThreads)
    if (s.keys[i] == kPadKey && (i == 0 || s.keys[i - 1] != kPadKey)) *n_cand = i;
  __syncthreads();
  const int n = *n_cand;
  __syncthreads();
  if (n == 0) {
    if (threadIdx.x == 0) class_count[img * K + k] = 0;
    return;
  }
  // torchvision batched_nms: coordinate trick while the image has few candidates, plain per-class NMS above
  float off = 0.0f;
  if (image_stats[2 * img] <= cfg.coord_trick_max)
    off = __fmul_rn((float)k, __fadd_rn(__int_as_float(image_stats[2 * img + 1]), 1.0f));
  const int Kb = cfg.agnostic ? 1 : K;
  const float4* cb = reinterpret_cast<const float4*>(cand_boxes);
  for (int p = threadIdx.x; p < npad; p += kPostThreads) {
    s.alive[p] = p < n;
    if (p < n) {
      float4 b = cb[(size_t)(r0 + (int)(uint32_t)s.keys[p]) * Kb + (cfg.agnostic ? 0 : k)];
      b.x = __fadd_rn(b.x, off);
      b.y = __fadd_rn(b.y, off);
      b.z = __fadd_rn(b.z, off);
      b.w = __fadd_rn(b.w, off);
      s.sb[p] = b;
    }
  }
  block_greedy<1>(s.sb, nullptr, s.alive, n, cfg.nms_thresh, reinterpret_cast<uint32_t*>(s.warp_tot));
  const uint64_t* keys = s.keys;
  const int cnt = block_compact(s.alive, n, s.warp_tot, [&](int rank, int pos) { kout[rank] = (int32_t)(uint32_t)keys[pos]; });
  if (threadIdx.x == 0) class_count[img * K + k] = cnt;
}

__global__ void __launch_bounds__(kPostThreads, 1)
box_topk_kernel(BoxCfg cfg, const float* __restrict__ probs, const float* __restrict__ cand_boxes,
                const int32_t* __restrict__ seg_offsets, int max_per_image, const int32_t* __restrict__ class_keep,
                const int32_t* __restrict__ class_count, int npad, int npad_small, int cap, float* __restrict__ out_boxes,
                float* __restrict__ out_scores, int64_t* __restrict__ out_classes, int32_t* __restrict__ out_index,
                float* __restrict__ out_scores_full, int32_t* __restrict__ out_count) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  uint64_t* keys = reinterpret_cast<uint64_t*>(smem_raw);            // [npad]
  uint64_t* small = keys + npad;                                     // [npad_small]: the selected top-k keys
  int* offs = reinterpret_cast<int*>(smem_raw + (size_t)(npad + npad_small) * 8);   // [K + 1]
  const int img = blockIdx.x, K = cfg.K;
  const int r0 = seg_offsets[img];
  const int len = min(seg_offsets[img + 1] - r0, max_per_image);
  const int per_class = cfg.topk >= 0 ? min(cfg.topk, max_per_image) : max_per_image;
  int* wsum = offs + K + 1;                                          // [33]
  {   // offs = exclusive scan over the classes of min(survivors, top-k): K <= 1024 = one value per thread
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c = (threadIdx.x < K && len > 0) ? min(class_count[img * K + threadIdx.x], per_class) : 0;
    int incl = c;
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xFFFFFFFFu, incl, o);
      if (lane >= o) incl += y;
    }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      const int v = wsum[lane];
      int wi = v;
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xFFFFFFFFu, wi, o);
        if (lane >= o) wi += y;
      }
      wsum[lane] = wi - v;
      if (lane == 31) wsum[32] = wi;
    }
    __syncthreads();
    if (threadIdx.x < K) offs[threadIdx.x] = wsum[warp] + incl - c;
    if (threadIdx.x == 0) offs[K] = wsum[32];
  }
  __syncthreads();
  const int total = offs[K];
  npad = min(npad, next_pow2(total));      // sort what is there, not the worst case the buffer was sized for
  for (int i = total + threadIdx.x; i < npad; i += kPostThreads) keys[i] = kPadKey;
  // a class list is already in decreasing score order: only its first top-k entries can reach the image's top-k
  for (int k = threadIdx.x >> 5; k < K; k += kPostWarps) {
    const int c = offs[k + 1] - offs[k];
    const int32_t* kl = class_keep + ((size_t)img * K + k) * max_per_image;
    for (int j = threadIdx.x & 31; j < c; j += 32) {
      const int r = kl[j];
      const float sc = probs[(size_t)(r0 + r) * K + k];
      keys[offs[k] + j] = ((uint64_t)score_key_desc(sc) << 32) | (uint32_t)(r * K + k);   // ties: (r, k) row-major,
    }                                                                                      // the order of nonzero()
  }
  const int n_out = cfg.topk >= 0 ? min(total, cfg.topk) : total;
  if (n_out > 0 && n_out < total && next_pow2(n_out) <= npad_small) {
    // Far more survivors than top-k (5000 against 100 at the shipped sizes): radix-select the top-k-th key, eight
    // 8-bit passes over the unsorted keys with a 256-bin shared-memory histogram, then sort only the selected keys
    // (sorting all 8192 slots took 84 us of the branch's 126).  Keys are unique, so exactly n_out keys are <= it.
    uint32_t* hist = reinterpret_cast<uint32_t*>(wsum + 33);         // [256]
    int* sel = reinterpret_cast<int*>(hist + 256);                   // [0] bin, [1] rank left inside the bin, [2] slot counter
    uint64_t prefix = 0;
    int want = n_out;
    __syncthreads();
    for (int shift = 56; shift >= 0; shift -= 8) {
      for (int i = threadIdx.x; i < 256; i += kPostThreads) hist[i] = 0;
      __syncthreads();
      const uint64_t hi_mask = shift == 56 ? 0ull : (~0ull << (shift + 8));
      for (int i = threadIdx.x; i < total; i += kPostThreads) {
        const uint64_t k = keys[i];
        if ((k & hi_mask) == prefix) atomicAdd(&hist[(uint32_t)(k >> shift) & 255u], 1u);
      }
      __syncthreads();
      if (threadIdx.x < 32) {
        const int lane = threadIdx.x;
        uint32_t h[8], loc = 0;
#pragma unroll
        for (int b = 0; b < 8; ++b) {
          h[b] = hist[lane * 8 + b];
          loc += h[b];
        }
        uint32_t incl = loc;
        for (int o = 1; o < 32; o <<= 1) {
          const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, incl, o);
          if (lane >= o) incl += y;
        }
        uint32_t before = incl - loc;
        if (before < (uint32_t)want && (uint32_t)want <= incl) {      // the crossing is inside this lane's 8 bins
#pragma unroll
          for (int b = 0; b < 8; ++b) {
            if (before < (uint32_t)want && (uint32_t)want <= before + h[b]) {
              sel[0] = lane * 8 + b;
              sel[1] = want - (int)before;
            }
            before += h[b];
          }
        }
      }
      __syncthreads();
      prefix |= (uint64_t)(uint32_t)sel[0] << shift;
      want = sel[1];
      __syncthreads();
    }
    const int np2 = next_pow2(n_out);
    if (threadIdx.x == 0) sel[2] = 0;
    for (int i = threadIdx.x; i < np2; i += kPostThreads) small[i] = kPadKey;
    __syncthreads();
    for (int i = threadIdx.x; i < total; i += kPostThreads) {
      const uint64_t k = keys[i];
      if (k <= prefix) small[atomicAdd(&sel[2], 1)] = k;
    }
    block_sort(small, np2);
    keys = small;
  } else {
    block_sort(keys, npad);
  }
  const int Kb = cfg.agnostic ? 1 : K;
  const float4* cb = reinterpret_cast<const float4*>(cand_boxes);
  for (int j = threadIdx.x; j < n_out; j += kPostThreads) {
    const uint32_t id = (uint32_t)keys[j];
    const int r = id / K, k = id - r * K;
    const size_t o = (size_t)img * cap + j;
    reinterpret_cast<float4*>(out_boxes)[o] = cb[(size_t)(r0 + r) * Kb + (cfg.agnostic ? 0 : k)];
    out_scores[o] = probs[(size_t)(r0 + r) * K + k];
    out_classes[o] = k;
    out_index[o] = r;
  }
  if (out_scores_full)
    for (int e = threadIdx.x; e < n_out * K; e += kPostThreads) {
      const int j = e / K, k = e - j * K;
      const int r = (uint32_t)keys[j] / K;
      out_scores_full[((size_t)img * cap + j) * K + k] = probs[(size_t)(r0 + r) * K + k];
    }
  if (threadIdx.x == 0) out_count[img] = n_out;
}

// Cluster size for the per-image NMS kernels: 8 (portable maximum) unless OVM_NMS_CLUSTER=1 asks for single CTAs.
int nms_cluster_size() {
  static int c = -1;
  if (c < 0) {
    const char* e = std::getenv("OVM_NMS_CLUSTER");
    c = (e && std::atoi(e) == 1) ? 1 : 8;
  }
  return c;
}

template <typename... KArgs, typename... Args>
cudaError_t launch_cluster(void (*kernel)(KArgs...), int clusters, int csize, size_t smem, cudaStream_t stream,
                           Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(clusters * csize);
  cfg.blockDim = dim3(kPostThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = csize;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

int host_next_pow2(int n) {
  int p = 32;
  while (p < n) p <<= 1;
  return p;
}

}  // namespace

void post_set_trace(unsigned long long* buf) { g_post_trace = buf; }

int nms_launch(const float* boxes, const float* scores, const int64_t* classes, const int32_t* seg_offsets,
               int32_t num_segments, int32_t max_segment_len, double iou_threshold, int32_t mode, int32_t* keep,
               int32_t* keep_count, cudaStream_t stream) {
  OVM_REQUIRE(num_segments >= 0 && max_segment_len >= 0, "ovm_nms: negative size");
  if (num_segments == 0) return OVM_OK;
  OVM_REQUIRE(boxes && scores && seg_offsets && keep && keep_count, "ovm_nms: NULL pointer");
  OVM_REQUIRE(mode == OVM_NMS_AGNOSTIC || mode == OVM_NMS_COORD_TRICK || mode == OVM_NMS_PER_CLASS,
              "ovm_nms: unknown mode %d", mode);
  OVM_REQUIRE(mode == OVM_NMS_AGNOSTIC || classes, "ovm_nms: this mode needs classes");
  OVM_REQUIRE(max_segment_len <= OVM_NMS_MAX_SEGMENT, "ovm_nms: segments of %d boxes exceed the %d one CTA holds",
              max_segment_len, OVM_NMS_MAX_SEGMENT);
  OVM_REQUIRE((reinterpret_cast<uintptr_t>(boxes) & 15) == 0, "ovm_nms: boxes must be 16-byte aligned");
  const size_t smem = post_smem_bytes(host_next_pow2(max_segment_len));
  static SmemOptIn opt[2];
  const int csize = nms_cluster_size();
  {
    int rc = csize > 1 ? opt[1].ensure(nms_kernel<8>, smem) : opt[0].ensure(nms_kernel<1>, smem);
    if (rc != OVM_OK) return rc;
  }
  OVM_CUDA(launch_cluster(csize > 1 ? nms_kernel<8> : nms_kernel<1>, num_segments, csize, smem, stream, boxes, scores,
                          classes, seg_offsets, max_segment_len, iou_threshold, mode, keep, keep_count));
  OVM_LAUNCHED("nms_kernel");
  return OVM_OK;
}

int gdino_post_launch(const float* logits, const float* boxes_cxcywh, const int32_t* phrase_ranges, int32_t batch,
                      int32_t nq, int32_t T, int32_t P, int32_t image_height, int32_t image_width,
                      float box_threshold, double nms_threshold, float* out_boxes, float* out_scores,
                      int64_t* out_classes, int32_t* out_query, int32_t* out_count, cudaStream_t stream) {
  OVM_REQUIRE(batch >= 0 && nq >= 0 && T > 0 && P > 0, "ovm_gdino_post: bad sizes (batch %d, queries %d, tokens %d, "
              "phrases %d)", batch, nq, T, P);
  if (batch == 0) return OVM_OK;
  OVM_REQUIRE(logits && boxes_cxcywh && phrase_ranges && out_boxes && out_scores && out_classes && out_count,
              "ovm_gdino_post: NULL pointer");
  OVM_REQUIRE(nq <= OVM_NMS_MAX_SEGMENT, "ovm_gdino_post: %d queries exceed the %d one CTA holds", nq,
              OVM_NMS_MAX_SEGMENT);
  OVM_REQUIRE(T <= 512, "ovm_gdino_post: %d text tokens (at most 512)", T);
  OVM_REQUIRE(((reinterpret_cast<uintptr_t>(boxes_cxcywh) | reinterpret_cast<uintptr_t>(out_boxes)) & 15) == 0,
              "ovm_gdino_post: boxes must be 16-byte aligned");
  if (nq == 0) {
    OVM_CUDA(cudaMemsetAsync(out_count, 0, sizeof(int32_t) * batch, stream));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return OVM_OK;
  }
  const size_t smem_score = (size_t)kScoreWarps * T * 4 + (size_t)P * 8;
  OVM_REQUIRE(smem_score <= 200 * 1024, "ovm_gdino_post: %d phrases need %zu bytes of shared memory", P, smem_score);
  static SmemOptIn opt_score;
  {
    int rc = opt_score.ensure(gdino_score_kernel, smem_score);
    if (rc != OVM_OK) return rc;
  }
  gdino_score_kernel<<<dim3((nq + kScoreWarps - 1) / kScoreWarps, batch), kScoreWarps * 32, smem_score, stream>>>(
      logits, phrase_ranges, nq, T, P, out_scores, out_classes);
  OVM_LAUNCHED("gdino_score_kernel");
  const int npad = host_next_pow2(nq);
  const size_t smem = post_smem_bytes(npad) + (size_t)npad * 8;
  static SmemOptIn opt[2];
  const int csize = nms_cluster_size();
  {
    int rc = csize > 1 ? opt[1].ensure(gdino_post_kernel<8>, smem) : opt[0].ensure(gdino_post_kernel<1>, smem);
    if (rc != OVM_OK) return rc;
  }
  OVM_CUDA(launch_cluster(csize > 1 ? gdino_post_kernel<8> : gdino_post_kernel<1>, batch, csize, smem, stream,
                          boxes_cxcywh, nq, (float)image_width, (float)image_height, box_threshold, nms_threshold,
                          out_boxes, out_scores, out_classes, out_query, out_count, g_post_trace));
  OVM_LAUNCHED("gdino_post_kernel");
  return OVM_OK;
}

int box_inference_launch(const ovm_box_inference_cfg_t* c, const float* logits, int64_t ld_logits, const float* deltas,
                         int64_t ld_deltas, const float* proposals, const int32_t* seg_offsets,
                         const int32_t* image_sizes, int32_t N, int32_t R, int32_t max_per_image,
                         const ovm_box_inference_ws_t* ws, float* out_boxes, float* out_scores, int64_t* out_classes,
                         int32_t* out_index, float* out_scores_full, int32_t* out_count, cudaStream_t stream) {
  OVM_REQUIRE(c && ws, "ovm_box_inference: NULL descriptor");
  OVM_REQUIRE(N >= 0 && R >= 0 && max_per_image >= 0, "ovm_box_inference: negative size");
  if (N == 0) return OVM_OK;
  OVM_REQUIRE(out_count, "ovm_box_inference: out_count is NULL");
  if (R == 0 || max_per_image == 0) {
    OVM_CUDA(cudaMemsetAsync(out_count, 0, sizeof(int32_t) * N, stream));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return OVM_OK;
  }
  const int K = c->num_classes;
  OVM_REQUIRE(K >= 1 && K <= 1024, "ovm_box_inference: %d classes (1..1024)", K);
  OVM_REQUIRE(logits && deltas && proposals && seg_offsets && image_sizes && out_boxes && out_scores && out_classes &&
              out_index, "ovm_box_inference: NULL pointer");
  OVM_REQUIRE(ws->probs && ws->cand_boxes && ws->class_keep && ws->class_count && ws->image_stats,
              "ovm_box_inference: incomplete workspace");
  OVM_REQUIRE(ld_logits >= K + 1 && ld_deltas >= (c->cls_agnostic_bbox_reg ? 4 : 4 * K),
              "ovm_box_inference: logits need %d columns, deltas %d", K + 1, c->cls_agnostic_bbox_reg ? 4 : 4 * K);
  OVM_REQUIRE(max_per_image <= OVM_NMS_MAX_SEGMENT, "ovm_box_inference: %d proposals per image exceed the %d one CTA "
              "holds", max_per_image, OVM_NMS_MAX_SEGMENT);
  OVM_REQUIRE(((reinterpret_cast<uintptr_t>(proposals) | reinterpret_cast<uintptr_t>(ws->cand_boxes) |
                reinterpret_cast<uintptr_t>(out_boxes)) & 15) == 0, "ovm_box_inference: boxes must be 16-byte aligned");
  const int per_class = c->topk_per_image >= 0 ? std::min(c->topk_per_image, max_per_image) : max_per_image;
  const int64_t merged = (int64_t)K * per_class;
  if (merged > 16384) {
    set_error("ovm_box_inference: %d classes x %d survivors per class = %lld merge candidates per image (at most "
              "16384 in this build)", K, per_class, (long long)merged);
    return OVM_EUNSUPPORTED;
  }
  const int cap = c->topk_per_image >= 0 ? c->topk_per_image : (int)merged;
  BoxCfg b;
  b.K = K;
  b.agnostic = c->cls_agnostic_bbox_reg != 0;
  b.wx = c->reg_weights[0];
  b.wy = c->reg_weights[1];
  b.ww = c->reg_weights[2];
  b.wh = c->reg_weights[3];
  b.scale_clamp = c->scale_clamp;
  b.score_thresh = c->score_thresh;
  b.nms_thresh = c->nms_thresh;
  b.topk = c->topk_per_image;
  b.coord_trick_max = c->coord_trick_max;
  OVM_CUDA(cudaMemsetAsync(ws->image_stats, 0, sizeof(int32_t) * 2 * N, stream));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  box_predict_kernel<<<(R + 7) / 8, 256, 0, stream>>>(b, logits, ld_logits, deltas, ld_deltas, proposals, seg_offsets,
                                                     image_sizes, N, R, ws->probs, ws->cand_boxes, ws->image_stats);
  OVM_LAUNCHED("box_predict_kernel");
  const size_t smem_nms = post_smem_bytes(host_next_pow2(max_per_image));
  static SmemOptIn opt_nms;
  {
    int rc = opt_nms.ensure(box_class_nms_kernel, smem_nms);
    if (rc != OVM_OK) return rc;
  }
  box_class_nms_kernel<<<dim3(K, N), kPostThreads, smem_nms, stream>>>(b, ws->probs, ws->cand_boxes, seg_offsets,
                                                                      ws->image_stats, max_per_image, ws->class_keep,
                                                                      ws->class_count);
  OVM_LAUNCHED("box_class_nms_kernel");
  const int npad = host_next_pow2((int)std::max<int64_t>(merged, 1));
  const int npad_small = c->topk_per_image >= 0 && c->topk_per_image <= 2048 ? host_next_pow2(c->topk_per_image) : 0;
  const size_t smem_topk = (size_t)(npad + npad_small) * 8 + (size_t)(K + 1 + 33 + 256 + 4) * 4;
  static SmemOptIn opt_topk;
  {
    int rc = opt_topk.ensure(box_topk_kernel, smem_topk);
    if (rc != OVM_OK) return rc;
  }
  box_topk_kernel<<<N, kPostThreads, smem_topk, stream>>>(b, ws->probs, ws->cand_boxes, seg_offsets, max_per_image,
                                                          ws->class_keep, ws->class_count, npad, npad_small, cap, out_boxes,
                                                          out_scores, out_classes, out_index, out_scores_full,
                                                          out_count);
  OVM_LAUNCHED("box_topk_kernel");
  return OVM_OK;
}

}  // namespace ovm
