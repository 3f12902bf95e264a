// Exact 3-D IoU of oriented cuboids at evaluation.
//
// Replaces `box3d_overlap` = `_check_coplanar` + `_check_nonzero` + `pytorch3d._C.iou_box3d` (reference
// cubercnn/evaluation/omni3d_evaluation.py:68-169), which the reference forces onto the CPU
// (MAX_DTS_CROSS_GTS_FOR_IOU3D = 0, :65) where pytorch3d walks every (detection, ground truth) pair serially.
//
// The pairs are independent: one warp per pair.  pytorch3d clips the 12 triangles of each box by the 6 face planes of
// the other, de-duplicates coplanar triangles pairwise and sums |tetrahedra| around the vertex mean.  Here the same
// polyhedron -- the intersection of two convex cells -- is built from its FACES: face f of one box, clipped by the six
// half-spaces of the other (Sutherland-Hodgman on a <= 10-gon, lane f of the warp), is exactly the part of the
// intersection's boundary that lies in the plane of f; the volume is the divergence-theorem sum
// (1/3) sum_f (n_f . x_f) area(f) with outward unit normals.  A face of the second box that lies in the same plane and
// faces the same way as a face of the first would be counted twice and is dropped (pytorch3d's coplanar-triangle rule);
// opposite-facing coplanar faces cancel by themselves.  Everything is translated to the first box's centre so that fp32
// products stay at the size of the boxes, not of the scene.
#include <cfloat>

#include "common.cuh"

namespace ovm {

namespace {

constexpr int kIouWarps = 4;
constexpr int kMaxPoly = 12;      // a quad clipped by six planes has at most 10 vertices

__constant__ int c_planes[6][4] = {{0, 1, 2, 3}, {3, 2, 6, 7}, {0, 1, 5, 4}, {0, 3, 7, 4}, {1, 2, 6, 5}, {4, 5, 6, 7}};
__constant__ int c_tris[12][3] = {{0, 1, 2}, {0, 3, 2}, {4, 5, 6}, {4, 6, 7}, {1, 5, 6}, {1, 6, 2},
                                  {0, 4, 7}, {0, 7, 3}, {3, 2, 6}, {3, 6, 7}, {0, 1, 5}, {0, 4, 5}};

struct V3 {
  float x, y, z;
};
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 operator*(float s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
__device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
__device__ __forceinline__ float norm(V3 a) { return sqrtf(dot(a, a)); }
// torch.nn.functional.normalize: x / max(|x|, 1e-12)
__device__ __forceinline__ V3 normalize12(V3 a) {
  const float n = fmaxf(norm(a), 1e-12f);
  return {a.x / n, a.y / n, a.z / n};
}
__device__ __forceinline__ V3 load3(const float* p) { return {p[0], p[1], p[2]}; }

// `_check_coplanar` / `_check_nonzero` of the reference on the detections: flags[i] bit 0 = not coplanar, bit 1 = a
// triangle with (almost) no area.  One thread per box.
__global__ void iou3d_check_kernel(const float* __restrict__ boxes, int n, float eps_coplanar, float eps_nonzero,
                                   int* __restrict__ flags) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* b = boxes + (size_t)i * 24;
  // sum over the six faces of (v3 - v0) . normal, THEN the absolute value (omni3d_evaluation.py:86-89: one bmm over P*3)
  float s = 0.f;
  for (int p = 0; p < 6; ++p) {
    const V3 v0 = load3(b + 3 * c_planes[p][0]), v1 = load3(b + 3 * c_planes[p][1]);
    const V3 v2 = load3(b + 3 * c_planes[p][2]), v3 = load3(b + 3 * c_planes[p][3]);
    const V3 nrm = normalize12(cross(normalize12(v1 - v0), normalize12(v2 - v0)));
    s += dot(v3 - v0, nrm);
  }
  int f = (fabsf(s) < eps_coplanar) ? 0 : 1;
  bool nonzero = true;
  for (int t = 0; t < 12; ++t) {
    const V3 v0 = load3(b + 3 * c_tris[t][0]), v1 = load3(b + 3 * c_tris[t][1]), v2 = load3(b + 3 * c_tris[t][2]);
    nonzero = nonzero && (norm(cross(v1 - v0, v2 - v0)) / 2.f > eps_nonzero);
  }
  if (!nonzero) f |= 2;
  flags[i] = f;
}

struct WarpScratch {
  float box[2][8][3];              // both boxes, translated to the first one's centre
  float plane[2][6][4];            // outward unit normal, offset d: inside is n . x <= d
  float poly[2][kMaxPoly][3][12];  // ping-pong polygons of the 12 face lanes, lane-minor
};

__global__ void __launch_bounds__(kIouWarps * 32)
iou3d_kernel(const float* __restrict__ dt, const float* __restrict__ gt, int N, int M, const int* __restrict__ dt_flags,
             float* __restrict__ iou, float* __restrict__ vol_out) {
  __shared__ WarpScratch scratch[kIouWarps];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  WarpScratch& s = scratch[warp];
  const long long pairs = (long long)N * M;
  for (long long pair = (long long)blockIdx.x * kIouWarps + warp; pair < pairs; pair += (long long)gridDim.x * kIouWarps) {
    const int i = (int)(pair / M), j = (int)(pair - (long long)i * M);
    if (dt_flags && dt_flags[i]) {   // offending detections get zero IoU (omni3d_evaluation.py:160-167)
      if (lane == 0) {
        iou[pair] = 0.f;
        if (vol_out) vol_out[pair] = 0.f;
      }
      continue;
    }
    __syncwarp();
    // ---- both boxes relative to the first one's centre
    const float* a = dt + (size_t)i * 24;
    const float* b = gt + (size_t)j * 24;
    float ca = lane < 24 ? a[lane] : 0.f;               // lane = 3 * vertex + coordinate
    float cx = ca;
    cx += __shfl_down_sync(0xffffffffu, cx, 12);        // (v, c) + (v + 4, c)
    cx += __shfl_down_sync(0xffffffffu, cx, 6);         // + (v + 2, c) ...
    cx += __shfl_down_sync(0xffffffffu, cx, 3);         // + (v + 1, c): lanes 0..2 hold the coordinate sums
    const float centre = __shfl_sync(0xffffffffu, cx, lane % 3) * 0.125f;
    if (lane < 24) {
      s.box[0][lane / 3][lane % 3] = ca - centre;
      s.box[1][lane / 3][lane % 3] = b[lane] - centre;
    }
    __syncwarp();
    // ---- the 12 face planes (lane f < 6: first box, 6 <= f < 12: second box) and the boxes' own volumes
    float own_vol = 0.f;
    if (lane < 12) {
      const int which = lane / 6, f = lane % 6;
      V3 v[4], c8 = {0.f, 0.f, 0.f};
      for (int k = 0; k < 8; ++k) c8 = c8 + load3(s.box[which][k]);
      c8 = 0.125f * c8;
      for (int k = 0; k < 4; ++k) v[k] = load3(s.box[which][c_planes[f][k]]);
      const V3 fc = 0.25f * (v[0] + v[1] + v[2] + v[3]);
      V3 nv = cross(v[2] - v[0], v[3] - v[1]);          // diagonals: area-weighted normal of a planar quad
      const float twice_area = norm(nv);
      nv = (1.f / fmaxf(twice_area, FLT_MIN)) * nv;
      if (dot(nv, fc - c8) < 0.f) nv = -1.f * nv;       // outward
      s.plane[which][f][0] = nv.x; s.plane[which][f][1] = nv.y; s.plane[which][f][2] = nv.z;
      s.plane[which][f][3] = dot(nv, fc);
      own_vol = dot(nv, fc - c8) * (0.5f * twice_area) * (1.f / 3.f);
      for (int k = 0; k < 4; ++k) {
        s.poly[0][k][0][lane] = v[k].x; s.poly[0][k][1][lane] = v[k].y; s.poly[0][k][2][lane] = v[k].z;
      }
    }
    __syncwarp();
    float va = lane < 6 ? own_vol : 0.f, vb = (lane >= 6 && lane < 12) ? own_vol : 0.f;
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) {
      va += __shfl_xor_sync(0xffffffffu, va, o);
      vb += __shfl_xor_sync(0xffffffffu, vb, o);
    }
    // ---- clip face `lane` against the six half-spaces of the other box
    float contrib = 0.f;
    if (lane < 12) {
      const int which = lane / 6, f = lane % 6, other = 1 - which;
      const float scale = cbrtf(fmaxf(fmaxf(va, vb), FLT_MIN));
      const float eps = 1e-6f * scale;                  // a vertex this close to a plane is on it
      const V3 nf = {s.plane[which][f][0], s.plane[which][f][1], s.plane[which][f][2]};
      const float df = s.plane[which][f][3];
      bool skip = false;
      if (which == 1) {   // same plane, same side as a face of the first box: that face already carries this polygon
        for (int p = 0; p < 6; ++p) {
          const V3 np = {s.plane[0][p][0], s.plane[0][p][1], s.plane[0][p][2]};
          skip = skip || (dot(np, nf) > 1.f - 1e-6f && fabsf(s.plane[0][p][3] - df) < 1e-5f * scale);
        }
      }
      int m = skip ? 0 : 4, cur = 0;
      for (int p = 0; p < 6 && m >= 3; ++p) {
        const V3 np = {s.plane[other][p][0], s.plane[other][p][1], s.plane[other][p][2]};
        const float dp = s.plane[other][p][3];
        int out = 0;
        V3 prev = {s.poly[cur][m - 1][0][lane], s.poly[cur][m - 1][1][lane], s.poly[cur][m - 1][2][lane]};
        float hprev = dot(np, prev) - dp;
        for (int k = 0; k < m; ++k) {
          const V3 q = {s.poly[cur][k][0][lane], s.poly[cur][k][1][lane], s.poly[cur][k][2][lane]};
          const float h = dot(np, q) - dp;
          const bool in_q = h <= eps, in_p = hprev <= eps;
          if (in_q != in_p && out < kMaxPoly) {         // the edge crosses the plane
            const float t = hprev / (hprev - h);
            const V3 x = prev + t * (q - prev);
            s.poly[cur ^ 1][out][0][lane] = x.x; s.poly[cur ^ 1][out][1][lane] = x.y; s.poly[cur ^ 1][out][2][lane] = x.z;
            ++out;
          }
          if (in_q && out < kMaxPoly) {
            s.poly[cur ^ 1][out][0][lane] = q.x; s.poly[cur ^ 1][out][1][lane] = q.y; s.poly[cur ^ 1][out][2][lane] = q.z;
            ++out;
          }
          prev = q;
          hprev = h;
        }
        m = out;
        cur ^= 1;
      }
      if (m >= 3) {
        const V3 q0 = {s.poly[cur][0][0][lane], s.poly[cur][0][1][lane], s.poly[cur][0][2][lane]};
        V3 area2 = {0.f, 0.f, 0.f};
        V3 e1 = V3{s.poly[cur][1][0][lane], s.poly[cur][1][1][lane], s.poly[cur][1][2][lane]} - q0;
        for (int k = 2; k < m; ++k) {
          const V3 e2 = V3{s.poly[cur][k][0][lane], s.poly[cur][k][1][lane], s.poly[cur][k][2][lane]} - q0;
          area2 = area2 + cross(e1, e2);
          e1 = e2;
        }
        contrib = df * (0.5f * norm(area2)) * (1.f / 3.f);   // (n_f . x) is d_f everywhere on the face
      }
    }
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) contrib += __shfl_xor_sync(0xffffffffu, contrib, o);
    if (lane == 0) {
      float vol = fminf(fmaxf(contrib, 0.f), fminf(va, vb));
      iou[pair] = vol / (va + vb - vol);
      if (vol_out) vol_out[pair] = vol;
    }
  }
}

}  // namespace

int iou_box3d_launch(const float* boxes_dt, const float* boxes_gt, int32_t N, int32_t M, float eps_coplanar,
                     float eps_nonzero, float* iou, float* vol, int32_t* dt_flags, cudaStream_t stream) {
  OVM_REQUIRE(N >= 0 && M >= 0, "ovm_iou_box3d: N=%d M=%d", N, M);
  if (N == 0 || M == 0) return OVM_OK;
  OVM_REQUIRE(boxes_dt && boxes_gt && iou, "ovm_iou_box3d: NULL pointer");
  if (dt_flags) {
    iou3d_check_kernel<<<(N + 127) / 128, 128, 0, stream>>>(boxes_dt, N, eps_coplanar, eps_nonzero, dt_flags);
    OVM_LAUNCHED("iou3d_check_kernel");
  }
  const long long pairs = (long long)N * M;
  long long blocks = (pairs + kIouWarps - 1) / kIouWarps;
  const long long cap = (long long)kNumSMs * 8;
  if (blocks > cap) blocks = cap;
  iou3d_kernel<<<(unsigned)blocks, kIouWarps * 32, 0, stream>>>(boxes_dt, boxes_gt, N, M, dt_flags, iou, vol);
  OVM_LAUNCHED("iou3d_kernel");
  return OVM_OK;
}

}  // namespace ovm
