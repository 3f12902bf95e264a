// The cube decode of one box from its selected head outputs: shared by head_decode_kernel (one warp per box,
// head_decode.cu) and fc_tail_kernel (fc2 + heads + decode in one launch, one thread per box, fc_tail.cu).
//
// Replaces (reference paths relative to the checkout):
//   * pose -> rotation matrix                        cubercnn/modeling/roi_heads/cube_head.py:176-186 with pytorch3d
//                                                    rotation_6d_to_matrix / quaternion_to_matrix /
//                                                    euler_angles_to_matrix("XYZ")
//   * per-box camera / scale prep                    roi_heads.py:378-408, math_util.py:581-592
//   * xy, dims, allocentric->egocentric, z           roi_heads.py:414-419,480-549, math_util.py:651-679,969-978
//   * back-projection, confidence, score merge       roi_heads.py:798-829
//   * the 8 corners                                  math_util.py:116-193 (get_cuboid_verts_faces)
// No fast-math: acosf/sinf/cosf/expf are the accurate versions, divisions are IEEE, and NaNs propagate exactly where
// the reference lets them.
#pragma once

#include <climits>
#include <cmath>

#include "common.cuh"

namespace ovm {

constexpr int kMaxHeadRows = 13;   // deltas 2 + dims 3 + pose <= 6 + z 1 + uncertainty 1

struct DecodeArgs {
  ovm_decode_cfg_t cfg;
  ovm_head_weights_t w;
  const float* feats;
  long long ld_feats;
  const float* boxes;
  const long long* classes;
  const float* scores;
  const int* batch_idx;
  const float* cams;
  int num_images, n;
  float* records;
  float* raw;
  unsigned long long* trace;   // optional [grid, 8] %globaltimer stamps of warp 0 (ovm_head_set_trace), else NULL
  int* error_flag;             // optional device int: set to 1 when a class id / image index is out of range
  // fused gather (ovm_lift_forward_peers): every record is also stored into these buffers -- the gather slots of this
  // rank in every rank's exchange buffer, mapped into this device's address space (NVLink peer mappings)
  float* peer_records[OVM_MAX_PEERS];
  int num_peers;
};

__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + expf(-x)); }
// torch .clip(min): NaN-transparent, unlike fmaxf
__device__ __forceinline__ float clip_min(float x, float lo) { return x < lo ? lo : x; }

// pytorch3d quaternion_to_matrix (real part first), row-major 3x3.
__device__ __forceinline__ void quat_to_matrix(float r, float i, float j, float k, float* R) {
  const float two_s = 2.f / (r * r + i * i + j * j + k * k);
  R[0] = 1.f - two_s * (j * j + k * k);
  R[1] = two_s * (i * j - k * r);
  R[2] = two_s * (i * k + j * r);
  R[3] = two_s * (i * j + k * r);
  R[4] = 1.f - two_s * (i * i + k * k);
  R[5] = two_s * (j * k - i * r);
  R[6] = two_s * (i * k - j * r);
  R[7] = two_s * (j * k + i * r);
  R[8] = 1.f - two_s * (i * i + j * j);
}

__device__ __forceinline__ void matmul3(const float* A, const float* B, float* C) {
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int c = 0; c < 3; ++c) C[3 * r + c] = A[3 * r] * B[c] + A[3 * r + 1] * B[3 + c] + A[3 * r + 2] * B[6 + c];
}

// What a box needs from the call's inputs (not from the FC stack): loaded ahead of the wait on the FC kernel.
// `valid` is false when the class id or the image index is out of range (the reference raises IndexError there):
// such a box is decoded to a NaN record and the call's error flag is raised.
struct BoxIn {
  float x1, y1, x2, y2;
  float k00, k02, k11, k12, ratio, h_cur;
  float score;
  int cls;
  bool valid;
};
__device__ __forceinline__ BoxIn load_box_in(const DecodeArgs& a, int box) {
  BoxIn b;
  int image = a.batch_idx[box];
  const long long cls = a.classes[box];
  const bool uses_class = a.cfg.dims_priors_enabled || a.cfg.cluster_bins > 1;
  b.valid = image >= 0 && image < a.num_images && (!uses_class || (cls >= 0 && cls < a.cfg.num_classes));
  if (!b.valid) image = 0;
  b.cls = b.valid ? (int)cls : 0;
  b.x1 = a.boxes[4 * box + 0]; b.y1 = a.boxes[4 * box + 1];
  b.x2 = a.boxes[4 * box + 2]; b.y2 = a.boxes[4 * box + 3];
  b.score = a.scores ? a.scores[box] : 0.f;
  const float* cam = a.cams + 12 * image;
  b.k00 = cam[0]; b.k02 = cam[2]; b.k11 = cam[4]; b.k12 = cam[5];
  b.ratio = cam[9]; b.h_cur = cam[10];
  if (!b.valid) b.cls = (int)(cls < INT_MIN ? INT_MIN : (cls > INT_MAX ? INT_MAX : cls));   // reported in the record
  return b;
}

// ---- the decode in three parts, so that fc_tail_kernel can run the two independent chains of a box on two threads.
// decode_record (below) runs them in sequence on one thread: the arithmetic is the same either way.

struct BoxCam {   // derived from the call's inputs only (roi_heads.py:378-419): computed before the wait on the FC layer
  float src_w, src_h, ctr_x, ctr_y, fx, fy, px, py, virtual_to_real;
};
__device__ __forceinline__ BoxCam box_camera(const ovm_decode_cfg_t& cfg, const BoxIn& in) {
  BoxCam c;
  // roi_heads.py:414-419
  c.src_w = in.x2 - in.x1; c.src_h = in.y2 - in.y1;
  c.ctr_x = in.x1 + 0.5f * c.src_w; c.ctr_y = in.y1 + 0.5f * c.src_h;
  // roi_heads.py:378-408: K / ratio with [2,2] forced to 1; focal is the un-scaled K[1,1]
  const float ratio = in.ratio, h_cur = in.h_cur;
  c.fx = in.k00 / ratio; c.fy = in.k11 / ratio; c.px = in.k02 / ratio; c.py = in.k12 / ratio;
  const float focal = in.k11;
  const float h_orig = h_cur * ratio;
  c.virtual_to_real = cfg.virtual_depth ? (h_cur * focal) / (cfg.virtual_focal * h_orig) : 1.f;
  return c;
}

// Chain A: raw pose -> view rotation R (rows; cube_head.py:176-186) and raw dims -> dims (roi_heads.py:483-507).
__device__ __forceinline__ void decode_pose_dims(const DecodeArgs& a, const BoxIn& in, const float (&pose_raw)[6],
                                                 const float (&dims_raw)[3], float (&R)[9], float (&dims)[3]) {
  const ovm_decode_cfg_t& cfg = a.cfg;
  const bool per_class = cfg.dims_priors_enabled != 0;
  const int cls = in.valid ? in.cls : 0;
  if (cfg.pose_type == OVM_POSE_6D) {
    const float a1x = pose_raw[0], a1y = pose_raw[1], a1z = pose_raw[2];
    const float a2x = pose_raw[3], a2y = pose_raw[4], a2z = pose_raw[5];
    const float n1 = fmaxf(sqrtf(a1x * a1x + a1y * a1y + a1z * a1z), 1e-12f);
    const float b1x = a1x / n1, b1y = a1y / n1, b1z = a1z / n1;
    const float d = b1x * a2x + b1y * a2y + b1z * a2z;
    float b2x = a2x - d * b1x, b2y = a2y - d * b1y, b2z = a2z - d * b1z;
    const float n2 = fmaxf(sqrtf(b2x * b2x + b2y * b2y + b2z * b2z), 1e-12f);
    b2x /= n2; b2y /= n2; b2z /= n2;
    R[0] = b1x; R[1] = b1y; R[2] = b1z;
    R[3] = b2x; R[4] = b2y; R[5] = b2z;
    R[6] = b1y * b2z - b1z * b2y;
    R[7] = b1z * b2x - b1x * b2z;
    R[8] = b1x * b2y - b1y * b2x;
  } else if (cfg.pose_type == OVM_POSE_QUATERNION) {
    // quats / _copysign(sqrt(sum q^2), q0): flips the sign of the norm when q0 < 0
    const float q0 = pose_raw[0], q1 = pose_raw[1], q2 = pose_raw[2], q3 = pose_raw[3];
    float nq = sqrtf(q0 * q0 + q1 * q1 + q2 * q2 + q3 * q3);
    if ((nq < 0.f) != (q0 < 0.f)) nq = -nq;
    quat_to_matrix(q0 / nq, q1 / nq, q2 / nq, q3 / nq, R);
  } else {
    // euler_angles_to_matrix(., "XYZ") = Rx(a0) @ Ry(a1) @ Rz(a2)
    const float cx = cosf(pose_raw[0]), sx = sinf(pose_raw[0]);
    const float cy = cosf(pose_raw[1]), sy = sinf(pose_raw[1]);
    const float cz = cosf(pose_raw[2]), sz = sinf(pose_raw[2]);
    const float Rx[9] = {1.f, 0.f, 0.f, 0.f, cx, -sx, 0.f, sx, cx};
    const float Ry[9] = {cy, 0.f, sy, 0.f, 1.f, 0.f, -sy, 0.f, cy};
    const float Rz[9] = {cz, -sz, 0.f, sz, cz, 0.f, 0.f, 0.f, 1.f};
    float T[9];
    matmul3(Rx, Ry, T);
    matmul3(T, Rz, R);
  }
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    if (per_class) {
      const float mean = a.w.priors_dims_per_cat[(long long)cls * 6 + j];
      const float sd = a.w.priors_dims_per_cat[(long long)cls * 6 + 3 + j];
      if (cfg.dims_priors_func == OVM_DIMS_SIGMOID) {
        const float lo = clip_min(mean - 3.f * sd, 0.f), hi = mean + 3.f * sd;
        dims[j] = lo + (hi - lo) * sigmoidf_(dims_raw[j]);
      } else {
        dims[j] = expf(dims_raw[j] > 5.f ? 5.f : dims_raw[j]) * mean;
      }
    } else {
      dims[j] = expf(dims_raw[j] > 5.f ? 5.f : dims_raw[j]);
    }
  }
}

// Chain B: deltas -> projected centre, allocentric rotation M (math_util.py:657-679; has_m false: the view rotation is
// kept), z (roi_heads.py:515-549), back-projection, confidence and score (roi_heads.py:802-829).
struct CentreOut {
  float M[9];
  int has_m;
  float X, Y, z, u_out, v_out, score;
};
__device__ __forceinline__ CentreOut decode_centre(const DecodeArgs& a, const BoxIn& in, const BoxCam& c, const int bin,
                                                   const float dx, const float dy, const float z_raw, const float uncert) {
  const ovm_decode_cfg_t& cfg = a.cfg;
  const int cls = in.valid ? in.cls : 0;
  CentreOut o;
  // ---- xy (roi_heads.py:480-481)
  const float u = c.ctr_x + c.src_w * dx;
  const float v = c.ctr_y + c.src_h * dy;
  // ---- allocentric -> egocentric (math_util.py:657-679)
  o.has_m = 0;
  if (cfg.allocentric_pose) {
    float rx = (u - c.px) / c.fx, ry = (v - c.py) / c.fy, rz = 1.f;
    const float rn = sqrtf(rx * rx + ry * ry + rz * rz);
    rx /= rn; ry /= rn; rz /= rn;
    const float angle = acosf(rz);
    if (angle > 0.f) {  // false for 0 and NaN: R_view is kept (boolean-mask assignment :678-679)
      const float ax = 0.f - ry, ay = 0.f + rx;
      const float an = sqrtf(ax * ax + ay * ay + 0.f);
      const float vx = angle * ax / an, vy = angle * ay / an, vz = angle * 0.f / an;
      // pytorch3d axis_angle_to_quaternion -> quaternion_to_matrix
      const float angles = sqrtf(vx * vx + vy * vy + vz * vz);
      const float half = angles * 0.5f;
      const float s = fabsf(angles) < 1e-6f ? 0.5f - (angles * angles) / 48.f : sinf(half) / angles;
      quat_to_matrix(cosf(half), vx * s, vy * s, vz * s, o.M);
      o.has_m = 1;
    }
  }
  // ---- z (roi_heads.py:515-549)
  float z = z_raw;
  if (cfg.z_type == OVM_Z_SIGMOID) {
    z = sigmoidf_(z) * 100.f;
  } else if (cfg.z_type == OVM_Z_LOG) {
    z = expf(z);
  } else if (cfg.z_type == OVM_Z_CLUSTERS) {
    const float* st = a.w.priors_z_stats + ((long long)cls * cfg.cluster_bins + bin) * 2;
    const float zlo = clip_min(st[0] - 3.f * st[1], 0.f), zhi = st[0] + 3.f * st[1];
    z = zlo + (zhi - zlo) * sigmoidf_(z);
  }
  if (cfg.virtual_depth) z = z * c.virtual_to_real;
  // ---- back-projection, confidence, score (roi_heads.py:802-829)
  o.X = z * (u - c.px) / c.fx;
  o.Y = z * (v - c.py) / c.fy;
  o.z = z;
  o.u_out = u * in.ratio;
  o.v_out = v * in.ratio;
  const float last = cfg.use_confidence ? expf(-uncert) : o.v_out;  // cube_3D[:, -1]
  o.score = a.scores ? sqrtf(in.score * last) : last;
  return o;
}

// Corners (math_util.py:172-193: x <-> L, y <-> H, z <-> W; box order is X,Y,Z,W,H,L), the 48-float record, the optional
// raw head outputs, the optional peer copies.
__device__ __forceinline__ void write_record(const DecodeArgs& a, const int box, const BoxIn& in, const float (&R)[9],
                                             const float (&dims)[3], const float X, const float Y, const float z,
                                             const float u_out, const float v_out, const float score,
                                             const float (&raw13)[13], const bool store_peers) {
  const float hw = dims[0] / 2.f, hh = dims[1] / 2.f, hl = dims[2] / 2.f;
  float rec[OVM_RECORD_WIDTH];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const float lx = (k == 0 || k == 3 || k == 4 || k == 7) ? -hl : hl;
    const float ly = (k == 0 || k == 1 || k == 4 || k == 5) ? -hh : hh;
    const float lz = (k < 4) ? -hw : hw;
    rec[3 * k + 0] = (R[0] * lx + R[1] * ly + R[2] * lz) + X;
    rec[3 * k + 1] = (R[3] * lx + R[4] * ly + R[5] * lz) + Y;
    rec[3 * k + 2] = (R[6] * lx + R[7] * ly + R[8] * lz) + z;
  }
  rec[24] = X; rec[25] = Y; rec[26] = z;
  rec[27] = u_out; rec[28] = v_out;
  rec[29] = dims[0]; rec[30] = dims[1]; rec[31] = dims[2];
#pragma unroll
  for (int i = 0; i < 9; ++i) rec[32 + i] = R[i];
  rec[41] = in.x1; rec[42] = in.y1; rec[43] = in.x2; rec[44] = in.y2;
  rec[45] = score;
  rec[46] = (float)in.cls;
  rec[47] = 0.f;
  if (!in.valid) {   // class id / image index out of range: the reference raises IndexError; here NaN + error flag
    const float qnan = __int_as_float(0x7fc00000);
#pragma unroll
    for (int i = 0; i < 41; ++i) rec[i] = qnan;
    rec[45] = qnan;
    if (a.error_flag) *a.error_flag = 1;
  }
  float4* out = reinterpret_cast<float4*>(a.records + (long long)box * OVM_RECORD_WIDTH);
#pragma unroll
  for (int i = 0; i < OVM_RECORD_WIDTH / 4; ++i) out[i] = make_float4(rec[4 * i], rec[4 * i + 1], rec[4 * i + 2], rec[4 * i + 3]);
  for (int p = 0; store_peers && p < a.num_peers; ++p) {   // the exchange: 192 bytes per box and peer, posted writes over NVLink
    float4* po = reinterpret_cast<float4*>(a.peer_records[p] + (long long)box * OVM_RECORD_WIDTH);
#pragma unroll
    for (int i = 0; i < OVM_RECORD_WIDTH / 4; ++i) po[i] = make_float4(rec[4 * i], rec[4 * i + 1], rec[4 * i + 2], rec[4 * i + 3]);
  }
  if (a.raw) {   // deltas 2, z, dims 3, uncertainty (clipped), pose <= 6
    float4* ro = reinterpret_cast<float4*>(a.raw + (long long)box * 16);
    ro[0] = make_float4(raw13[0], raw13[1], raw13[2], raw13[3]);
    ro[1] = make_float4(raw13[4], raw13[5], raw13[6], raw13[7]);
    ro[2] = make_float4(raw13[8], raw13[9], raw13[10], raw13[11]);
    ro[3] = make_float4(raw13[12], 0.f, 0.f, 0.f);
  }
}

// hv = the selected head outputs (bias added) in the row order deltas 0-1, dims 2-4, z 5, uncertainty 6 (if
// USE_CONFIDENCE), pose after that  ->  the pieces the chains take
struct HeadOut {
  float deltas[2], dims_raw[3], z_raw, uncert, pose_raw[6];
};
__device__ __forceinline__ HeadOut split_heads(const ovm_decode_cfg_t& cfg, const float (&hv)[kMaxHeadRows]) {
  const int pose_n = cfg.pose_type == OVM_POSE_6D ? 6 : (cfg.pose_type == OVM_POSE_QUATERNION ? 4 : 3);
  const int conf_rows = cfg.use_confidence ? 1 : 0;
  HeadOut h;
  h.deltas[0] = hv[0]; h.deltas[1] = hv[1];
  h.dims_raw[0] = hv[2]; h.dims_raw[1] = hv[3]; h.dims_raw[2] = hv[4];
  h.z_raw = hv[5];
  h.uncert = conf_rows ? clip_min(hv[6], 0.01f) : 0.f;   // .clip(0.01), cube_head.py:164
#pragma unroll
  for (int j = 0; j < 6; ++j) h.pose_raw[j] = j < pose_n ? (conf_rows ? hv[7 + j] : hv[6 + j]) : 0.f;
  return h;
}

// Decode one box on one thread.  `bin` = the box's depth bin (roi_heads.py:436-444; 0 when CLUSTER_BINS == 1).  Every
// calling thread computes everything; the record (and the optional raw head outputs) are written only when `write`.
__device__ __forceinline__ void decode_record(const DecodeArgs& a, const int box, const BoxIn& in, const int bin,
                                              const float (&hv)[kMaxHeadRows], const bool write,
                                              const bool store_peers = true) {
  const HeadOut h = split_heads(a.cfg, hv);
  const BoxCam c = box_camera(a.cfg, in);
  float R[9], dims[3];
  decode_pose_dims(a, in, h.pose_raw, h.dims_raw, R, dims);
  const CentreOut o = decode_centre(a, in, c, bin, h.deltas[0], h.deltas[1], h.z_raw, h.uncert);
  if (o.has_m) {
    float T[9];
    matmul3(o.M, R, T);
#pragma unroll
    for (int i = 0; i < 9; ++i) R[i] = T[i];
  }
  if (!write) return;
  const float raw13[13] = {h.deltas[0], h.deltas[1], h.z_raw, h.dims_raw[0], h.dims_raw[1], h.dims_raw[2], h.uncert,
                           h.pose_raw[0], h.pose_raw[1], h.pose_raw[2], h.pose_raw[3], h.pose_raw[4], h.pose_raw[5]};
  write_record(a, box, in, R, dims, o.X, o.Y, o.z, o.u_out, o.v_out, o.score, raw13, store_peers);
}

}  // namespace ovm
