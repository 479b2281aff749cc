// b200slam — Gauss-Newton Hessian build, stage 1 (per-point work) for sm_100a.
//
// Replaces ray_align_kernel / calib_proj_kernel / point_align_kernel of the reference
// (gn_kernels.cu:812-1137, :1230-1542, :454-722) with a restructured algebra:
//
//   The reference maps every residual row J0 (1x7, expressed in camera i) through
//   apply_Sim3_adj_inv(T_i) (gn_kernels.cu:276-296), which is LINEAR in J0 with an
//   edge-constant 7x7 matrix M, and sets Ji = -Jj (:998-999).  Hence, per edge,
//       H_jj = M^T (sum_k w J0^T J0) M = M^T H0 M,  H_ii = H_jj,  H_ij = H_ji = -H_jj,
//       g_j  = M^T (sum_k w r J0^T)   = M^T g0,     g_i  = -g_j.
//   Stage 1 therefore accumulates only H0 (28 unique entries) + g0 (7) per point, using the
//   sparsity of J0 (4-5 non-zeros per row): ~95 FMAs/point instead of ~1500 flops, which moves
//   the kernel from FP32-ALU-bound to HBM-bound.  M is applied once per edge in stage 2 (fp64).
//
// Execution shape: grid (S, Ecount): S CTAs split the n points of one edge; a thread handles
// groups of 4 consecutive points with 16-byte loads (valid u32, idx 2x16B, Q float4, Xj 3x
// float4, Cj float4) and scalar gathers for the matched point of keyframe i.  35 fp32
// accumulators per thread -> warp shuffle tree -> per-CTA fp64 partial written to the
// workspace (no atomics, fixed order => deterministic).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200slam.h"

namespace b200gn {

constexpr int BUILD_THREADS = 256;
constexpr int NACC = 35;  // 28 (upper triangle of 7x7 H0, row-major a<=b) + 7 (g0)

struct EdgePose {  // relative pose T_ij = T_i^-1 T_j  (relSim3, gn_kernels.cu:251-271)
  float t[3], q[4], s;
};

__device__ __forceinline__ void rotate(const float *q, const float *X, float *Y) {
  // actSO3, gn_kernels.cu:194-204
  const float uv0 = 2.0f * (q[1] * X[2] - q[2] * X[1]);
  const float uv1 = 2.0f * (q[2] * X[0] - q[0] * X[2]);
  const float uv2 = 2.0f * (q[0] * X[1] - q[1] * X[0]);
  Y[0] = X[0] + q[3] * uv0 + (q[1] * uv2 - q[2] * uv1);
  Y[1] = X[1] + q[3] * uv1 + (q[2] * uv0 - q[0] * uv2);
  Y[2] = X[2] + q[3] * uv2 + (q[0] * uv1 - q[1] * uv0);
}

__device__ __forceinline__ EdgePose rel_pose(const float *Ti, const float *Tj) {
  EdgePose P;
  const float si_inv = 1.0f / Ti[7];
  P.s = si_inv * Tj[7];
  const float qi[4] = {-Ti[3], -Ti[4], -Ti[5], Ti[6]};
  const float *qj = Tj + 3;
  P.q[0] = qi[3] * qj[0] + qi[0] * qj[3] + qi[1] * qj[2] - qi[2] * qj[1];
  P.q[1] = qi[3] * qj[1] - qi[0] * qj[2] + qi[1] * qj[3] + qi[2] * qj[0];
  P.q[2] = qi[3] * qj[2] + qi[0] * qj[1] - qi[1] * qj[0] + qi[2] * qj[3];
  P.q[3] = qi[3] * qj[3] - qi[0] * qj[0] - qi[1] * qj[1] - qi[2] * qj[2];
  const float d[3] = {Tj[0] - Ti[0], Tj[1] - Ti[1], Tj[2] - Ti[2]};
  rotate(qi, d, P.t);
  P.t[0] *= si_inv; P.t[1] *= si_inv; P.t[2] *= si_inv;
  return P;
}

// The Hessian contract is 1e-5 relative (not bit-exactness), so reciprocals, square roots and the
// Huber quotient use the single-MUFU approximations (<= 2 ulp) instead of the IEEE sequences
// (8-10 instructions each; ncu: the build kernel was 51 % issue-bound with the XU pipe at 18 %).
__device__ __forceinline__ float fast_sqrt(float x) {
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float huber_w(float r) {  // gn_kernels.cu:171-174
  const float a = fabsf(r);
  return a < 1.345f ? 1.0f : __fdividef(1.345f, a);
}

// acc += w * J^T J (upper triangle) and w*err*J, for the non-zero columns in MASK only.
template <unsigned MASK>
__device__ __forceinline__ void accum_row(float (&acc)[NACC], const float (&J)[7], float w, float err) {
  float wJ[7];
#pragma unroll
  for (int a = 0; a < 7; a++) wJ[a] = (MASK >> a & 1u) ? w * J[a] : 0.0f;
  int l = 0;
#pragma unroll
  for (int a = 0; a < 7; a++) {
#pragma unroll
    for (int b = a; b < 7; b++, l++)
      if ((MASK >> a & 1u) && (MASK >> b & 1u)) acc[l] = fmaf(wJ[a], J[b], acc[l]);
  }
#pragma unroll
  for (int a = 0; a < 7; a++)
    if (MASK >> a & 1u) acc[28 + a] = fmaf(wJ[a], err, acc[28 + a]);
}

struct BuildConsts {
  float sa_inv, sb_inv, C_thresh, Q_thresh;
  float fx, fy, cx, cy, z_eps, ulo, uhi, vlo, vhi;
  int width;
  const float *Kptr;  // device (3,3) intrinsics, read by the calib kernel itself (no host sync)
  // two-way edge lists without the caller's torch.cat (global_opt.py:104-110): directed edges e >= E1 read
  // their per-edge arrays from the second set (pointers pre-offset by -E1 rows); E1 = INT_MAX: one set only
  int E1 = 0x7fffffff;
  const int64_t *idx2 = nullptr;
  const uint8_t *valid2 = nullptr;
  const float *Q2 = nullptr;
  // 16-byte loads are legal: n % 4 == 0 AND every base pointer is 16-byte aligned (valid: 4-byte); a
  // contiguous view with an odd storage offset takes the scalar path instead of faulting
  int vec_ok = 0;
  // optional rank -> buffer slot map for Xs / Cs (b200slam_gn_params::kf_slots); nullptr = identity
  const int64_t *slots = nullptr;
};

// One point.  Xi/ci = matched point of keyframe i, Xj/cj = point k of keyframe j.
template <int KIND>
__device__ __forceinline__ void point_contrib(float (&acc)[NACC], const EdgePose &P, const BuildConsts &C,
                                              const float *Xi, float ci, const float *Xj, float cj,
                                              float q, long long ind) {
  if (!((q > C.Q_thresh) & (ci > C.C_thresh) & (cj > C.C_thresh))) return;
  float Pj[3];
  rotate(P.q, Xj, Pj);  // actSim3, gn_kernels.cu:206-218
  Pj[0] = Pj[0] * P.s + P.t[0]; Pj[1] = Pj[1] * P.s + P.t[1]; Pj[2] = Pj[2] * P.s + P.t[2];
  const float sq = fast_sqrt(q);
  if (KIND == B200SLAM_GN_RAYS) {
    const float n2i = Xi[0] * Xi[0] + Xi[1] * Xi[1] + Xi[2] * Xi[2];
    const float n1i_inv = rsqrtf(n2i), n1i = n2i * n1i_inv;
    const float n2j = Pj[0] * Pj[0] + Pj[1] * Pj[1] + Pj[2] * Pj[2];
    const float n1j_inv = rsqrtf(n2j), n1j = n2j * n1j_inv;
    const float r0 = n1j_inv * Pj[0], r1 = n1j_inv * Pj[1], r2 = n1j_inv * Pj[2];
    const float e0 = r0 - n1i_inv * Xi[0], e1 = r1 - n1i_inv * Xi[1], e2 = r2 - n1i_inv * Xi[2];
    const float e3 = n1j - n1i;
    const float swr = C.sa_inv * sq, swd = C.sb_inv * sq;
    const float wr = swr * swr, wd = swd * swd;
    const float w0 = huber_w(swr * e0) * wr, w1 = huber_w(swr * e1) * wr, w2 = huber_w(swr * e2) * wr;
    const float w3 = huber_w(swd * e3) * wd;
    // d r / d P = (I - r r^T)/|P|   (gn_kernels.cu:981-987)
    const float dxx = n1j_inv - r0 * r0 * n1j_inv, dyy = n1j_inv - r1 * r1 * n1j_inv,
                dzz = n1j_inv - r2 * r2 * n1j_inv;
    const float dxy = -r0 * r1 * n1j_inv, dxz = -r0 * r2 * n1j_inv, dyz = -r1 * r2 * n1j_inv;
    { const float J[7] = {dxx, dxy, dxz, 0.f, r2, -r1, 0.f}; accum_row<0x37>(acc, J, w0, e0); }
    { const float J[7] = {dxy, dyy, dyz, -r2, 0.f, r0, 0.f}; accum_row<0x2F>(acc, J, w1, e1); }
    { const float J[7] = {dxz, dyz, dzz, r1, -r0, 0.f, 0.f}; accum_row<0x1F>(acc, J, w2, e2); }
    { const float J[7] = {r0, r1, r2, 0.f, 0.f, 0.f, n1j}; accum_row<0x47>(acc, J, w3, e3); }
  } else if (KIND == B200SLAM_GN_CALIB) {
    if (!((Pj[2] > C.z_eps) && (Xi[2] > C.z_eps))) return;  // valid_z, :1367
    const float zinv = __fdividef(1.0f, Pj[2]);
    const float xz = Pj[0] * zinv, yz = Pj[1] * zinv;
    const float u = C.fx * xz + C.cx, v = C.fy * yz + C.cy;
    if (!((u > C.ulo) && (u < C.uhi) && (v > C.vlo) && (v < C.vhi))) return;  // :1381-1382
    const int ut = (int)(ind % C.width), vt = (int)(ind / C.width);          // :1360-1361
    const float e0 = u - (float)ut, e1 = v - (float)vt, e2 = logf(Pj[2]) - logf(Xi[2]);
    const float swp = C.sa_inv * sq, swd = C.sb_inv * sq;
    const float w0 = huber_w(swp * e0) * (swp * swp), w1 = huber_w(swp * e1) * (swp * swp),
                w2 = huber_w(swd * e2) * (swd * swd);
    { const float J[7] = {C.fx * zinv, 0.f, -C.fx * xz * zinv, -C.fx * xz * yz, C.fx * (1.f + xz * xz),
                          -C.fx * yz, 0.f};
      accum_row<0x3D>(acc, J, w0, e0); }
    { const float J[7] = {0.f, C.fy * zinv, -C.fy * yz * zinv, -C.fy * (1.f + yz * yz), C.fy * xz * yz,
                          C.fy * xz, 0.f};
      accum_row<0x3E>(acc, J, w1, e1); }
    { const float J[7] = {0.f, 0.f, zinv, yz, -xz, 0.f, 1.f}; accum_row<0x5C>(acc, J, w2, e2); }
  } else {
    const float e0 = Pj[0] - Xi[0], e1 = Pj[1] - Xi[1], e2 = Pj[2] - Xi[2];
    const float sw = C.sa_inv * sq, wc = sw * sw;
    const float w0 = huber_w(sw * e0) * wc, w1 = huber_w(sw * e1) * wc, w2 = huber_w(sw * e2) * wc;
    { const float J[7] = {1.f, 0.f, 0.f, 0.f, Pj[2], -Pj[1], Pj[0]}; accum_row<0x71>(acc, J, w0, e0); }
    { const float J[7] = {0.f, 1.f, 0.f, -Pj[2], 0.f, Pj[0], Pj[1]}; accum_row<0x6A>(acc, J, w1, e1); }
    { const float J[7] = {0.f, 0.f, 1.f, Pj[1], -Pj[0], 0.f, Pj[2]}; accum_row<0x5C>(acc, J, w2, e2); }
  }
}

}  // namespace b200gn
