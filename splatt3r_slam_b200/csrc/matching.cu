// b200slam — matching kernels for sm_100a: iter_proj and refine_matches.
//
// Both kernels are bit-exact re-implementations of the reference's arithmetic
// (splatt3r_slam/backend/src/matching_kernels.cu) with a different execution shape:
// 128-thread CTAs (the reference uses 16), bounds guards, vector loads, 64-bit offsets,
// explicit stream.  The rounding sequence of iter_proj is pinned with explicit
// __fmaf_rn/__fmul_rn/... intrinsics so that neither NVVM nor ptxas can re-associate it:
// it reproduces, operation for operation, the FFMA/FMUL/FADD/DMUL sequence of the sm_100a
// SASS that nvcc 12.9 emits for the reference kernel (documented in oracle/oracle_matching.c).
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200slam.h"

namespace {

__device__ __forceinline__ float clampf(float x, float lo, float hi) {
  return fminf(fmaxf(x, lo), hi);
}

struct Weights {
  float w11, w12, w21, w22;
};

// matching_kernels.cu:156-163.  The reference forms (1.0-du)*dv etc. in fp64 and rounds to fp32.
// u, v >= 1 after the clamp, so du, dv are multiples of 2^-23 in [0,1): 1-du is EXACT in fp32, and
// the fp64 product of two fp32 values is exact, so its rounding to fp32 equals one IEEE fp32
// multiply.  The weights are therefore bit-identical without any fp64 instruction.
__device__ __forceinline__ Weights bilinear_weights(float du, float dv) {
  Weights k;
  const float omdu = __fsub_rn(1.0f, du), omdv = __fsub_rn(1.0f, dv);
  k.w11 = __fmul_rn(du, dv);
  k.w12 = __fmul_rn(omdu, dv);
  k.w21 = __fmul_rn(du, omdv);
  k.w22 = __fmul_rn(omdu, omdv);
  return k;
}

// w11*a11 + w12*a12 + w21*a21 + w22*a22 in the reference's fused order.
__device__ __forceinline__ float tap(const Weights &k, float a11, float a12, float a21, float a22) {
  float s = __fmul_rn(k.w12, a12);
  s = __fmaf_rn(k.w11, a11, s);
  s = __fmaf_rn(k.w21, a21, s);
  s = __fmaf_rn(k.w22, a22, s);
  return s;
}

// a0*b0 + a1*b1 + a2*b2 : FMUL(a1,b1), FFMA(a0,b0), FFMA(a2,b2)
__device__ __forceinline__ float dot3(float a0, float a1, float a2, float b0, float b1, float b2) {
  float s = __fmul_rn(a1, b1);
  s = __fmaf_rn(a0, b0, s);
  s = __fmaf_rn(a2, b2, s);
  return s;
}

constexpr int IP_THREADS = 128;

// Pre-pass: (h*w, 9) interleaved -> 9 planes of h*w floats (per batch row), through shared memory
// so both the read and the write are fully coalesced.  In the planar layout the 32 lanes of a
// warp (neighbouring query pixels -> neighbouring targets) touch 1-2 cache lines per load instead
// of ~10 in the interleaved layout, which is what bounds this kernel (L1 wavefronts).
__global__ void __launch_bounds__(IP_THREADS)
repack_planar_kernel(const float *__restrict__ rays_img, float *__restrict__ planes, int hw) {
  __shared__ float sm[IP_THREADS * 9];
  const size_t img_off = (size_t)blockIdx.y * hw * 9;
  const int p0 = blockIdx.x * IP_THREADS;
  const int npx = min(IP_THREADS, hw - p0);
  const float *src = rays_img + img_off + (size_t)p0 * 9;
  for (int t = threadIdx.x; t < npx * 9; t += IP_THREADS) sm[t] = __ldg(src + t);
  __syncthreads();
  if (threadIdx.x < npx) {
    float *dst = planes + img_off + p0 + threadIdx.x;
#pragma unroll
    for (int c = 0; c < 9; c++) dst[(size_t)c * hw] = sm[threadIdx.x * 9 + c];
  }
}

// One thread per query pixel; planes (9, h*w) of batch row blockIdx.y.
__global__ void __launch_bounds__(IP_THREADS)
iter_proj_kernel(const float *__restrict__ planes, const float *__restrict__ pts,
                 const float *__restrict__ p_init, float *__restrict__ p_new,
                 uint8_t *__restrict__ converged, int h, int w, int n, int max_iter,
                 float lambda_init, float cost_thresh) {
  const int q = blockIdx.x * IP_THREADS + threadIdx.x;
  if (q >= n) return;
  const size_t bq = (size_t)blockIdx.y * n + q;
  const int hw = h * w;
  const float *img = planes + (size_t)blockIdx.y * hw * 9;

  const float2 pi = __ldg(reinterpret_cast<const float2 *>(p_init) + bq);
  const float umax = (float)(w - 2), vmax = (float)(h - 2);
  float u = clampf(pi.x, 1.0f, umax);
  float v = clampf(pi.y, 1.0f, vmax);
  const float p0 = __ldg(pts + bq * 3 + 0), p1 = __ldg(pts + bq * 3 + 1), p2 = __ldg(pts + bq * 3 + 2);

  float lambda = lambda_init;
  bool conv = false;
  for (int it = 0; it < max_iter; it++) {
    const int u11 = (int)floorf(u), v11 = (int)floorf(v);
    const Weights k = bilinear_weights(__fsub_rn(u, (float)u11), __fsub_rn(v, (float)v11));
    const float *top = img + v11 * w + u11;  // (v,u); +1 = (v,u+1); +w = (v+1,u); +w+1 = (v+1,u+1)
    float c[9];
#pragma unroll
    for (int j = 0; j < 9; j++) {
      const float *pj = top + (size_t)j * hw;
      c[j] = tap(k, __ldg(pj + w + 1), __ldg(pj + w), __ldg(pj + 1), __ldg(pj));
    }
    const float nrm = __fsqrt_rn(dot3(c[0], c[1], c[2], c[0], c[1], c[2]));
    const float inv = __frcp_rn(nrm);
    const float e0 = __fmaf_rn(c[0], inv, -p0), e1 = __fmaf_rn(c[1], inv, -p1),
                e2 = __fmaf_rn(c[2], inv, -p2);
    const float cost = dot3(e0, e1, e2, e0, e1, e2);
    float A00 = dot3(c[3], c[4], c[5], c[3], c[4], c[5]);
    const float A01 = dot3(c[3], c[4], c[5], c[6], c[7], c[8]);
    float A11 = dot3(c[6], c[7], c[8], c[6], c[7], c[8]);
    const float s0 = dot3(e0, e1, e2, c[3], c[4], c[5]);
    const float s1 = dot3(e0, e1, e2, c[6], c[7], c[8]);
    A00 = __fadd_rn(A00, lambda);
    A11 = __fadd_rn(A11, lambda);
    const float det = __fmaf_rn(A00, A11, -__fmul_rn(A01, A01));
    const float dinv = __frcp_rn(det);
    const float nu = __fmaf_rn(A01, s1, -__fmul_rn(s0, A11));
    const float nv = __fmaf_rn(A01, s0, -__fmul_rn(s1, A00));
    const float un = clampf(__fmaf_rn(nu, dinv, u), 1.0f, umax);
    const float vn = clampf(__fmaf_rn(nv, dinv, v), 1.0f, vmax);

    const int u2 = (int)floorf(un), v2 = (int)floorf(vn);
    const Weights k2 = bilinear_weights(__fsub_rn(un, (float)u2), __fsub_rn(vn, (float)v2));
    const float *top2 = img + v2 * w + u2;
    float d[3];
#pragma unroll
    for (int j = 0; j < 3; j++) {
      const float *pj = top2 + (size_t)j * hw;
      d[j] = tap(k2, __ldg(pj + w + 1), __ldg(pj + w), __ldg(pj + 1), __ldg(pj));
    }
    const float nrm2 = __fsqrt_rn(dot3(d[0], d[1], d[2], d[0], d[1], d[2]));
    const float inv2 = __frcp_rn(nrm2);
    const float f0 = __fmaf_rn(d[0], inv2, -p0), f1 = __fmaf_rn(d[1], inv2, -p1),
                f2 = __fmaf_rn(d[2], inv2, -p2);
    const float new_cost = dot3(f0, f1, f2, f0, f1, f2);
    if (new_cost < cost) {
      u = un;
      v = vn;
      lambda = __double2float_rn(__dmul_rn((double)lambda, 0.1));  // matching_kernels.cu:261
      conv = new_cost < cost_thresh;
    } else {
      lambda = __fmul_rn(lambda, 10.0f);
      conv = cost < cost_thresh;
    }
  }
  reinterpret_cast<float2 *>(p_new)[bq] = make_float2(u, v);
  converged[bq] = conv ? 1 : 0;
}

}  // namespace

extern "C" int b200slam_version(void) { return 100; }

extern "C" size_t b200slam_iter_proj_workspace_bytes(int b, int h, int w) {
  if (b < 0 || h < 0 || w < 0) return 0;
  return (size_t)b * h * w * 9 * sizeof(float);
}

extern "C" int b200slam_iter_proj(const float *rays_img, const float *pts_3d_norm,
                                  const float *p_init, float *p_new, uint8_t *converged, int b,
                                  int h, int w, int n, int max_iter, float lambda_init,
                                  float cost_thresh, void *workspace, size_t workspace_bytes,
                                  void *stream) {
  if (b < 0 || n < 0) return B200SLAM_EINVAL;
  if (b == 0 || n == 0) return B200SLAM_OK;  // empty inputs: nothing to do (pointers may be null)
  if (!rays_img || !pts_3d_norm || !p_init || !p_new || !converged) return B200SLAM_EINVAL;
  if (h < 3 || w < 3 || (long long)h * w * 9 > 0x7fffffffLL) return B200SLAM_EINVAL;
  if (!workspace || workspace_bytes < b200slam_iter_proj_workspace_bytes(b, h, w)) return B200SLAM_EWORKSPACE;
  cudaStream_t st = (cudaStream_t)stream;
  float *planes = (float *)workspace;
  const int hw = h * w;
  repack_planar_kernel<<<dim3((hw + IP_THREADS - 1) / IP_THREADS, b), IP_THREADS, 0, st>>>(rays_img, planes, hw);
  dim3 grid((n + IP_THREADS - 1) / IP_THREADS, b);
  iter_proj_kernel<<<grid, IP_THREADS, 0, st>>>(planes, pts_3d_norm, p_init, p_new, converged, h, w, n,
                                                max_iter, lambda_init, cost_thresh);
  return (int)cudaGetLastError();
}
