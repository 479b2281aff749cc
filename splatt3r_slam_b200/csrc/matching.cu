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
#ifndef IP_MIN_CTAS
#define IP_MIN_CTAS 6  // resident CTAs per SM the LM kernels are compiled for (A/B r02: 4 = 96 registers, 6 = 80, 8 = 64 with 100 B of spills)
#endif

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

#include "iter_proj_kernel.cuh"
#include "match_fused.cuh"

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

// ---- allocation-free iter_proj (no counterpart in the reference) ----
// For callers that cannot allocate (CUDA graph capture, tight serving loops): the planar copy of the ray image lives
// in a file-scope device array instead of a caller-provided workspace.  Every GPU that loads the library holds these
// B200SLAM_ITER_PROJ_STATIC_MAX_BYTES until exit.  Nothing in it has to start at zero (repack_planar_kernel
// overwrites every word iter_proj_kernel reads), but ALL calls on a device share it: calls that overlap on the GPU
// (different streams) race.
__device__ __align__(256) float g_iter_proj_static_planes[B200SLAM_ITER_PROJ_STATIC_MAX_BYTES / sizeof(float)];

namespace {
// In host code the array's name is only a handle: the device address comes from cudaGetSymbolAddress, looked up once
// per device with that device current.
int iter_proj_static_planes(float **out) {
  static float *addr[64] = {};
  int d = 0;
  cudaError_t err = cudaGetDevice(&d);
  if (err != cudaSuccess) return (int)err;
  if (d < 0 || d >= 64) return B200SLAM_EINVAL;
  if (!addr[d]) {
    void *p = nullptr;
    err = cudaGetSymbolAddress(&p, g_iter_proj_static_planes);
    if (err != cudaSuccess) return (int)err;
    if (!p) return B200SLAM_EINVAL;
    addr[d] = (float *)p;
  }
  *out = addr[d];
  return B200SLAM_OK;
}
}  // namespace

extern "C" int b200slam_iter_proj_static(const float *rays_img, const float *pts_3d_norm, const float *p_init,
                                         float *p_new, uint8_t *converged, int b, int h, int w, int n,
                                         int max_iter, float lambda_init, float cost_thresh, void *stream) {
  if (b < 0 || n < 0) return B200SLAM_EINVAL;
  if (b == 0 || n == 0) return B200SLAM_OK;
  if (!rays_img || !pts_3d_norm || !p_init || !p_new || !converged) return B200SLAM_EINVAL;
  if (h < 3 || w < 3 || (long long)h * w * 9 > 0x7fffffffLL) return B200SLAM_EINVAL;
  if ((unsigned long long)b * h * w * 9ull * sizeof(float) > (unsigned long long)B200SLAM_ITER_PROJ_STATIC_MAX_BYTES)
    return B200SLAM_ESTATIC_LIMIT;
  float *planes = nullptr;
  const int rc = iter_proj_static_planes(&planes);
  if (rc != B200SLAM_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const int hw = h * w;
  repack_planar_kernel<<<dim3((hw + IP_THREADS - 1) / IP_THREADS, b), IP_THREADS, 0, st>>>(rays_img, planes, hw);
  dim3 grid((n + IP_THREADS - 1) / IP_THREADS, b);
  iter_proj_kernel<<<grid, IP_THREADS, 0, st>>>(planes, pts_3d_norm, p_init, p_new, converged, h, w, n, max_iter,
                                                lambda_init, cost_thresh);
  return (int)cudaGetLastError();
}

// ---- fused match_iterative_proj --------------------------------------------
extern "C" int b200slam_internal_refine_f16(const void *D11, const void *D21, const int64_t *p1,
                                            int64_t *p1_new, int b, int h, int w, int n, int fdim,
                                            int radius, int dilation_max, int lin_out, int *used_tile,
                                            void *stream);

namespace {
struct MatchLayout {
  size_t planes, p1, p1b, d11h, d21h, total;
};
MatchLayout match_layout(int b, int h, int w, int fdim, int desc_f32) {
  const size_t hw = (size_t)h * w, al = 256;
  auto up = [&](size_t x) { return (x + al - 1) / al * al; };
  MatchLayout L;
  size_t off = 0;
  L.planes = off; off += up((size_t)b * hw * 9 * sizeof(float));
  L.p1 = off;     off += up((size_t)b * hw * 2 * sizeof(int64_t));
  L.p1b = off;    off += up((size_t)b * hw * 2 * sizeof(int64_t));
  L.d11h = off;   off += desc_f32 ? up((size_t)b * hw * fdim * sizeof(__half)) : 0;
  L.d21h = off;   off += desc_f32 ? up((size_t)b * hw * fdim * sizeof(__half)) : 0;
  L.total = off;
  return L;
}
}  // namespace

extern "C" size_t b200slam_match_workspace_bytes(int b, int h, int w, int fdim, int desc_f32) {
  if (b < 0 || h < 0 || w < 0 || fdim < 0) return 0;
  return match_layout(b, h, w, fdim, desc_f32).total;
}

extern "C" int b200slam_match(const float *X11, const float *X21, const void *D11, const void *D21,
                              int desc_f32, const int64_t *idx_init, int64_t *idx_1_to_2,
                              uint8_t *valid, int b, int h, int w, int fdim, int max_iter,
                              float lambda_init, float cost_thresh, float dist_thresh, int radius,
                              int dilation_max, void *workspace, size_t workspace_bytes, void *stream) {
  if (b < 0 || h < 0 || w < 0) return B200SLAM_EINVAL;
  if (b == 0 || h == 0 || w == 0) return B200SLAM_OK;
  if (!X11 || !X21 || !idx_1_to_2 || !valid) return B200SLAM_EINVAL;
  if (radius > 0 && (!D11 || !D21 || fdim < 1)) return B200SLAM_EINVAL;
  if (h < 3 || w < 3 || (long long)h * w * 9 > 0x7fffffffLL) return B200SLAM_EINVAL;
  const MatchLayout L = match_layout(b, h, w, fdim, radius > 0 ? desc_f32 : 0);
  if (!workspace || workspace_bytes < L.total) return B200SLAM_EWORKSPACE;
  cudaStream_t st = (cudaStream_t)stream;
  char *ws = (char *)workspace;
  float *planes = (float *)(ws + L.planes);
  int64_t *p1 = (int64_t *)(ws + L.p1), *p1b = (int64_t *)(ws + L.p1b);
  const int hw = h * w;
  const size_t total = (size_t)b * hw;

  prep_planes_kernel<<<dim3((w + PREP_TW - 1) / PREP_TW, (h + PREP_TH - 1) / PREP_TH, b), PREP_THREADS, 0, st>>>(
      X11, planes, h, w);
  match_proj_kernel<<<dim3((hw + IP_THREADS - 1) / IP_THREADS, b), IP_THREADS, 0, st>>>(
      planes, X11, X21, idx_init, p1, valid, h, w, max_iter, lambda_init, cost_thresh, dist_thresh);
  const int64_t *pairs = p1;
  if (radius > 0) {
    const void *a = D11, *q = D21;
    if (desc_f32) {  // matching.py:80-81: D11.half(), D21.half()
      const size_t nd = total * fdim;
      const unsigned blocks = (unsigned)((nd + 1023) / 1024);
      to_half_kernel<<<blocks, 256, 0, st>>>((const float *)D11, (__half *)(ws + L.d11h), nd);
      to_half_kernel<<<blocks, 256, 0, st>>>((const float *)D21, (__half *)(ws + L.d21h), nd);
      a = ws + L.d11h;
      q = ws + L.d21h;
    }
    int used_tile = -1;
    int rc = b200slam_internal_refine_f16(a, q, p1, idx_1_to_2, b, h, w, hw, fdim, radius, dilation_max, 1,
                                          &used_tile, stream);
    if (rc == B200SLAM_OK && used_tile == 1) return (int)cudaGetLastError();
    if (used_tile != 0) return rc;  // a real argument error
    rc = b200slam_internal_refine_f16(a, q, p1, p1b, b, h, w, hw, fdim, radius, dilation_max, 0, &used_tile, stream);
    if (rc != B200SLAM_OK) return rc;
    pairs = p1b;
  }
  pixel_to_lin_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(pairs, idx_1_to_2, total, w);
  return (int)cudaGetLastError();
}
