// b200slam — refine_matches for sm_100a (replaces matching_kernels.cu:24-115 of the reference).
//
// Per query: coarse-to-fine dilated (2r+1)^2 window scan around the current match, score =
// sequential fp16-rounded dot product of the query descriptor with the candidate descriptor,
// strict-> argmax carried across dilation levels (seed 0).  Bit-exact with the reference's
// c10::Half arithmetic: a*b = half(float(a)*float(b)) == HMUL (the product of two fp16 is
// exact in fp32, so one rounding), s+p = half(float(s)+float(p)) == HADD (the fp32 sum of two
// fp16 values is either exact or far from an fp16 tie), compare in fp32.
//
// Execution shape: 128-thread CTAs, query descriptor held in registers as packed half2,
// candidate rows fetched with 16-byte loads, products formed two at a time with HMUL2 and
// folded sequentially with HADD (1.5 issue slots per MAC instead of ~8 in the reference).
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <stdlib.h>
#include "../../include/b200slam.h"
#include "dev_cache.cuh"
#include "refine_tile.cuh"

#ifndef RT_QPQ
#define RT_QPQ 0  // A/B builds: the quarter-warp-per-query-pair kernel (experiments/refine_qpq.cuh, measured slower)
#endif
#if RT_QPQ
#include "../../experiments/refine_qpq.cuh"
#endif

#ifndef RT_BAND_FIXED
#define RT_BAND_FIXED 6
#endif
namespace {

constexpr int RF_THREADS = 128;

template <int FDIM>
__device__ __forceinline__ __half score_row(const __half2 (&q)[FDIM / 2], const uint4 *__restrict__ cand) {
  static_assert(FDIM % 8 == 0, "FDIM must be a multiple of 8");
  __half s = __ushort_as_half(0);
#pragma unroll
  for (int v = 0; v < FDIM / 8; v++) {
    const uint4 c = __ldg(cand + v);
    const uint32_t cw[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
    for (int t = 0; t < 4; t++) {
      const __half2 pr = __hmul2(q[v * 4 + t], *reinterpret_cast<const __half2 *>(&cw[t]));
      s = __hadd(s, __low2half(pr));
      s = __hadd(s, __high2half(pr));
    }
  }
  return s;
}

template <int FDIM>
__global__ void __launch_bounds__(RF_THREADS)
refine_f16_kernel(const __half *__restrict__ D11, const __half *__restrict__ D21,
                  const int64_t *__restrict__ p1, int64_t *__restrict__ p1_new, int h, int w, int n,
                  int radius, int dilation_max) {
  const int qi = blockIdx.x * RF_THREADS + threadIdx.x;
  if (qi >= n) return;
  const size_t bq = (size_t)blockIdx.y * n + qi;
  const __half *img = D11 + (size_t)blockIdx.y * h * w * FDIM;

  __half2 q[FDIM / 2];
  {
    const uint4 *src = reinterpret_cast<const uint4 *>(D21 + bq * FDIM);
#pragma unroll
    for (int v = 0; v < FDIM / 8; v++) {
      const uint4 c = __ldg(src + v);
      q[v * 4 + 0] = *reinterpret_cast<const __half2 *>(&c.x);
      q[v * 4 + 1] = *reinterpret_cast<const __half2 *>(&c.y);
      q[v * 4 + 2] = *reinterpret_cast<const __half2 *>(&c.z);
      q[v * 4 + 3] = *reinterpret_cast<const __half2 *>(&c.w);
    }
  }
  const longlong2 pin = __ldg(reinterpret_cast<const longlong2 *>(p1) + bq);
  long long u0 = pin.x, v0 = pin.y;
  long long u_new = u0, v_new = v0;
  float max_score = 0.0f;  // numeric_limits<c10::Half>::min() == 0 in the reference build

  for (int d = dilation_max; d > 0; d--) {
    const int rd = radius * d;
    for (int i = -rd; i <= rd; i += d) {
      const long long u = u0 + i;
      if (u < 0 || u >= w) continue;
      for (int j = -rd; j <= rd; j += d) {
        const long long v = v0 + j;
        if (v < 0 || v >= h) continue;
        const uint4 *cand = reinterpret_cast<const uint4 *>(img + ((size_t)v * w + (size_t)u) * FDIM);
        const float s = __half2float(score_row<FDIM>(q, cand));
        if (s > max_score) {
          max_score = s;
          u_new = u;
          v_new = v;
        }
      }
    }
    u0 = u_new;
    v0 = v_new;
  }
  reinterpret_cast<longlong2 *>(p1_new)[bq] = make_longlong2(u_new, v_new);
}

// Generic fdim (any value, 2-byte aligned rows): same arithmetic, scalar loads.
__global__ void __launch_bounds__(RF_THREADS)
refine_f16_generic_kernel(const __half *__restrict__ D11, const __half *__restrict__ D21,
                          const int64_t *__restrict__ p1, int64_t *__restrict__ p1_new, int h, int w,
                          int n, int fdim, int radius, int dilation_max) {
  const int qi = blockIdx.x * RF_THREADS + threadIdx.x;
  if (qi >= n) return;
  const size_t bq = (size_t)blockIdx.y * n + qi;
  const __half *img = D11 + (size_t)blockIdx.y * h * w * fdim;
  const __half *q = D21 + bq * fdim;
  long long u0 = p1[2 * bq], v0 = p1[2 * bq + 1];
  long long u_new = u0, v_new = v0;
  float max_score = 0.0f;
  for (int d = dilation_max; d > 0; d--) {
    const int rd = radius * d;
    for (int i = -rd; i <= rd; i += d) {
      const long long u = u0 + i;
      if (u < 0 || u >= w) continue;
      for (int j = -rd; j <= rd; j += d) {
        const long long v = v0 + j;
        if (v < 0 || v >= h) continue;
        const __half *cand = img + ((size_t)v * w + (size_t)u) * fdim;
        __half s = __ushort_as_half(0);
        for (int k = 0; k < fdim; k++) s = __hadd(s, __hmul(__ldg(q + k), __ldg(cand + k)));
        const float sf = __half2float(s);
        if (sf > max_score) { max_score = sf; u_new = u; v_new = v; }
      }
    }
    u0 = u_new;
    v0 = v_new;
  }
  p1_new[2 * bq] = u_new;
  p1_new[2 * bq + 1] = v_new;
}

// float / double instantiations: score = fma(q_k, c_k, score) (nvcc contracts the reference's MAC into one
// FFMA / DFMA), seed numeric_limits<scalar>::min() = FLT_MIN / DBL_MIN (matching_kernels.cu:46).
__device__ __forceinline__ float mac_rn(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ double mac_rn(double a, double b, double c) { return __fma_rn(a, b, c); }
template <typename real>
__global__ void __launch_bounds__(RF_THREADS)
refine_real_kernel(const real *__restrict__ D11, const real *__restrict__ D21,
                   const int64_t *__restrict__ p1, int64_t *__restrict__ p1_new, int h, int w, int n,
                   int fdim, int radius, int dilation_max, real seed) {
  const int qi = blockIdx.x * RF_THREADS + threadIdx.x;
  if (qi >= n) return;
  const size_t bq = (size_t)blockIdx.y * n + qi;
  const real *img = D11 + (size_t)blockIdx.y * h * w * fdim;
  const real *q = D21 + bq * fdim;
  long long u0 = p1[2 * bq], v0 = p1[2 * bq + 1];
  long long u_new = u0, v_new = v0;
  real max_score = seed;
  for (int d = dilation_max; d > 0; d--) {
    const int rd = radius * d;
    for (int i = -rd; i <= rd; i += d) {
      const long long u = u0 + i;
      if (u < 0 || u >= w) continue;
      for (int j = -rd; j <= rd; j += d) {
        const long long v = v0 + j;
        if (v < 0 || v >= h) continue;
        const real *cand = img + ((size_t)v * w + (size_t)u) * fdim;
        real s = 0;
        for (int k = 0; k < fdim; k++) s = mac_rn(__ldg(q + k), __ldg(cand + k), s);
        if (s > max_score) { max_score = s; u_new = u; v_new = v; }
      }
    }
    u0 = u_new;
    v0 = v_new;
  }
  p1_new[2 * bq] = u_new;
  p1_new[2 * bq + 1] = v_new;
}

}  // namespace

// lin_out != 0: p1_new is (b,n) linear indices u + w*v instead of (b,n,2) pixels (tiled kernel only;
// otherwise EINVAL and the caller converts).  *used_tile reports which path ran.
extern "C" int b200slam_internal_refine_f16(const void *D11, const void *D21, const int64_t *p1,
                                            int64_t *p1_new, int b, int h, int w, int n, int fdim,
                                            int radius, int dilation_max, int lin_out, int *used_tile,
                                            void *stream) {
  if (b < 0 || n < 0) return B200SLAM_EINVAL;
  if (b == 0 || n == 0) return B200SLAM_OK;
  if (!D11 || !D21 || !p1 || !p1_new) return B200SLAM_EINVAL;
  if (h < 1 || w < 1 || fdim < 0 || radius < 0) return B200SLAM_EINVAL;
  dim3 grid((n + RF_THREADS - 1) / RF_THREADS, b);
  cudaStream_t st = (cudaStream_t)stream;
  const __half *a = (const __half *)D11, *q = (const __half *)D21;
  const bool aligned = ((uintptr_t)D11 % 16 == 0) && ((uintptr_t)D21 % 16 == 0);
  // Tiled kernels with bulk-copy staged windows: queries form the (h, w) pixel grid of view 2
  // (n == h*w, as in matching.py:79-85), radius 3, and the nominal box of the coarsest level fits.
  const int reach = radius * dilation_max;
  if (fdim == 24 && aligned && radius == 3 && n == h * w && dilation_max >= 1 && w >= rtile::TW + 2 * reach + 2 &&
      (long long)h * w * 3 < 0x7fffffffLL) {
    // band count: fewest waves x rows-per-tile (time of a tile ~ its query rows); see refine_tile.cuh
    const int n_sm = b200dev::info().num_sms;
    const int tiles_x = (w + rtile::TW - 1) / rtile::TW, bands_min = (h + rtile::TH - 1) / rtile::TH;
    int bands = bands_min;
#ifndef RT_NO_BANDS
    long long best = -1;
    for (int nb = bands_min; nb <= bands_min + bands_min / 2 && nb <= h; nb++) {
      // time of a tile ~ its query rows + a fixed part (box staging, barriers, halo) worth RT_BAND_FIXED rows
      const long long waves = ((long long)tiles_x * nb * b + n_sm - 1) / n_sm, rows = (h + nb - 1) / nb + RT_BAND_FIXED;
      if (best < 0 || waves * rows <= best) { best = waves * rows; bands = nb; }  // ties: more, smaller tiles
    }
    static int forced = -1;  // tuning aid: B200SLAM_REFINE_BANDS=<n> overrides the choice
    if (forced < 0) { const char *e = getenv("B200SLAM_REFINE_BANDS"); forced = e ? atoi(e) : 0; }
    if (forced >= bands_min && forced <= h) bands = forced;
#endif
    const int tiles = tiles_x * bands;
    // a level whose full nominal box does not fit is scanned in two column passes of half the reach in x
    long long nominal = (long long)(rtile::TW + 2 * reach + rtile::SLACK) * (rtile::TH + 2 * reach + rtile::SLACK) * fdim * 2;
#if !RT_LINEAR
    if (nominal > rtile::BOX_BYTES) nominal = (long long)(rtile::TW + reach + 2) * (rtile::TH + 2 * reach + 2) * fdim * 2;
#endif
#if RT_QPQ
    if (h < 16000 && w < 16000 &&
        (long long)rtile::qpq_pitch16(rtile::TW + 2 * reach + rtile::SLACK, 3) * 16 * (rtile::TH + 2 * reach + rtile::SLACK) <=
            rtile::QPQ_BOX_BYTES) {
      constexpr int smem = rtile::QPQ_BOX_BYTES + rtile::THREADS / 2 * 24 * 4;
      static b200dev::Once attr_q;  // function attributes are per device
      if (attr_q.first())
        cudaFuncSetAttribute(rtile::refine_qpq_kernel<24, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
      rtile::refine_qpq_kernel<24, 3><<<dim3(tiles, b), rtile::THREADS, smem, st>>>(a, q, p1, p1_new, h, w, n, w, bands,
                                                                                    dilation_max, lin_out);
      if (used_tile) *used_tile = 1;
      return (int)cudaGetLastError();
    }
#endif
    if (nominal <= rtile::BOX_BYTES) {
      static b200dev::Once attr;  // function attributes are per device
      if (attr.first())
        cudaFuncSetAttribute(rtile::refine_tile_kernel<24, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             rtile::BOX_BYTES);
      rtile::refine_tile_kernel<24, 3><<<dim3(tiles, b), rtile::THREADS, rtile::BOX_BYTES, st>>>(
          a, q, p1, p1_new, h, w, n, w, bands, dilation_max, lin_out);
      if (used_tile) *used_tile = 1;
      return (int)cudaGetLastError();
    }
  }
  if (used_tile) *used_tile = 0;
  if (lin_out) return B200SLAM_EINVAL;  // caller converts (b200slam_match keeps a pair buffer for this case)
  if (fdim == 24 && aligned)
    refine_f16_kernel<24><<<grid, RF_THREADS, 0, st>>>(a, q, p1, p1_new, h, w, n, radius, dilation_max);
  else if (fdim == 16 && aligned)
    refine_f16_kernel<16><<<grid, RF_THREADS, 0, st>>>(a, q, p1, p1_new, h, w, n, radius, dilation_max);
  else if (fdim == 32 && aligned)
    refine_f16_kernel<32><<<grid, RF_THREADS, 0, st>>>(a, q, p1, p1_new, h, w, n, radius, dilation_max);
  else
    refine_f16_generic_kernel<<<grid, RF_THREADS, 0, st>>>(a, q, p1, p1_new, h, w, n, fdim, radius,
                                                           dilation_max);
  return (int)cudaGetLastError();
}

#if RT_STATS
// A/B builds only (-DRT_STATS=1): read and clear the per-level counters of the tiled kernel.
extern "C" int b200slam_internal_refine_stats(unsigned long long *out32) {
  cudaError_t err = cudaMemcpyFromSymbol(out32, rtile::g_rt_stats, sizeof(unsigned long long) * 32);
  if (err != cudaSuccess) return (int)err;
  static unsigned long long zeros[32] = {};
  return (int)cudaMemcpyToSymbol(rtile::g_rt_stats, zeros, sizeof(zeros));
}
#endif

extern "C" int b200slam_refine_matches_f16(const void *D11, const void *D21, const int64_t *p1,
                                           int64_t *p1_new, int b, int h, int w, int n, int fdim,
                                           int radius, int dilation_max, void *stream) {
  return b200slam_internal_refine_f16(D11, D21, p1, p1_new, b, h, w, n, fdim, radius, dilation_max, 0, nullptr, stream);
}

extern "C" int b200slam_refine_matches_f32(const float *D11, const float *D21, const int64_t *p1,
                                           int64_t *p1_new, int b, int h, int w, int n, int fdim,
                                           int radius, int dilation_max, void *stream) {
  if (b < 0 || n < 0) return B200SLAM_EINVAL;
  if (b == 0 || n == 0) return B200SLAM_OK;
  if (!D11 || !D21 || !p1 || !p1_new) return B200SLAM_EINVAL;
  if (h < 1 || w < 1 || fdim < 0 || radius < 0) return B200SLAM_EINVAL;
  dim3 grid((n + RF_THREADS - 1) / RF_THREADS, b);
  refine_real_kernel<float><<<grid, RF_THREADS, 0, (cudaStream_t)stream>>>(D11, D21, p1, p1_new, h, w, n, fdim, radius,
                                                                           dilation_max, 1.17549435e-38f);
  return (int)cudaGetLastError();
}

extern "C" int b200slam_refine_matches_f64(const double *D11, const double *D21, const int64_t *p1,
                                           int64_t *p1_new, int b, int h, int w, int n, int fdim,
                                           int radius, int dilation_max, void *stream) {
  if (b < 0 || n < 0) return B200SLAM_EINVAL;
  if (b == 0 || n == 0) return B200SLAM_OK;
  if (!D11 || !D21 || !p1 || !p1_new) return B200SLAM_EINVAL;
  if (h < 1 || w < 1 || fdim < 0 || radius < 0) return B200SLAM_EINVAL;
  dim3 grid((n + RF_THREADS - 1) / RF_THREADS, b);
  refine_real_kernel<double><<<grid, RF_THREADS, 0, (cudaStream_t)stream>>>(D11, D21, p1, p1_new, h, w, n, fdim, radius,
                                                                            dilation_max, 2.2250738585072014e-308);
  return (int)cudaGetLastError();
}
