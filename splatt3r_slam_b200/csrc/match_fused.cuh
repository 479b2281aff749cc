// b200slam — fused front/back glue of match_iterative_proj (reference
// splatt3r_slam/matching.py:25-90 and image.py:5-38).  Included by matching.cu.
//
// The reference runs ~12 PyTorch kernels around the two backend calls: F.normalize (norm, clamp,
// div) of both pointmaps, two grouped 3x3 convolutions on a reflect-padded copy, cat + permute +
// contiguous to the 9-channel image, lin_to_pixel + float for the start pixels, .long(), an
// advanced-index gather + norm + compare for the occlusion test, and pixel_to_lin.  Here:
//   prep_planes_kernel   X11 -> the 9 PLANES iter_proj gathers from (the interleaved 9-channel image
//                        and its repack are never materialised): reads 12 B/px, writes 36 B/px
//   match_proj_kernel    iter_proj with X21 normalised in registers, the start pixel derived from
//                        idx_init (or the identity map), and the occlusion test + .long() folded
//                        into the epilogue: reads 12 (+8) B/px, writes 17 B/px
// The float operation ORDER follows what torch 2.11's CUDA kernels produce for that glue on B200
// (probed by tools/glue_probe.py, bit-identical on every pixel):
//   norm of a 3-vector      sqrt((x*x + z*z) + y*y), every product and sum rounded (the reduce
//                           kernel gives one lane per element and combines lanes 0+2, then +1)
//   normalize               x / max(norm, 1e-12)   (IEEE division)
//   grouped 3x3 conv        acc = 0; acc = fma(w[kh][kw], in, acc), kh outer, kw inner
// so the fused path returns the same matches as the unfused one, bit for bit (tests/test_gpu_match_fused.py).
#pragma once

constexpr int PREP_TW = 32, PREP_TH = 8, PREP_THREADS = PREP_TW * PREP_TH;

__device__ __forceinline__ float norm3_torch(float x, float y, float z) {
  return __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(z, z)), __fmul_rn(y, y)));
}

__device__ __forceinline__ int reflect_idx(int i, int n) {  // F.pad(mode="reflect") with pad 1
  if (i < 0) i = -i;
  if (i >= n) i = 2 * n - 2 - i;
  return min(max(i, 0), n - 1);  // only reached by threads beyond the image edge (results unused)
}

// X11 (b,h,w,3) -> planes (b, 9, h*w): ray(3), d/dx(3), d/dy(3).
__global__ void __launch_bounds__(PREP_THREADS)
prep_planes_kernel(const float *__restrict__ X11, float *__restrict__ planes, int h, int w) {
  __shared__ float ray[3][PREP_TH + 2][PREP_TW + 2];
  const int x0 = blockIdx.x * PREP_TW, y0 = blockIdx.y * PREP_TH;
  const size_t hw = (size_t)h * w;
  const float *src = X11 + (size_t)blockIdx.z * hw * 3;
  for (int t = threadIdx.x; t < (PREP_TH + 2) * (PREP_TW + 2); t += PREP_THREADS) {
    const int ly = t / (PREP_TW + 2), lx = t % (PREP_TW + 2);
    const int gy = reflect_idx(y0 + ly - 1, h), gx = reflect_idx(x0 + lx - 1, w);
    const float *p = src + ((size_t)gy * w + gx) * 3;
    const float x = __ldg(p), y = __ldg(p + 1), z = __ldg(p + 2);
    const float den = fmaxf(norm3_torch(x, y, z), 1e-12f);
    ray[0][ly][lx] = __fdiv_rn(x, den);
    ray[1][ly][lx] = __fdiv_rn(y, den);
    ray[2][ly][lx] = __fdiv_rn(z, den);
  }
  __syncthreads();
  const int lx = threadIdx.x % PREP_TW, ly = threadIdx.x / PREP_TW;
  const int x = x0 + lx, y = y0 + ly;
  if (x >= w || y >= h) return;
  float *dst = planes + (size_t)blockIdx.z * hw * 9 + (size_t)y * w + x;
  constexpr float k3 = 3.0f / 32.0f, k10 = 10.0f / 32.0f;
#pragma unroll
  for (int c = 0; c < 3; c++) {
    const float(*r)[PREP_TW + 2] = ray[c];
    const float a00 = r[ly][lx], a01 = r[ly][lx + 1], a02 = r[ly][lx + 2];
    const float a10 = r[ly + 1][lx], a11 = r[ly + 1][lx + 1], a12 = r[ly + 1][lx + 2];
    const float a20 = r[ly + 2][lx], a21 = r[ly + 2][lx + 1], a22 = r[ly + 2][lx + 2];
    float gx = __fmul_rn(-k3, a00);
    gx = __fmaf_rn(k3, a02, gx);
    gx = __fmaf_rn(-k10, a10, gx);
    gx = __fmaf_rn(k10, a12, gx);
    gx = __fmaf_rn(-k3, a20, gx);
    gx = __fmaf_rn(k3, a22, gx);
    float gy = __fmul_rn(-k3, a00);
    gy = __fmaf_rn(-k10, a01, gy);
    gy = __fmaf_rn(-k3, a02, gy);
    gy = __fmaf_rn(k3, a20, gy);
    gy = __fmaf_rn(k10, a21, gy);
    gy = __fmaf_rn(k3, a22, gy);
    dst[(size_t)c * hw] = a11;
    dst[(size_t)(3 + c) * hw] = gx;
    dst[(size_t)(6 + c) * hw] = gy;
  }
}

// One thread per query pixel of view 2 (n == h*w).  p1 (b,n,2) int64 = truncated pixel, valid (b,n).
__global__ void __launch_bounds__(IP_THREADS, IP_MIN_CTAS)
match_proj_kernel(const float *__restrict__ planes, const float *__restrict__ X11,
                  const float *__restrict__ X21, const int64_t *__restrict__ idx_init,
                  int64_t *__restrict__ p1, uint8_t *__restrict__ valid, int h, int w, int max_iter,
                  float lambda_init, float cost_thresh, float dist_thresh) {
  const int hw = h * w;
  const int q = blockIdx.x * IP_THREADS + threadIdx.x;
  if (q >= hw) return;
  const size_t bq = (size_t)blockIdx.y * hw + q;
  const float *img = planes + (size_t)blockIdx.y * hw * 9;

  // matching.py:20-24,44-49: p_init = (idx % w, idx // w).float() (python floor semantics)
  float pu, pv;
  if (idx_init) {
    const long long id = __ldg(idx_init + bq);
    long long qv = id / w, rv = id % w;
    if (rv < 0) { rv += w; qv -= 1; }
    pu = (float)rv;
    pv = (float)qv;
  } else {
    pu = (float)(q % w);
    pv = (float)(q / w);
  }
  const float umax = (float)(w - 2), vmax = (float)(h - 2);
  float u = clampf(pu, 1.0f, umax);
  float v = clampf(pv, 1.0f, vmax);
  const float x2 = __ldg(X21 + bq * 3), y2 = __ldg(X21 + bq * 3 + 1), z2 = __ldg(X21 + bq * 3 + 2);
  const float den = fmaxf(norm3_torch(x2, y2, z2), 1e-12f);
  const float p0 = __fdiv_rn(x2, den), p1n = __fdiv_rn(y2, den), p2 = __fdiv_rn(z2, den);
  const bool conv = iter_proj_core(img, hw, w, umax, vmax, u, v, p0, p1n, p2, max_iter, lambda_init, cost_thresh);

  // matching.py:66-76: p1.long(), then ||X11[p1] - X21|| < dist_thresh
  const int iu = (int)u, iv = (int)v;  // u, v in [1, w-2] x [1, h-2]: truncation == .long()
  const float *xs = X11 + ((size_t)blockIdx.y * hw + (size_t)iv * w + iu) * 3;
  const float dx = __fsub_rn(__ldg(xs), x2), dy = __fsub_rn(__ldg(xs + 1), y2), dz = __fsub_rn(__ldg(xs + 2), z2);
  const bool near = norm3_torch(dx, dy, dz) < dist_thresh;
  reinterpret_cast<longlong2 *>(p1)[bq] = make_longlong2(iu, iv);
  valid[bq] = (conv && near) ? 1 : 0;
}

__global__ void pixel_to_lin_kernel(const int64_t *__restrict__ p, int64_t *__restrict__ idx, size_t total, int w) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const longlong2 v = __ldg(reinterpret_cast<const longlong2 *>(p) + i);
  idx[i] = v.x + (long long)w * v.y;
}

__global__ void to_half_kernel(const float *__restrict__ src, __half *__restrict__ dst, size_t total) {
  const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i + 3 < total) {
    const float4 v = __ldg(reinterpret_cast<const float4 *>(src + i));
    reinterpret_cast<__half2 *>(dst + i)[0] = __floats2half2_rn(v.x, v.y);
    reinterpret_cast<__half2 *>(dst + i)[1] = __floats2half2_rn(v.z, v.w);
  } else {
    for (size_t k = i; k < total; k++) dst[k] = __float2half_rn(src[k]);
  }
}
