// b200slam — iter_proj main kernel (included by matching.cu inside its anonymous namespace).
//
// Same arithmetic as the reference's loop (matching_kernels.cu:147-271), restructured around two
// identities that hold bit for bit:
//  (1) the 3-channel sample the reference takes at the proposed pixel (:224-249) is, when the
//      step is accepted, exactly the first three channels of the 9-channel sample it takes at the
//      top of the next iteration (:154-183) — same taps, same weights, same operation sequence —
//      and when the step is rejected the next iteration re-samples the unchanged pixel.  So ONE
//      9-channel sample per iteration, at the proposed pixel, replaces both, and the dependent
//      load rounds per iteration drop from two to one.
//  (2) while the proposal stays inside the last sampled integer pixel cell (the usual case after the
//      first few LM steps) the 36 raw taps are unchanged, so they are kept in registers and only
//      re-weighted: no loads at all.
#pragma once

// Raw bilinear taps of NCH channels at base pixel (bu, bv):
// t[j] = {(v+1,u+1), (v+1,u), (v,u+1), (v,u)} of channel plane j.
template <int NCH>
__device__ __forceinline__ void load_taps(float (&t)[9][4], const float *img, int hw, int w, int bu, int bv) {
  const float *top = img + bv * w + bu;
#pragma unroll
  for (int j = 0; j < NCH; j++) {
    const float *pj = top + (size_t)j * hw;
    t[j][0] = __ldg(pj + w + 1);
    t[j][1] = __ldg(pj + w);
    t[j][2] = __ldg(pj + 1);
    t[j][3] = __ldg(pj);
  }
}

// Residual of the normalised interpolated ray against the target ray, and its squared norm
// (matching_kernels.cu:185-197 and :240-254: the same operation sequence in both places).
struct Resid {
  float e0, e1, e2, cost;
};
__device__ __forceinline__ Resid residual(float c0, float c1, float c2, float p0, float p1, float p2) {
  Resid r;
  const float nrm = __fsqrt_rn(dot3(c0, c1, c2, c0, c1, c2));
  const float inv = __frcp_rn(nrm);
  r.e0 = __fmaf_rn(c0, inv, -p0);
  r.e1 = __fmaf_rn(c1, inv, -p1);
  r.e2 = __fmaf_rn(c2, inv, -p2);
  r.cost = dot3(r.e0, r.e1, r.e2, r.e0, r.e1, r.e2);
  return r;
}

// LM loop of one query: img = the 9 planes of its batch row, (u, v) = clamped start pixel (in/out),
// (p0, p1, p2) = target ray.  Returns the converged flag.
//
// Register budget: ONE set of 36 raw taps, the taps of the LAST SAMPLED cell (tu, tv), whether or not that
// proposal was accepted; the accepted state is only its 9 interpolated channels + residual.  (The first version
// kept the accepted cell's taps AND the proposal's: 72 tap registers, 119 in total, 4 CTAs per SM.)  A proposal
// in the sampled cell re-weights the registers, any other cell reloads — the same values either way, so the
// result does not depend on the policy.
__device__ __forceinline__ bool iter_proj_core(const float *__restrict__ img, int hw, int w, float umax, float vmax,
                                               float &u, float &v, float p0, float p1, float p2, int max_iter,
                                               float lambda_init, float cost_thresh) {
  float lambda = lambda_init;
  bool conv = false;
  if (max_iter > 0) {
    float T[9][4], c[9];
    int tu = (int)floorf(u), tv = (int)floorf(v);
    load_taps<9>(T, img, hw, w, tu, tv);
    {
      const Weights k = bilinear_weights(__fsub_rn(u, (float)tu), __fsub_rn(v, (float)tv));
#pragma unroll
      for (int j = 0; j < 9; j++) c[j] = tap(k, T[j][0], T[j][1], T[j][2], T[j][3]);
    }
    Resid r = residual(c[0], c[1], c[2], p0, p1, p2);
    for (int it = 0; it < max_iter; it++) {
      float A00 = dot3(c[3], c[4], c[5], c[3], c[4], c[5]);
      const float A01 = dot3(c[3], c[4], c[5], c[6], c[7], c[8]);
      float A11 = dot3(c[6], c[7], c[8], c[6], c[7], c[8]);
      const float s0 = dot3(r.e0, r.e1, r.e2, c[3], c[4], c[5]);
      const float s1 = dot3(r.e0, r.e1, r.e2, c[6], c[7], c[8]);
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
#ifndef IP_NO_RELOAD  // -DIP_NO_RELOAD: WRONG results, timing bound only — what a free tap reload (any staging scheme) could gain at most
#define IP_NO_RELOAD 0
#endif
      if (!IP_NO_RELOAD && (u2 != tu || v2 != tv)) {  // another cell: reload (usually an L1 hit: neighbouring lanes sample neighbouring cells)
        load_taps<9>(T, img, hw, w, u2, v2);
        tu = u2;
        tv = v2;
      }
      float cn[9];
#pragma unroll
      for (int j = 0; j < 9; j++) cn[j] = tap(k2, T[j][0], T[j][1], T[j][2], T[j][3]);
      const Resid rn = residual(cn[0], cn[1], cn[2], p0, p1, p2);
      if (rn.cost < r.cost) {
        u = un;
        v = vn;
#pragma unroll
        for (int j = 0; j < 9; j++) c[j] = cn[j];
        r = rn;
        lambda = __double2float_rn(__dmul_rn((double)lambda, 0.1));  // matching_kernels.cu:261
        conv = rn.cost < cost_thresh;
      } else {
        lambda = __fmul_rn(lambda, 10.0f);
        conv = r.cost < cost_thresh;
      }
    }
  }
  return conv;
}

// One thread per query pixel; planes (9, h*w) of batch row blockIdx.y.
__global__ void __launch_bounds__(IP_THREADS, IP_MIN_CTAS)
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
  const bool conv = iter_proj_core(img, hw, w, umax, vmax, u, v, p0, p1, p2, max_iter, lambda_init, cost_thresh);
  reinterpret_cast<float2 *>(p_new)[bq] = make_float2(u, v);
  converged[bq] = conv ? 1 : 0;
}
