/* ORACLE — TEST INFRASTRUCTURE ONLY.  Never linked or called by the product path.
 *
 * Plain-C CPU restatement of the reference's matching kernels
 *   /root/reference/splatt3r_slam/backend/src/matching_kernels.cu
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library.
 *
 * Parity status: pinned against outputs of the reference CUDA build itself
 * (oracle/_ref, run on a B200; fixtures in tests/golden/, generator
 * tests/golden/make_golden.py).  The reference ships no golden vectors of its own.
 *
 * Rounding contract (bit-exact target).  The reference is compiled with nvcc's default
 * -fmad=true, so which multiply/add pairs are fused is decided by NVVM *and* ptxas.  The
 * sequence below was read off the sm_100a SASS of the reference kernel (nvcc 12.9):
 * every fmaf() here is an FFMA there, every plain * / + / - a lone FMUL / FADD, the
 * bilinear weights go through double (matching_kernels.cu:161-163), `1.0/x` compiles to
 * an IEEE float reciprocal (rcp.rn), sqrtf is IEEE (sqrt.rn).
 * Compile with -ffp-contract=off so the C compiler adds no fusions of its own.
 */
#include <math.h>
#include <stdint.h>
#include <stddef.h>

static inline float clampf(float x, float lo, float hi) {
  /* matching_kernels.cu:20-22  fmin(fmax(x,min),max) */
  return fminf(fmaxf(x, lo), hi);
}

/* One bilinear tap set.  matching_kernels.cu:154-169 (and :224-238).
 * Corner naming is crossed on purpose in the reference: w11 (=du*dv) multiplies the
 * (v+1,u+1) pixel, w12 the (v+1,u) pixel, w21 the (v,u+1) pixel, w22 the (v,u) pixel. */
typedef struct {
  float w11, w12, w21, w22;
  const float *r11, *r12, *r21, *r22;
} taps_t;

static inline taps_t make_taps(const float *img, int w, int c, float u, float v) {
  taps_t t;
  int u11 = (int)floorf(u);
  int v11 = (int)floorf(v);
  float du = u - (float)u11;
  float dv = v - (float)v11;
  double omdu = 1.0 - (double)du;
  double omdv = 1.0 - (double)dv;
  t.w11 = du * dv;
  t.w12 = (float)(omdu * (double)dv);
  t.w21 = (float)((double)du * omdv);
  t.w22 = (float)(omdu * omdv);
  t.r11 = img + ((size_t)(v11 + 1) * w + (u11 + 1)) * c;
  t.r12 = img + ((size_t)(v11 + 1) * w + u11) * c;
  t.r21 = img + ((size_t)v11 * w + (u11 + 1)) * c;
  t.r22 = img + ((size_t)v11 * w + u11) * c;
  return t;
}

static inline float tap(const taps_t *t, int j) {
  /* matching_kernels.cu:173  w11*r11 + w12*r12 + w21*r21 + w22*r22
   * SASS: FMUL(w12,r12) ; FFMA(w11,r11,.) ; FFMA(w21,r21,.) ; FFMA(w22,r22,.) */
  float s = t->w12 * t->r12[j];
  s = fmaf(t->w11, t->r11[j], s);
  s = fmaf(t->w21, t->r21[j], s);
  s = fmaf(t->w22, t->r22[j], s);
  return s;
}

static inline float dot3_ref(float a0, float a1, float a2, float b0, float b1, float b2) {
  /* a0*b0 + a1*b1 + a2*b2 -> SASS: FMUL(a1,b1) ; FFMA(a0,b0,.) ; FFMA(a2,b2,.) */
  float s = a1 * b1;
  s = fmaf(a0, b0, s);
  s = fmaf(a2, b2, s);
  return s;
}

/* iter_proj for one batch row.  matching_kernels.cu:118-274.
 * rays_img (h,w,9) f32, pts (n,3) f32, p_init (n,2) f32 -> p_new (n,2) f32, converged (n) u8 */
static void iter_proj_row(const float *img, const float *pts, const float *p_init,
                          float *p_new, uint8_t *conv, int h, int w, int n, int max_iter,
                          float lambda_init, float cost_thresh) {
  const int c = 9;
  const float umax = (float)(w - 2), vmax = (float)(h - 2);
#pragma omp parallel for schedule(static)
  for (int q = 0; q < n; q++) {
    float u = clampf(p_init[2 * q + 0], 1.0f, umax); /* :142-143 */
    float v = clampf(p_init[2 * q + 1], 1.0f, vmax);
    const float p0 = pts[3 * q + 0], p1 = pts[3 * q + 1], p2 = pts[3 * q + 2];
    float lambda = lambda_init;
    uint8_t cv = 0; /* outputs are torch::zeros (:299) */
    for (int it = 0; it < max_iter; it++) {
      taps_t t = make_taps(img, w, c, u, v);
      float r0 = tap(&t, 0), r1 = tap(&t, 1), r2 = tap(&t, 2);
      float gx0 = tap(&t, 3), gx1 = tap(&t, 4), gx2 = tap(&t, 5);
      float gy0 = tap(&t, 6), gy1 = tap(&t, 7), gy2 = tap(&t, 8);
      /* :185-190 normalise; :194-196 err = r*inv - p fused by ptxas into one FFMA */
      float nrm = sqrtf(dot3_ref(r0, r1, r2, r0, r1, r2));
      float inv = 1.0f / nrm;
      float e0 = fmaf(r0, inv, -p0), e1 = fmaf(r1, inv, -p1), e2 = fmaf(r2, inv, -p2);
      float cost = dot3_ref(e0, e1, e2, e0, e1, e2); /* :197 */
      /* :201-209 */
      float A00 = dot3_ref(gx0, gx1, gx2, gx0, gx1, gx2);
      float A01 = dot3_ref(gx0, gx1, gx2, gy0, gy1, gy2);
      float A11 = dot3_ref(gy0, gy1, gy2, gy0, gy1, gy2);
      float s0 = dot3_ref(e0, e1, e2, gx0, gx1, gx2); /* = -b0 */
      float s1 = dot3_ref(e0, e1, e2, gy0, gy1, gy2); /* = -b1 */
      A00 = A00 + lambda;
      A11 = A11 + lambda;
      /* :212-214.  det = FFMA(A00,A11,-(A01*A01)); numerators FFMA(A01,s1,-(A11*s0)) etc. */
      float det = fmaf(A00, A11, -(A01 * A01));
      float dinv = 1.0f / det;
      float nu = fmaf(A01, s1, -(s0 * A11));
      float nv = fmaf(A01, s0, -(s1 * A00));
      float un = clampf(fmaf(nu, dinv, u), 1.0f, umax); /* :217-220 */
      float vn = clampf(fmaf(nv, dinv, v), 1.0f, vmax);
      /* :224-255 re-sample 3 channels at the candidate pixel */
      taps_t t2 = make_taps(img, w, c, un, vn);
      float q0 = tap(&t2, 0), q1 = tap(&t2, 1), q2 = tap(&t2, 2);
      float nrm2 = sqrtf(dot3_ref(q0, q1, q2, q0, q1, q2));
      float inv2 = 1.0f / nrm2;
      float f0 = fmaf(q0, inv2, -p0), f1 = fmaf(q1, inv2, -p1), f2 = fmaf(q2, inv2, -p2);
      float new_cost = dot3_ref(f0, f1, f2, f0, f1, f2);
      /* :258-267.  (new_cost < cost) is false for NaN -> else branch */
      if (new_cost < cost) {
        u = un;
        v = vn;
        lambda = (float)((double)lambda * 0.1);
        cv = new_cost < cost_thresh;
      } else {
        lambda = lambda * 10.0f;
        cv = cost < cost_thresh;
      }
    }
    p_new[2 * q + 0] = u; /* :271-272 */
    p_new[2 * q + 1] = v;
    conv[q] = cv;
  }
}

void oracle_iter_proj(const float *rays_img, const float *pts, const float *p_init,
                      float *p_new, uint8_t *conv, int b, int h, int w, int n,
                      int max_iter, float lambda_init, float cost_thresh) {
  for (int i = 0; i < b; i++)
    iter_proj_row(rays_img + (size_t)i * h * w * 9, pts + (size_t)i * n * 3,
                  p_init + (size_t)i * n * 2, p_new + (size_t)i * n * 2, conv + (size_t)i * n,
                  h, w, n, max_iter, lambda_init, cost_thresh);
}

/* refine_matches, c10::Half instantiation.  matching_kernels.cu:24-80.
 * c10::Half arithmetic (torch/headeronly/util/Half.h): a*b = half(float(a)*float(b)),
 * s += p = half(float(s)+float(p)); compare in float; seed
 * ::cuda::std::numeric_limits<c10::Half>::min() == 0 (PTX: mov.u16 0).
 * D11 (b,h,w,fdim) f16, D21 (b,n,fdim) f16, p1 (b,n,2) i64 -> p1_new (b,n,2) i64 */
void oracle_refine_matches_f16(const uint16_t *D11_, const uint16_t *D21_, const int64_t *p1,
                               int64_t *p1_new, int b, int h, int w, int n, int fdim,
                               int radius, int dilation_max) {
  const _Float16 *D11 = (const _Float16 *)D11_;
  const _Float16 *D21 = (const _Float16 *)D21_;
#pragma omp parallel for schedule(static)
  for (long bq = 0; bq < (long)b * n; bq++) {
    const int bi = (int)(bq / n);
    const _Float16 *q = D21 + (size_t)bq * fdim;
    const _Float16 *img = D11 + (size_t)bi * h * w * fdim;
    int64_t u0 = p1[2 * bq + 0], v0 = p1[2 * bq + 1];
    _Float16 max_score = (_Float16)0.0f;
    int64_t u_new = u0, v_new = v0;
    for (int d = dilation_max; d > 0; d--) {
      const int rd = radius * d;
      const int diam = 2 * rd + 1;
      for (int i = 0; i < diam; i += d) {
        for (int j = 0; j < diam; j += d) {
          const int64_t u = u0 - rd + i;
          const int64_t v = v0 - rd + j;
          if (v >= 0 && v < h && u >= 0 && u < w) {
            const _Float16 *cnd = img + ((size_t)v * w + (size_t)u) * fdim;
            _Float16 score = (_Float16)0.0f;
            for (int k = 0; k < fdim; k++) {
              _Float16 prod = (_Float16)((float)q[k] * (float)cnd[k]);
              score = (_Float16)((float)score + (float)prod);
            }
            if ((float)score > (float)max_score) {
              max_score = score;
              u_new = u;
              v_new = v;
            }
          }
        }
      }
      u0 = u_new;
      v0 = v_new;
    }
    p1_new[2 * bq + 0] = u_new;
    p1_new[2 * bq + 1] = v_new;
  }
}

/* float instantiation (seed FLT_MIN = numeric_limits<float>::min(); plain float MACs;
 * nvcc contracts score += a*b into an FFMA). */
void oracle_refine_matches_f32(const float *D11, const float *D21, const int64_t *p1,
                               int64_t *p1_new, int b, int h, int w, int n, int fdim,
                               int radius, int dilation_max) {
#pragma omp parallel for schedule(static)
  for (long bq = 0; bq < (long)b * n; bq++) {
    const int bi = (int)(bq / n);
    const float *q = D21 + (size_t)bq * fdim;
    const float *img = D11 + (size_t)bi * h * w * fdim;
    int64_t u0 = p1[2 * bq + 0], v0 = p1[2 * bq + 1];
    float max_score = 1.17549435e-38f;
    int64_t u_new = u0, v_new = v0;
    for (int d = dilation_max; d > 0; d--) {
      const int rd = radius * d;
      const int diam = 2 * rd + 1;
      for (int i = 0; i < diam; i += d)
        for (int j = 0; j < diam; j += d) {
          const int64_t u = u0 - rd + i, v = v0 - rd + j;
          if (v >= 0 && v < h && u >= 0 && u < w) {
            const float *cnd = img + ((size_t)v * w + (size_t)u) * fdim;
            float score = 0.0f;
            for (int k = 0; k < fdim; k++) score = fmaf(q[k], cnd[k], score);
            if (score > max_score) { max_score = score; u_new = u; v_new = v; }
          }
        }
      u0 = u_new;
      v0 = v_new;
    }
    p1_new[2 * bq + 0] = u_new;
    p1_new[2 * bq + 1] = v_new;
  }
}

/* double instantiation of refine_matches_kernel (AT_DISPATCH_FLOATING_TYPES_AND_HALF, matching_kernels.cu:102):
 * same scan, seed ::cuda::std::numeric_limits<double>::min() = DBL_MIN (:46), `score += a * b` contracted to one
 * DFMA by nvcc exactly like the float instantiation's FFMA. */
void oracle_refine_matches_f64(const double *D11, const double *D21, const int64_t *p1,
                               int64_t *p1_new, int b, int h, int w, int n, int fdim,
                               int radius, int dilation_max) {
#pragma omp parallel for schedule(static)
  for (long bq = 0; bq < (long)b * n; bq++) {
    const int bi = (int)(bq / n);
    const double *q = D21 + (size_t)bq * fdim;
    const double *img = D11 + (size_t)bi * h * w * fdim;
    int64_t u0 = p1[2 * bq + 0], v0 = p1[2 * bq + 1];
    double max_score = 2.2250738585072014e-308;
    int64_t u_new = u0, v_new = v0;
    for (int d = dilation_max; d > 0; d--) {
      const int rd = radius * d;
      const int diam = 2 * rd + 1;
      for (int i = 0; i < diam; i += d)
        for (int j = 0; j < diam; j += d) {
          const int64_t u = u0 - rd + i, v = v0 - rd + j;
          if (v >= 0 && v < h && u >= 0 && u < w) {
            const double *cnd = img + ((size_t)v * w + (size_t)u) * fdim;
            double score = 0.0;
            for (int k = 0; k < fdim; k++) score = fma(q[k], cnd[k], score);
            if (score > max_score) { max_score = score; u_new = u; v_new = v; }
          }
        }
      u0 = u_new;
      v0 = v_new;
    }
    p1_new[2 * bq + 0] = u_new;
    p1_new[2 * bq + 1] = v_new;
  }
}
