/* ORACLE — TEST INFRASTRUCTURE ONLY.  Never linked or called by the product path.
 *
 * Plain-C CPU restatement of the reference's Sim3 Gauss-Newton backend
 *   /root/reference/splatt3r_slam/backend/src/gn_kernels.cu
 * Per-point arithmetic is fp32 in the reference's expression order; the per-edge sums,
 * which the reference forms in fp32 (256 strided partials + tree, :909,:1095-1136), are
 * accumulated in fp64 here and rounded to fp32 once — the oracle is therefore at least
 * as accurate as the reference and agrees with it to ~1e-6 relative; the contract for
 * this part of the path is 1e-5 relative (BASELINE.json north_star), not bit-exactness.
 *
 * Parity status: pinned against outputs of the reference CUDA build (oracle/_ref run on a
 * B200; fixtures under tests/golden/).  The reference has no golden vectors of its own.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define GN_EPS 1e-6 /* gn_kernels.cu:33 */

/* ---- Sim3 helpers, gn_kernels.cu:171-296 ---------------------------------------- */
static inline float huber(float r) { /* :171-174 */
  const float a = fabsf(r);
  return a < 1.345 ? 1.0f : (float)(1.345 / a);
}
static void quat_comp(const float *qi, const float *qj, float *o) { /* :177-183 */
  o[0] = qi[3] * qj[0] + qi[0] * qj[3] + qi[1] * qj[2] - qi[2] * qj[1];
  o[1] = qi[3] * qj[1] - qi[0] * qj[2] + qi[1] * qj[3] + qi[2] * qj[0];
  o[2] = qi[3] * qj[2] + qi[0] * qj[1] - qi[1] * qj[0] + qi[2] * qj[3];
  o[3] = qi[3] * qj[3] - qi[0] * qj[0] - qi[1] * qj[1] - qi[2] * qj[2];
}
static void actSO3(const float *q, const float *X, float *Y) { /* :194-204 */
  float uv[3];
  uv[0] = (float)(2.0 * (q[1] * X[2] - q[2] * X[1]));
  uv[1] = (float)(2.0 * (q[2] * X[0] - q[0] * X[2]));
  uv[2] = (float)(2.0 * (q[0] * X[1] - q[1] * X[0]));
  float y0 = X[0] + q[3] * uv[0] + (q[1] * uv[2] - q[2] * uv[1]);
  float y1 = X[1] + q[3] * uv[1] + (q[2] * uv[0] - q[0] * uv[2]);
  float y2 = X[2] + q[3] * uv[2] + (q[0] * uv[1] - q[1] * uv[0]);
  Y[0] = y0; Y[1] = y1; Y[2] = y2;
}
static void actSim3(const float *t, const float *q, const float *s, const float *X, float *Y) {
  actSO3(q, X, Y); /* :206-218 */
  Y[0] *= s[0]; Y[1] *= s[0]; Y[2] *= s[0];
  Y[0] += t[0]; Y[1] += t[1]; Y[2] += t[2];
}
static void relSim3(const float *ti, const float *qi, const float *si, const float *tj,
                    const float *qj, const float *sj, float *tij, float *qij, float *sij) {
  /* :251-271  T_ij = T_i^-1 T_j */
  float si_inv = (float)(1.0 / si[0]);
  sij[0] = si_inv * sj[0];
  float qi_inv[4] = {-qi[0], -qi[1], -qi[2], qi[3]};
  quat_comp(qi_inv, qj, qij);
  float d[3] = {tj[0] - ti[0], tj[1] - ti[1], tj[2] - ti[2]};
  actSO3(qi_inv, d, tij);
  tij[0] *= si_inv; tij[1] *= si_inv; tij[2] *= si_inv;
}
static void apply_Sim3_adj_inv(const float *t, const float *q, const float *s, const float *X,
                               float *Y) { /* :276-296 */
  const float s_inv = (float)(1.0 / s[0]);
  float Ra[3];
  actSO3(q, &X[0], Ra);
  Y[0] = s_inv * Ra[0]; Y[1] = s_inv * Ra[1]; Y[2] = s_inv * Ra[2];
  actSO3(q, &X[3], &Y[3]);
  Y[3] += s_inv * (t[1] * Ra[2] - t[2] * Ra[1]);
  Y[4] += s_inv * (t[2] * Ra[0] - t[0] * Ra[2]);
  Y[5] += s_inv * (t[0] * Ra[1] - t[1] * Ra[0]);
  Y[6] = X[6] + (s_inv * (t[0] * Ra[0] + t[1] * Ra[1] + t[2] * Ra[2]));
}

/* accumulate one residual row: hij (105, lower-tri of 14x14), vi, vj.  :1001-1012 */
static inline void accum_row(double *hij, double *vi, double *vj, const float *ti,
                             const float *qi, const float *si, const float *J0, float w,
                             float err) {
  float Jx[14];
  apply_Sim3_adj_inv(ti, qi, si, J0, &Jx[7]);
  for (int n = 0; n < 7; n++) Jx[n] = -Jx[7 + n];
  int l = 0;
  for (int n = 0; n < 14; n++)
    for (int m = 0; m <= n; m++) hij[l++] += (double)(w * Jx[n] * Jx[m]);
  for (int n = 0; n < 7; n++) {
    vi[n] += (double)(w * err * Jx[n]);
    vj[n] += (double)(w * err * Jx[7 + n]);
  }
}

typedef struct {
  int kind; /* 0 rays, 1 calib, 2 points */
  float sigma_a, sigma_b; /* rays: ray,dist; calib: pixel,depth; points: point,- */
  float C_thresh, Q_thresh;
  float fx, fy, cx, cy;
  int height, width, pixel_border;
  float z_eps;
} gn_params_t;

/* One directed edge: gn_kernels.cu ray_align :812-1137, calib_proj :1230-1542,
 * point_align :454-722.  ix/jx are ROWS into Twc/Xs/Cs. */
static void edge_build(const gn_params_t *P, const float *Twc, const float *Xs, const float *Cs,
                       int ix, int jx, const int64_t *idx, const uint8_t *valid, const float *Q,
                       int n, double *hij, double *vi, double *vj) {
  const float *ti = Twc + 8 * ix, *qi = ti + 3, *si = ti + 7;
  const float *tj = Twc + 8 * jx, *qj = tj + 3, *sj = tj + 7;
  float tij[3], qij[4], sij[1];
  relSim3(ti, qi, si, tj, qj, sj, tij, qij, sij);
  memset(hij, 0, 105 * sizeof(double));
  memset(vi, 0, 7 * sizeof(double));
  memset(vj, 0, 7 * sizeof(double));
  const float sa_inv = (float)(1.0 / P->sigma_a);
  const float sb_inv = (float)(1.0 / P->sigma_b);
  const float *Xi_base = Xs + (size_t)ix * n * 3, *Xj_base = Xs + (size_t)jx * n * 3;
  const float *Ci_base = Cs + (size_t)ix * n, *Cj_base = Cs + (size_t)jx * n;
  float J0[7];
  for (int k = 0; k < n; k++) {
    const int vm = valid[k] != 0;
    const int64_t ind = vm ? idx[k] : 0; /* :912-913 */
    const float *Xi = Xi_base + 3 * ind, *Xj = Xj_base + 3 * (size_t)k;
    float Pj[3];
    actSim3(tij, qij, sij, Xj, Pj);
    const float q = Q[k], ci = Ci_base[ind], cj = Cj_base[k];
    int ok = vm & (q > P->Q_thresh) & (ci > P->C_thresh) & (cj > P->C_thresh);
    if (P->kind == 0) {
      const float n2i = Xi[0] * Xi[0] + Xi[1] * Xi[1] + Xi[2] * Xi[2];
      const float n1i = sqrtf(n2i), n1i_inv = (float)(1.0 / n1i);
      float ri[3] = {n1i_inv * Xi[0], n1i_inv * Xi[1], n1i_inv * Xi[2]};
      const float n2j = Pj[0] * Pj[0] + Pj[1] * Pj[1] + Pj[2] * Pj[2];
      const float n1j = sqrtf(n2j), n1j_inv = (float)(1.0 / n1j);
      float rj[3] = {n1j_inv * Pj[0], n1j_inv * Pj[1], n1j_inv * Pj[2]};
      float err[4] = {rj[0] - ri[0], rj[1] - ri[1], rj[2] - ri[2], n1j - n1i};
      const float sq = sqrtf(q);
      const float swr = ok ? sa_inv * sq : 0, swd = ok ? sb_inv * sq : 0; /* :962-963 */
      float w[4] = {huber(swr * err[0]), huber(swr * err[1]), huber(swr * err[2]),
                    huber(swd * err[3])};
      const float wcr = swr * swr, wcd = swd * swd;
      w[0] *= wcr; w[1] *= wcr; w[2] *= wcr; w[3] *= wcd;
      if (!ok) continue; /* all four weights are exactly 0: contributes nothing */
      const float n3 = n1j_inv / n2j; /* :981-987 */
      const float dxx = n1j_inv - Pj[0] * Pj[0] * n3, dyy = n1j_inv - Pj[1] * Pj[1] * n3,
                  dzz = n1j_inv - Pj[2] * Pj[2] * n3;
      const float dxy = -Pj[0] * Pj[1] * n3, dxz = -Pj[0] * Pj[2] * n3, dyz = -Pj[1] * Pj[2] * n3;
      J0[0] = dxx; J0[1] = dxy; J0[2] = dxz; J0[3] = 0; J0[4] = rj[2]; J0[5] = -rj[1]; J0[6] = 0;
      accum_row(hij, vi, vj, ti, qi, si, J0, w[0], err[0]);
      J0[0] = dxy; J0[1] = dyy; J0[2] = dyz; J0[3] = -rj[2]; J0[4] = 0; J0[5] = rj[0]; J0[6] = 0;
      accum_row(hij, vi, vj, ti, qi, si, J0, w[1], err[1]);
      J0[0] = dxz; J0[1] = dyz; J0[2] = dzz; J0[3] = rj[1]; J0[4] = -rj[0]; J0[5] = 0; J0[6] = 0;
      accum_row(hij, vi, vj, ti, qi, si, J0, w[2], err[2]);
      J0[0] = rj[0]; J0[1] = rj[1]; J0[2] = rj[2]; J0[3] = 0; J0[4] = 0; J0[5] = 0; J0[6] = n1j;
      accum_row(hij, vi, vj, ti, qi, si, J0, w[3], err[3]);
    } else if (P->kind == 1) {
      const int u_t = (int)(ind % P->width), v_t = (int)(ind / P->width); /* :1360-1361 */
      const int valid_z = (Pj[2] > P->z_eps) && (Xi[2] > P->z_eps);
      const float zj_inv = valid_z ? (float)(1.0 / Pj[2]) : 0.0f;
      const float zj_log = valid_z ? logf(Pj[2]) : 0.0f;
      const float zi_log = valid_z ? logf(Xi[2]) : 0.0f;
      const float xz = Pj[0] * zj_inv, yz = Pj[1] * zj_inv;
      const float u = P->fx * xz + P->cx, v = P->fy * yz + P->cy;
      const int valid_u = (u > (float)P->pixel_border) && (u < (float)(P->width - 1 - P->pixel_border));
      const int valid_v = (v > (float)P->pixel_border) && (v < (float)(P->height - 1 - P->pixel_border));
      float err[3] = {u - (float)u_t, v - (float)v_t, zj_log - zi_log};
      ok = ok & valid_u & valid_v & valid_z;
      if (!ok) continue;
      const float sq = sqrtf(q);
      const float swp = sa_inv * sq, swd = sb_inv * sq;
      float w[3] = {huber(swp * err[0]), huber(swp * err[1]), huber(swd * err[2])};
      const float wcp = swp * swp, wcd = swd * swd;
      w[0] *= wcp; w[1] *= wcp; w[2] *= wcd;
      const float fx = P->fx, fy = P->fy; /* :1421-1478 */
      J0[0] = fx * zj_inv; J0[1] = 0; J0[2] = -fx * xz * zj_inv; J0[3] = -fx * xz * yz;
      J0[4] = fx * (1 + xz * xz); J0[5] = -fx * yz; J0[6] = 0;
      accum_row(hij, vi, vj, ti, qi, si, J0, w[0], err[0]);
      J0[0] = 0; J0[1] = fy * zj_inv; J0[2] = -fy * yz * zj_inv; J0[3] = -fy * (1 + yz * yz);
      J0[4] = fy * xz * yz; J0[5] = fy * xz; J0[6] = 0;
      accum_row(hij, vi, vj, ti, qi, si, J0, w[1], err[1]);
      J0[0] = 0; J0[1] = 0; J0[2] = zj_inv; J0[3] = yz; J0[4] = -xz; J0[5] = 0; J0[6] = 1.0f;
      accum_row(hij, vi, vj, ti, qi, si, J0, w[2], err[2]);
    } else {
      if (!ok) continue;
      float err[3] = {Pj[0] - Xi[0], Pj[1] - Xi[1], Pj[2] - Xi[2]}; /* :567-569 */
      const float sw = sa_inv * sqrtf(q);
      float w[3] = {huber(sw * err[0]), huber(sw * err[1]), huber(sw * err[2])};
      const float wc = sw * sw;
      w[0] *= wc; w[1] *= wc; w[2] *= wc;
      J0[0] = 1; J0[1] = 0; J0[2] = 0; J0[3] = 0; J0[4] = Pj[2]; J0[5] = -Pj[1]; J0[6] = Pj[0];
      accum_row(hij, vi, vj, ti, qi, si, J0, w[0], err[0]); /* :601-607 */
      J0[0] = 0; J0[1] = 1; J0[2] = 0; J0[3] = -Pj[2]; J0[4] = 0; J0[5] = Pj[0]; J0[6] = Pj[1];
      accum_row(hij, vi, vj, ti, qi, si, J0, w[1], err[1]);
      J0[0] = 0; J0[1] = 0; J0[2] = 1; J0[3] = Pj[1]; J0[4] = -Pj[0]; J0[5] = 0; J0[6] = Pj[2];
      accum_row(hij, vi, vj, ti, qi, si, J0, w[2], err[2]);
    }
  }
}

/* Build Hs (4,E,7,7) f32 and gs (2,E,7) f32 for all edges (scatter :1119-1132). */
void oracle_gn_build(int kind, const float *Twc, const float *Xs, const float *Cs, const float *K,
                     const int64_t *ii_row, const int64_t *jj_row, const int64_t *idx,
                     const uint8_t *valid, const float *Q, int E, int n, float sigma_a,
                     float sigma_b, float C_thresh, float Q_thresh, int height, int width,
                     int pixel_border, float z_eps, float *Hs, float *gs) {
  gn_params_t P;
  memset(&P, 0, sizeof(P));
  P.kind = kind; P.sigma_a = sigma_a; P.sigma_b = sigma_b;
  P.C_thresh = C_thresh; P.Q_thresh = Q_thresh;
  if (K) { P.fx = K[0]; P.fy = K[4]; P.cx = K[2]; P.cy = K[5]; } /* :1273-1278 */
  P.height = height; P.width = width; P.pixel_border = pixel_border; P.z_eps = z_eps;
#pragma omp parallel for schedule(dynamic, 1)
  for (int e = 0; e < E; e++) {
    double hij[105], vi[7], vj[7];
    edge_build(&P, Twc, Xs, Cs, (int)ii_row[e], (int)jj_row[e], idx + (size_t)e * n,
               valid + (size_t)e * n, Q + (size_t)e * n, n, hij, vi, vj);
    for (int k = 0; k < 7; k++) {
      gs[((size_t)0 * E + e) * 7 + k] = (float)vi[k];
      gs[((size_t)1 * E + e) * 7 + k] = (float)vj[k];
    }
    int l = 0;
    for (int a = 0; a < 14; a++)
      for (int m = 0; m <= a; m++, l++) {
        const float v = (float)hij[l];
        if (a < 7) {
          Hs[(((size_t)0 * E + e) * 7 + a) * 7 + m] = v;
          Hs[(((size_t)0 * E + e) * 7 + m) * 7 + a] = v;
        } else if (m < 7) {
          Hs[(((size_t)1 * E + e) * 7 + m) * 7 + (a - 7)] = v;
          Hs[(((size_t)2 * E + e) * 7 + (a - 7)) * 7 + m] = v;
        } else {
          Hs[(((size_t)3 * E + e) * 7 + (a - 7)) * 7 + (m - 7)] = v;
          Hs[(((size_t)3 * E + e) * 7 + (m - 7)) * 7 + (a - 7)] = v;
        }
      }
  }
}
