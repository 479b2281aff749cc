/* ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_gn.c header).
 *
 * Assembly + solve + retraction + GN loop of the reference backend:
 *   SparseBlock   gn_kernels.cu:56-158   (assembly, SimplicialLLT<double>, zero on failure)
 *   expSO3/expSim3/retrSim3/pose_retr_kernel   :298-452
 *   gauss_newton_*_cuda host loops   :724-810, :1139-1227, :1545-1637
 * The sparse LLT is restated as a dense fp64 Cholesky of the same matrix (same lower
 * triangle, same "not SPD => dx = 0" rule); results agree to fp64 rounding.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

void oracle_gn_build(int kind, const float *Twc, const float *Xs, const float *Cs, const float *K,
                     const int64_t *ii_row, const int64_t *jj_row, const int64_t *idx,
                     const uint8_t *valid, const float *Q, int E, int n, float sigma_a,
                     float sigma_b, float C_thresh, float Q_thresh, int height, int width,
                     int pixel_border, float z_eps, float *Hs, float *gs);

/* Assemble + solve.  ii_opt/jj_opt = row - num_fix (negative => fixed, skipped, :85,:106).
 * dx (M,7) f32 = -(A^-1 b), zeros when the factorisation fails (:141-149, :1208). */
int oracle_gn_solve(const float *Hs, const float *gs, const int64_t *ii_opt,
                    const int64_t *jj_opt, int E, int M, float *dx) {
  const int D = 7 * M;
  double *A = (double *)calloc((size_t)D * D, sizeof(double));
  double *b = (double *)calloc((size_t)D, sizeof(double));
  for (int blk = 0; blk < 4; blk++)
    for (int e = 0; e < E; e++) {
      const int64_t r = (blk < 2) ? ii_opt[e] : jj_opt[e];      /* cat(ii,ii,jj,jj) :1201 */
      const int64_t c = (blk % 2 == 0) ? ii_opt[e] : jj_opt[e]; /* cat(ii,jj,ii,jj) :1202 */
      if (r < 0 || c < 0) continue;
      const float *H = Hs + ((size_t)blk * E + e) * 49;
      for (int k = 0; k < 7; k++)
        for (int l = 0; l < 7; l++) A[(size_t)(7 * r + k) * D + (7 * c + l)] += (double)H[k * 7 + l];
    }
  for (int blk = 0; blk < 2; blk++)
    for (int e = 0; e < E; e++) {
      const int64_t r = blk == 0 ? ii_opt[e] : jj_opt[e];
      if (r < 0) continue;
      for (int k = 0; k < 7; k++) b[7 * r + k] += (double)gs[((size_t)blk * E + e) * 7 + k];
    }
  /* dense LLT on the lower triangle */
  int ok = 1;
  for (int j = 0; j < D && ok; j++) {
    double d = A[(size_t)j * D + j];
    for (int k = 0; k < j; k++) d -= A[(size_t)j * D + k] * A[(size_t)j * D + k];
    if (!(d > 0.0)) { ok = 0; break; }
    d = sqrt(d);
    A[(size_t)j * D + j] = d;
    for (int i = j + 1; i < D; i++) {
      double s = A[(size_t)i * D + j];
      for (int k = 0; k < j; k++) s -= A[(size_t)i * D + k] * A[(size_t)j * D + k];
      A[(size_t)i * D + j] = s / d;
    }
  }
  if (ok) {
    for (int i = 0; i < D; i++) {
      double s = b[i];
      for (int k = 0; k < i; k++) s -= A[(size_t)i * D + k] * b[k];
      b[i] = s / A[(size_t)i * D + i];
    }
    for (int i = D - 1; i >= 0; i--) {
      double s = b[i];
      for (int k = i + 1; k < D; k++) s -= A[(size_t)k * D + i] * b[k];
      b[i] = s / A[(size_t)i * D + i];
    }
    for (int i = 0; i < D; i++) dx[i] = -(float)b[i];
  } else {
    for (int i = 0; i < D; i++) dx[i] = -0.0f;
  }
  free(A);
  free(b);
  return ok;
}

static void quat_comp(const float *qi, const float *qj, float *o) {
  o[0] = qi[3] * qj[0] + qi[0] * qj[3] + qi[1] * qj[2] - qi[2] * qj[1];
  o[1] = qi[3] * qj[1] - qi[0] * qj[2] + qi[1] * qj[3] + qi[2] * qj[0];
  o[2] = qi[3] * qj[2] + qi[0] * qj[1] - qi[1] * qj[0] + qi[2] * qj[3];
  o[3] = qi[3] * qj[3] - qi[0] * qj[0] - qi[1] * qj[1] - qi[2] * qj[2];
}
static void actSO3(const float *q, const float *X, float *Y) {
  float uv[3];
  uv[0] = (float)(2.0 * (q[1] * X[2] - q[2] * X[1]));
  uv[1] = (float)(2.0 * (q[2] * X[0] - q[0] * X[2]));
  uv[2] = (float)(2.0 * (q[0] * X[1] - q[1] * X[0]));
  float y0 = X[0] + q[3] * uv[0] + (q[1] * uv[2] - q[2] * uv[1]);
  float y1 = X[1] + q[3] * uv[1] + (q[2] * uv[0] - q[0] * uv[2]);
  float y2 = X[2] + q[3] * uv[2] + (q[0] * uv[1] - q[1] * uv[0]);
  Y[0] = y0; Y[1] = y1; Y[2] = y2;
}
static void cross_inplace(const float *a, float *b) { /* :228-239 */
  float x[3] = {a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]};
  b[0] = x[0]; b[1] = x[1]; b[2] = x[2];
}
static void expSO3(const float *phi, float *q) { /* :298-320 */
  float theta_sq = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  float imag, real;
  if (theta_sq < 1e-6) {
    float theta_p4 = theta_sq * theta_sq;
    imag = (float)(0.5 - (1.0 / 48.0) * theta_sq + (1.0 / 3840.0) * theta_p4);
    real = (float)(1.0 - (1.0 / 8.0) * theta_sq + (1.0 / 384.0) * theta_p4);
  } else {
    float theta = sqrtf(theta_sq);
    imag = sinf((float)(0.5 * theta)) / theta;
    real = cosf((float)(0.5 * theta));
  }
  q[0] = imag * phi[0]; q[1] = imag * phi[1]; q[2] = imag * phi[2]; q[3] = real;
}
static void expSim3(const float *xi, float *t, float *q, float *s) { /* :322-389 */
  float tau[3] = {xi[0], xi[1], xi[2]};
  float phi[3] = {xi[3], xi[4], xi[5]};
  float sigma = xi[6];
  float scale = expf(sigma);
  expSO3(phi, q);
  s[0] = scale;
  float theta_sq = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  float theta = sqrtf(theta_sq);
  float A, B, C;
  const float one = 1.0f, half = 0.5f;
  if (fabsf(sigma) < 1e-6) {
    C = one;
    if (fabsf(theta) < 1e-6) {
      A = half;
      B = (float)(1.0 / 6.0);
    } else {
      A = (one - cosf(theta)) / theta_sq;
      B = (theta - sinf(theta)) / (theta_sq * theta);
    }
  } else {
    C = (scale - one) / sigma;
    if (fabsf(theta) < 1e-6) {
      float sigma_sq = sigma * sigma;
      A = ((sigma - one) * scale + one) / sigma_sq;
      B = (scale * half * sigma_sq + scale - one - sigma * scale) / (sigma_sq * sigma);
    } else {
      float a = scale * sinf(theta);
      float b = scale * cosf(theta);
      float c = theta_sq + sigma * sigma;
      A = (a * sigma + (one - b) * theta) / (theta * c);
      B = (C - ((b - one) * sigma + a * theta) / (c)) / (theta_sq);
    }
  }
  t[0] = C * tau[0]; t[1] = C * tau[1]; t[2] = C * tau[2];
  cross_inplace(phi, tau);
  t[0] += A * tau[0]; t[1] += A * tau[1]; t[2] += A * tau[2];
  cross_inplace(phi, tau);
  t[0] += B * tau[0]; t[1] += B * tau[1]; t[2] += B * tau[2];
}

/* pose_retr_kernel :414-452 with retrSim3 :391-412.  poses (N,8) in place, dx (N-num_fix,7) */
void oracle_pose_retr(float *poses, const float *dx, int N, int num_fix) {
  for (int k = num_fix; k < N; k++) {
    float *p = poses + 8 * k;
    float dt[3], dq[4], ds[1], t1[3], q1[4];
    expSim3(dx + 7 * (k - num_fix), dt, dq, ds);
    quat_comp(dq, p + 3, q1);
    actSO3(dq, p, t1);
    t1[0] *= ds[0]; t1[1] *= ds[0]; t1[2] *= ds[0];
    t1[0] += dt[0]; t1[1] += dt[1]; t1[2] += dt[2];
    p[0] = t1[0]; p[1] = t1[1]; p[2] = t1[2];
    p[3] = q1[0]; p[4] = q1[1]; p[5] = q1[2]; p[6] = q1[3];
    p[7] = ds[0] * p[7];
  }
}

static int cmp_i64(const void *a, const void *b) {
  int64_t x = *(const int64_t *)a, y = *(const int64_t *)b;
  return (x > y) - (x < y);
}

/* Full driver (:1139-1227 and siblings).  ii/jj are KEYFRAME IDS; rows come from
 * unique+searchsorted (:160-169); num_fix = 1 (:1156).  Twc updated in place.
 * Returns the number of iterations executed; dx_out (N-1,7) = last dx. */
int oracle_gauss_newton(int kind, float *Twc, const float *Xs, const float *Cs, const float *K,
                        const int64_t *ii, const int64_t *jj, const int64_t *idx,
                        const uint8_t *valid, const float *Q, int N, int E, int n,
                        float sigma_a, float sigma_b, float C_thresh, float Q_thresh, int height,
                        int width, int pixel_border, float z_eps, int max_iter,
                        float delta_thresh, float *dx_out) {
  const int num_fix = 1;
  int64_t *uniq = (int64_t *)malloc(sizeof(int64_t) * 2 * (size_t)E);
  memcpy(uniq, ii, sizeof(int64_t) * E);
  memcpy(uniq + E, jj, sizeof(int64_t) * E);
  qsort(uniq, 2 * (size_t)E, sizeof(int64_t), cmp_i64);
  int U = 0;
  for (int i = 0; i < 2 * E; i++)
    if (i == 0 || uniq[i] != uniq[i - 1]) uniq[U++] = uniq[i];
  int64_t *rows = (int64_t *)malloc(sizeof(int64_t) * 4 * (size_t)E);
  int64_t *ii_row = rows, *jj_row = rows + E, *ii_opt = rows + 2 * E, *jj_opt = rows + 3 * E;
  for (int e = 0; e < E; e++) {
    int a = 0, b = 0;
    while (uniq[a] < ii[e]) a++;
    while (uniq[b] < jj[e]) b++;
    ii_row[e] = a; jj_row[e] = b; ii_opt[e] = a - num_fix; jj_opt[e] = b - num_fix;
  }
  float *Hs = (float *)malloc(sizeof(float) * 4 * (size_t)E * 49);
  float *gs = (float *)malloc(sizeof(float) * 2 * (size_t)E * 7);
  int it = 0;
  for (; it < max_iter;) {
    oracle_gn_build(kind, Twc, Xs, Cs, K, ii_row, jj_row, idx, valid, Q, E, n, sigma_a, sigma_b,
                    C_thresh, Q_thresh, height, width, pixel_border, z_eps, Hs, gs);
    oracle_gn_solve(Hs, gs, ii_opt, jj_opt, E, N - num_fix, dx_out);
    oracle_pose_retr(Twc, dx_out, N, num_fix);
    it++;
    double nn = 0;
    for (int i = 0; i < 7 * (N - num_fix); i++) nn += (double)dx_out[i] * dx_out[i];
    if ((float)sqrt(nn) < delta_thresh) break; /* :1218-1221 */
  }
  free(uniq); free(rows); free(Hs); free(gs);
  return it;
}
