// b200slam — stage 3 of a GN iteration: deterministic assembly of the normal equations,
// fp64 Cholesky, back-substitution, Sim3 retraction and the convergence flag, all on the device.
// Replaces SparseBlock (gn_kernels.cu:56-158: host-side Eigen SimplicialLLT<double>),
// pose_retr_kernel (:414-452) and the host-side ||dx|| test (:1218-1221).
#pragma once

namespace b200gn {

constexpr int NUM_FIX = 1;  // gn_kernels.cu:1156

// M rows = apply_Sim3_adj_inv(T_i) applied to unit vectors (gn_kernels.cu:276-296), in fp64,
// row-major M[k*7+c]:  Jj = J0 * M.
__device__ inline void adj_inv_matrix(const float *Ti, double *M) {
  const double t[3] = {Ti[0], Ti[1], Ti[2]}, q[4] = {Ti[3], Ti[4], Ti[5], Ti[6]};
  const double s_inv = 1.0 / (double)Ti[7];
  double R[3][3];
  for (int c = 0; c < 3; c++) {
    double X[3] = {0, 0, 0};
    X[c] = 1.0;
    const double uv0 = 2.0 * (q[1] * X[2] - q[2] * X[1]);
    const double uv1 = 2.0 * (q[2] * X[0] - q[0] * X[2]);
    const double uv2 = 2.0 * (q[0] * X[1] - q[1] * X[0]);
    R[0][c] = X[0] + q[3] * uv0 + (q[1] * uv2 - q[2] * uv1);
    R[1][c] = X[1] + q[3] * uv1 + (q[2] * uv0 - q[0] * uv2);
    R[2][c] = X[2] + q[3] * uv2 + (q[0] * uv1 - q[1] * uv0);
  }
  for (int k = 0; k < 49; k++) M[k] = 0.0;
  for (int k = 0; k < 3; k++) {
    const double Ra[3] = {R[0][k], R[1][k], R[2][k]};
    M[k * 7 + 0] = s_inv * Ra[0]; M[k * 7 + 1] = s_inv * Ra[1]; M[k * 7 + 2] = s_inv * Ra[2];
    M[k * 7 + 3] = s_inv * (t[1] * Ra[2] - t[2] * Ra[1]);
    M[k * 7 + 4] = s_inv * (t[2] * Ra[0] - t[0] * Ra[2]);
    M[k * 7 + 5] = s_inv * (t[0] * Ra[1] - t[1] * Ra[0]);
    M[k * 7 + 6] = s_inv * (t[0] * Ra[0] + t[1] * Ra[1] + t[2] * Ra[2]);
    M[(3 + k) * 7 + 3] = R[0][k]; M[(3 + k) * 7 + 4] = R[1][k]; M[(3 + k) * 7 + 5] = R[2][k];
  }
  M[48] = 1.0;
}

constexpr int FIN_SCRATCH = 36 + 49 + 49;  // doubles of shared scratch per warp

// Warp-cooperative stage 2 for one edge: sum the S per-CTA partials in fixed order (fp64),
// then H_jj = M^T H0 M and g_j = M^T g0 -> out[0..48], out[49..55].
__device__ inline void finalize_edge(const float *Twc, const int64_t *ii_row, const double *partials, int S,
                                     int e, int slot, double *out, double *sc, int lane) {
  double *acc = sc, *M = sc + 36, *T = sc + 36 + 49;
  // lane L owns splits L, L+32, ... (all 35 sums independent -> loads pipeline), then a fixed
  // xor-shuffle tree combines the lanes: deterministic, no serial chain over S.
  {
    double s[NACC];
#pragma unroll
    for (int l = 0; l < NACC; l++) s[l] = 0.0;
    for (int k = lane; k < S; k += 32) {
      const double *p = partials + ((size_t)slot * S + k) * NACC;
#pragma unroll
      for (int l = 0; l < NACC; l++) s[l] += __ldcg(p + l);
    }
#pragma unroll
    for (int l = 0; l < NACC; l++) {
      double v = s[l];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) acc[l] = v;
    }
  }
  if (lane == 0) adj_inv_matrix(Twc + 8 * (int)ii_row[e], M);
  __syncwarp();
  for (int o = lane; o < 49; o += 32) {
    const int a = o / 7, c = o % 7;
    double s = 0.0;
    for (int k = 0; k < 7; k++) {
      const int lo = a < k ? a : k, hi = a < k ? k : a;
      s += acc[lo * 7 - lo * (lo - 1) / 2 + (hi - lo)] * M[k * 7 + c];
    }
    T[o] = s;
  }
  __syncwarp();
  for (int o = lane; o < 56; o += 32) {
    double s = 0.0;
    if (o < 49) {
      const int r = o / 7, c = o % 7;
      for (int k = 0; k < 7; k++) s += M[k * 7 + r] * T[k * 7 + c];
    } else {
      const int r = o - 49;
      for (int k = 0; k < 7; k++) s += M[k * 7 + r] * acc[28 + k];
    }
    out[o] = s;
  }
  __syncwarp();
}

// One CTA per free pose r: zero its 7 x D strip of A, then add every incident edge block in
// edge order (fixed order => deterministic, no atomics).  Mirrors update_lhs/update_rhs
// (:70-112): blocks with a fixed pose (index < 0) are skipped, duplicates are summed.
__global__ void __launch_bounds__(256)
assemble_kernel(const int64_t *__restrict__ ii_row, const int64_t *__restrict__ jj_row,
                const double *__restrict__ edge_blocks, int E, int D, double *__restrict__ A,
                double *__restrict__ rhs, const State *__restrict__ state, int gated) {
  if (state->done || (gated && !state->pcg_bad)) return;
  const int r = blockIdx.x;
  double *strip = A + (size_t)7 * r * D;
  for (int t = threadIdx.x; t < 7 * D; t += blockDim.x) strip[t] = 0.0;
  __syncthreads();
  // The CTA scans the edge list 256 edges at a time (ballot per warp); threads 0..55 then visit only the
  // incident edges of the chunk, in edge order (a serial scan of E = 8192 edges per pose cost 2.8 ms).
  __shared__ unsigned incident[8];
  const int t = threadIdx.x, k = t / 7, l = t % 7;  // t in [49,56): rhs component l
  double accb = 0.0;
  for (int base = 0; base < E; base += 256) {
    const int e = base + t;
    bool hit = false;
    if (e < E) hit = ((int)ii_row[e] - NUM_FIX == r) || ((int)jj_row[e] - NUM_FIX == r);
    const unsigned m = __ballot_sync(0xffffffffu, hit);
    if ((t & 31) == 0) incident[t >> 5] = m;
    __syncthreads();
    if (t < 56) {
      for (int w = 0; w < 8; w++) {
        unsigned mm = incident[w];
        while (mm) {
          const int e2 = base + 32 * w + __ffs(mm) - 1;
          mm &= mm - 1;
          const int i = (int)ii_row[e2] - NUM_FIX, j = (int)jj_row[e2] - NUM_FIX;
          const double v = edge_blocks[(size_t)e2 * 56 + t];
          if (t < 49) {
            if (i == r) {
              strip[(size_t)k * D + 7 * i + l] += v;
              if (j >= 0) strip[(size_t)k * D + 7 * j + l] -= v;
            }
            if (j == r) {
              strip[(size_t)k * D + 7 * j + l] += v;
              if (i >= 0) strip[(size_t)k * D + 7 * i + l] -= v;
            }
          } else {
            if (i == r) accb -= v;  // g_i = -g_j
            if (j == r) accb += v;
          }
        }
      }
    }
    __syncthreads();
  }
  if (t >= 49 && t < 56) rhs[7 * r + l] = accb;
}

// ---- Sim3 exponential / retraction (expSO3 :298-320, expSim3 :322-389, retrSim3 :391-412) -----
__device__ inline void retract_pose(float *p, const float *xi) {
  const float tau[3] = {xi[0], xi[1], xi[2]}, phi[3] = {xi[3], xi[4], xi[5]}, sigma = xi[6];
  const float scale = expf(sigma);
  const float theta_sq = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  const float theta = sqrtf(theta_sq);
  float imag, real;
  if (theta_sq < 1e-6f) {
    const float p4 = theta_sq * theta_sq;
    imag = 0.5f - (1.0f / 48.0f) * theta_sq + (1.0f / 3840.0f) * p4;
    real = 1.0f - (1.0f / 8.0f) * theta_sq + (1.0f / 384.0f) * p4;
  } else {
    imag = sinf(0.5f * theta) / theta;
    real = cosf(0.5f * theta);
  }
  const float dq[4] = {imag * phi[0], imag * phi[1], imag * phi[2], real};
  float A, B, C;
  if (fabsf(sigma) < 1e-6f) {
    C = 1.0f;
    if (fabsf(theta) < 1e-6f) {
      A = 0.5f;
      B = 1.0f / 6.0f;
    } else {
      A = (1.0f - cosf(theta)) / theta_sq;
      B = (theta - sinf(theta)) / (theta_sq * theta);
    }
  } else {
    C = (scale - 1.0f) / sigma;
    if (fabsf(theta) < 1e-6f) {
      const float s2 = sigma * sigma;
      A = ((sigma - 1.0f) * scale + 1.0f) / s2;
      B = (scale * 0.5f * s2 + scale - 1.0f - sigma * scale) / (s2 * sigma);
    } else {
      const float a = scale * sinf(theta), b = scale * cosf(theta), c = theta_sq + sigma * sigma;
      A = (a * sigma + (1.0f - b) * theta) / (theta * c);
      B = (C - ((b - 1.0f) * sigma + a * theta) / c) / theta_sq;
    }
  }
  // dt = (C I + A [phi]x + B [phi]x^2) tau
  const float c1[3] = {phi[1] * tau[2] - phi[2] * tau[1], phi[2] * tau[0] - phi[0] * tau[2],
                       phi[0] * tau[1] - phi[1] * tau[0]};
  const float c2[3] = {phi[1] * c1[2] - phi[2] * c1[1], phi[2] * c1[0] - phi[0] * c1[2],
                       phi[0] * c1[1] - phi[1] * c1[0]};
  const float dt[3] = {C * tau[0] + A * c1[0] + B * c2[0], C * tau[1] + A * c1[1] + B * c2[1],
                       C * tau[2] + A * c1[2] + B * c2[2]};
  // T <- Exp(xi) * T   (left composition)
  const float q[4] = {p[3], p[4], p[5], p[6]};
  float t1[3];
  rotate(dq, p, t1);
  p[0] = t1[0] * scale + dt[0];
  p[1] = t1[1] * scale + dt[1];
  p[2] = t1[2] * scale + dt[2];
  p[3] = dq[3] * q[0] + dq[0] * q[3] + dq[1] * q[2] - dq[2] * q[1];
  p[4] = dq[3] * q[1] - dq[0] * q[2] + dq[1] * q[3] + dq[2] * q[0];
  p[5] = dq[3] * q[2] + dq[0] * q[1] - dq[1] * q[0] + dq[2] * q[3];
  p[6] = dq[3] * q[3] - dq[0] * q[0] - dq[1] * q[1] - dq[2] * q[2];
  p[7] = scale * p[7];
}

// Lower-triangular storage accessors: packed in shared memory (small systems) or the full
// row-major matrix in global memory.
struct PackedLower {
  double *p;
  __device__ __forceinline__ double &at(int i, int k) const { return p[(size_t)i * (i + 1) / 2 + k]; }
};
struct FullLower {
  double *p;
  int D;
  __device__ __forceinline__ double &at(int i, int k) const { return p[(size_t)i * D + k]; }
};

// Right-looking Cholesky + two triangular solves by ONE CTA.  Returns false when a pivot is
// not positive (the reference's "solver.info() != Success" => dx = 0, :141-149).
// x (length D, shared) holds the rhs on entry and the solution on exit; col is scratch (D).
// WARP = true: executed by the 32 threads of warp 0 only (D <= 32), __syncwarp instead of barriers.
//
// Two barriers per column: every thread reads the pivot itself and forms 1/sqrt(d) (no
// broadcast step), L(j,j) stores that RECIPROCAL (the triangular solves multiply by it), and
// the forward substitution L y = b rides along as one more row of the trailing update.  The
// backward substitution is done by warp 0 alone with warp-level synchronisation (D <= 256).
constexpr int CHOL_UNROLL = 7;  // register-blocked trailing update for D <= 224
template <class Mat, bool WARP = false>
__device__ bool cholesky_solve(Mat L, int D, double *x, double *col, int *fail,
                               unsigned long long *stamp = nullptr) {
#define CHOL_SYNC() do { if (WARP) __syncwarp(); else __syncthreads(); } while (0)
  const int tid = threadIdx.x, T = WARP ? 32 : blockDim.x;
  const int lane = tid & 31, warp = tid >> 5, nwarp = T >> 5;
  if (tid == 0) *fail = 0;
  CHOL_SYNC();
  for (int j = 0; j < D; j++) {
    const double d = L.at(j, j);  // final: the previous column's trailing update is behind a barrier
    if (!(d > 0.0)) {             // uniform: every thread sees the same pivot
      if (tid == 0) *fail = 1;
      return false;
    }
    const double dinv = rsqrt(d);
    const double yj = x[j] * dinv;  // forward substitution, column-oriented
    for (int i = j + 1 + tid; i < D; i += T) {
      const double v = L.at(i, j) * dinv;
      L.at(i, j) = v;
      col[i] = v;
      x[i] -= v * yj;
    }
    CHOL_SYNC();  // col[] complete; everyone has read L(j,j) and x[j]
    if (tid == 0) {
      L.at(j, j) = dinv;
      x[j] = yj;
    }
    // trailing update: row i (one warp per row), k over lanes
    if (D <= 32 * CHOL_UNROLL) {
      // a lane owns columns k = j+1+lane+32m: its col[k] stay in registers for the whole column
      // step, and a row's loads are all issued before its FMAs/stores (the plain loop below
      // serialises on possible aliasing of L and col: measured 1.6 us per column at D = 217)
      double ck[CHOL_UNROLL];
#pragma unroll
      for (int m = 0; m < CHOL_UNROLL; m++) {
        const int k = j + 1 + lane + 32 * m;
        ck[m] = k < D ? col[k] : 0.0;
      }
      for (int i = j + 1 + warp; i < D; i += nwarp) {
        const double ci = col[i];
        double *row = &L.at(i, 0);
        double v[CHOL_UNROLL];
#pragma unroll
        for (int m = 0; m < CHOL_UNROLL; m++) {
          const int k = j + 1 + lane + 32 * m;
          if (k <= i) v[m] = row[k];
        }
#pragma unroll
        for (int m = 0; m < CHOL_UNROLL; m++) {
          const int k = j + 1 + lane + 32 * m;
          if (k <= i) row[k] = v[m] - ci * ck[m];
        }
      }
    } else {
      for (int i = j + 1 + warp; i < D; i += nwarp) {
        const double ci = col[i];
        for (int k = j + 1 + lane; k <= i; k += 32) L.at(i, k) -= ci * col[k];
      }
    }
    CHOL_SYNC();
  }
  if (stamp && tid == 0) *stamp = gtimer();
  // backward: L^T x = y
  if (WARP || D <= 256) {
    if (warp == 0) {
      for (int j = D - 1; j >= 0; j--) {
        const double xj = x[j] * L.at(j, j);
        __syncwarp();
        if (lane == 0) x[j] = xj;
        for (int i = lane; i < j; i += 32) x[i] -= L.at(j, i) * xj;
        __syncwarp();
      }
    }
    CHOL_SYNC();
  } else {
    for (int j = D - 1; j >= 0; j--) {
      const double xj = x[j] * L.at(j, j);
      CHOL_SYNC();
      if (tid == 0) x[j] = xj;
      for (int i = tid; i < j; i += T) x[i] -= L.at(j, i) * xj;
      CHOL_SYNC();
    }
  }
  return true;
#undef CHOL_SYNC
}

// dx = -x (f32; zeros on failure), ||dx||_2 taken on the f32 values like the reference's
// linalg_norm (:1218), retraction of the free poses, iteration bookkeeping.  One CTA.
__device__ inline void finish_update(bool ok, const double *x, double *scratch, int D, int N, float *Twc,
                                     float *dx, float delta_thresh, State *state) {
  const int tid = threadIdx.x;
  double part = 0.0;
  for (int i = tid; i < D; i += blockDim.x) {
    const float v = ok ? -(float)x[i] : 0.0f;
    dx[i] = v;
    part += (double)v * (double)v;
  }
  for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
  __syncthreads();
  if ((tid & 31) == 0) scratch[tid >> 5] = part;  // scratch: >= 32 doubles
  __syncthreads();
  for (int k = NUM_FIX + tid; k < N; k += blockDim.x) retract_pose(Twc + 8 * k, dx + 7 * (k - NUM_FIX));
  if (tid == 0) {
    double s = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); w++) s += scratch[w];
    state->iters += 1;
    if (!ok) state->chol_fail += 1;
    if ((float)sqrt(s) < delta_thresh) state->done = 1;
  }
}

}  // namespace b200gn
#include "gn_chol_blocked.cuh"
namespace b200gn {

constexpr int SOLVE_THREADS = 512;
constexpr int PACKED_MAX_D = B200SLAM_SINGLE_CTA_MAX_D;  // 217*218/2*8 = 189 KB of packed lower triangle in smem

// dynamic smem: [x: Dp][col: Dp][packed L: D(D+1)/2][Pt: CB*ldp][Dg: CB*CB + CB] if PACKED, else [x][col]
__host__ __device__ inline int chol_ldp(int D) { return ((D + 3) & ~3) + 4; }
inline size_t solve_smem_bytes(int D, int Dp, bool packed) {
  size_t n = (size_t)2 * Dp;
  if (packed) n += (size_t)D * (D + 1) / 2 + (size_t)CB * chol_ldp(D) + CB * CB + CB;
  return n * sizeof(double);
}

template <bool PACKED>
__global__ void __launch_bounds__(SOLVE_THREADS)
solve_retract_kernel(float *__restrict__ Twc, double *__restrict__ A, const double *__restrict__ rhs,
                     int N, int D, int Dp, float delta_thresh, float *__restrict__ dx, State *state) {
  if (state->done) return;
  extern __shared__ double sm[];
  double *x = sm, *col = sm + Dp;  // Dp = max(D, 32): col doubles as the cross-warp scratch
  __shared__ int fail;
  const int tid = threadIdx.x;
  if (tid == 0) state->t[0] = gtimer();
  for (int i = tid; i < D; i += blockDim.x) x[i] = rhs[i];
  bool ok;
  if (PACKED) {
    double *Lp = sm + 2 * Dp;
    // one thread per element of the lower triangle (flat index -> (i,k)), coalesced along rows; eight independent
    // loads in flight per thread (one dependent L2 round trip per element cost 22 us at D = 217: 92 trips per thread)
    for (int t0 = tid; t0 < D * D; t0 += 8 * blockDim.x) {
      double v[8];
#pragma unroll
      for (int u = 0; u < 8; u++) {
        const int t = t0 + u * blockDim.x;
        const int i = t / D, k = t - i * D;
        v[u] = (t < D * D && k <= i) ? __ldcg(A + t) : 0.0;
      }
#pragma unroll
      for (int u = 0; u < 8; u++) {
        const int t = t0 + u * blockDim.x;
        const int i = t / D, k = t - i * D;
        if (t < D * D && k <= i) Lp[(size_t)i * (i + 1) / 2 + k] = v[u];
      }
    }
    __syncthreads();
    if (tid == 0) state->t[1] = gtimer();
    double *Pt = Lp + (size_t)D * (D + 1) / 2;
    const int ldp = chol_ldp(D);
    ok = cholesky_blocked(Lp, D, x, Pt, ldp, Pt + (size_t)CB * ldp, &fail);
    if (tid == 0) state->t[2] = gtimer();
  } else {
    FullLower L{A, D};
    __syncthreads();
    ok = cholesky_solve(L, D, x, col, &fail);
  }
  finish_update(ok, x, col, D, N, Twc, dx, delta_thresh, state);
  if (tid == 0) state->t[3] = gtimer();
}

// ---- fused tail for small problems (D <= SMALL_MAX_D): executed by the LAST stage-1 CTA of an
// iteration (ticket counter), so a whole GN iteration is ONE kernel launch. ---------------------
constexpr int SMALL_MAX_D = 63;   // N <= 10 poses
constexpr int SMALL_MAX_E = 96;

// Column c of the rotation matrix of unit quaternion q (xyzw), fp64 (actSO3 on a unit vector).
__device__ __forceinline__ void rot_col(const double *q, int c, double *Y) {
  const double X[3] = {c == 0 ? 1.0 : 0.0, c == 1 ? 1.0 : 0.0, c == 2 ? 1.0 : 0.0};
  const double uv0 = 2.0 * (q[1] * X[2] - q[2] * X[1]);
  const double uv1 = 2.0 * (q[2] * X[0] - q[0] * X[2]);
  const double uv2 = 2.0 * (q[0] * X[1] - q[1] * X[0]);
  Y[0] = X[0] + q[3] * uv0 + (q[1] * uv2 - q[2] * uv1);
  Y[1] = X[1] + q[3] * uv1 + (q[2] * uv0 - q[0] * uv2);
  Y[2] = X[2] + q[3] * uv2 + (q[0] * uv1 - q[1] * uv0);
}
// One entry M[k][c] of adj_inv_matrix (same expressions), so that 49 threads build M at once.
__device__ __forceinline__ double adj_inv_entry(const float *Ti, int k, int c) {
  if (k == 6) return c == 6 ? 1.0 : 0.0;
  const double q[4] = {Ti[3], Ti[4], Ti[5], Ti[6]};
  double Ra[3];
  rot_col(q, k < 3 ? k : k - 3, Ra);
  if (k >= 3) return (c >= 3 && c < 6) ? Ra[c - 3] : 0.0;
  const double t[3] = {Ti[0], Ti[1], Ti[2]};
  const double s_inv = 1.0 / (double)Ti[7];
  switch (c) {
    case 0: case 1: case 2: return s_inv * Ra[c];
    case 3: return s_inv * (t[1] * Ra[2] - t[2] * Ra[1]);
    case 4: return s_inv * (t[2] * Ra[0] - t[0] * Ra[2]);
    case 5: return s_inv * (t[0] * Ra[1] - t[1] * Ra[0]);
    default: return s_inv * (t[0] * Ra[0] + t[1] * Ra[1] + t[2] * Ra[2]);
  }
}

// Stage 2 for up to 8 edges by the whole CTA (256 threads) instead of one warp per edge: the S
// partials of every (edge, accumulator) pair are summed by NC threads in interleaved chunks and
// combined in fixed order; M is built one entry per thread.  sc[e] = [acc 36 | M 49 | T 49].
// out: (E,56) doubles [H_jj | g_j].  Deterministic (fixed summation order).
constexpr int SMALL_FIN_E = 8;
template <int NT>
__device__ inline void finalize_small(const float *Twc, const int64_t *ii_row, const double *partials, int S, int E,
                                      double *out, double (*sc)[FIN_SCRATCH], double *tmp) {
  const int tid = threadIdx.x, items = E * NACC;
  const int NC = NT / items > 0 ? NT / items : 1;
  // The pose row this thread's M entry needs (below) is requested BEFORE the partial sums: its two dependent L2 round
  // trips (edge row, then pose) overlap the partials' round trip instead of following it.
  float Tpre[8];
  if (tid < E * 49) {
    const float *Tp = Twc + 8 * (int)ii_row[tid / 49];
    if ((reinterpret_cast<uintptr_t>(Tp) & 15) == 0) {  // rows of a (N,8) tensor are 32-byte aligned unless it is an offset view
      const float4 lo = reinterpret_cast<const float4 *>(Tp)[0], hi = reinterpret_cast<const float4 *>(Tp)[1];
      Tpre[0] = lo.x; Tpre[1] = lo.y; Tpre[2] = lo.z; Tpre[3] = lo.w; Tpre[4] = hi.x; Tpre[5] = hi.y; Tpre[6] = hi.z; Tpre[7] = hi.w;
    } else {
#pragma unroll
      for (int k = 0; k < 8; k++) Tpre[k] = Tp[k];
    }
  }
  for (int t = tid; t < items * NC; t += NT) {
    const int item = t % items, chunk = t / items, e = item / NACC, l = item - e * NACC;
    const double *p = partials + (size_t)e * S * NACC + l;
    double sum = 0.0;
    for (int k0 = chunk; k0 < S; k0 += 12 * NC) {  // 12 independent loads in flight per round trip (tracker: S = 74, NC = 7: one trip)
      double v[12];
#pragma unroll
      for (int u = 0; u < 12; u++) {
        const int k = k0 + u * NC;
        v[u] = k < S ? __ldcg(p + (size_t)k * NACC) : 0.0;
      }
#pragma unroll
      for (int u = 0; u < 12; u++) sum += v[u];
    }
    tmp[t] = sum;
  }
  for (int t = tid; t < E * 49; t += NT) {
    const int e = t / 49, o = t - e * 49;
    sc[e][36 + o] = t == tid ? adj_inv_entry(Tpre, o / 7, o % 7) : adj_inv_entry(Twc + 8 * (int)ii_row[e], o / 7, o % 7);
  }
  __syncthreads();
  for (int t = tid; t < items; t += NT) {
    double s = 0.0;
    for (int c = 0; c < NC; c++) s += tmp[c * items + t];
    sc[t / NACC][t % NACC] = s;
  }
  __syncthreads();
  for (int t = tid; t < E * 49; t += NT) {  // T = H0 * M
    const int e = t / 49, o = t - e * 49, a = o / 7, c = o % 7;
    const double *acc = sc[e], *M = sc[e] + 36;
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < 7; k++) {
      const int lo = a < k ? a : k, hi = a < k ? k : a;
      s += acc[lo * 7 - lo * (lo - 1) / 2 + (hi - lo)] * M[k * 7 + c];
    }
    sc[e][36 + 49 + o] = s;
  }
  __syncthreads();
  for (int t = tid; t < E * 56; t += NT) {  // H_jj = M^T T, g_j = M^T g0
    const int e = t / 56, o = t - e * 56;
    const double *acc = sc[e], *M = sc[e] + 36, *T = sc[e] + 36 + 49;
    double s = 0.0;
    if (o < 49) {
      const int r = o / 7, c = o % 7;
#pragma unroll
      for (int k = 0; k < 7; k++) s += M[k * 7 + r] * T[k * 7 + c];
    } else {
      const int r = o - 49;
#pragma unroll
      for (int k = 0; k < 7; k++) s += M[k * 7 + r] * acc[28 + k];
    }
    out[t] = s;
  }
  __syncthreads();
}

// 7x7 SPD solve in fp64 by one thread; H upper triangle row-major (28).  false on a non-positive pivot.
__device__ inline bool solve7(const double *Hu, const double *g, double *x) {
  double L[7][7];
  bool ok = true;
#pragma unroll
  for (int a = 0; a < 7; a++) {
#pragma unroll
    for (int b = a; b < 7; b++) L[b][a] = Hu[a * 7 - a * (a - 1) / 2 + (b - a)];
  }
#pragma unroll
  for (int j = 0; j < 7; j++) {
    double d = L[j][j];
#pragma unroll
    for (int k = 0; k < j; k++) d -= L[j][k] * L[j][k];
    ok = ok && (d > 0.0);
    const double dinv = rsqrt(d);  // one MUFU + Newton steps instead of a square root AND a division per column (as gn_chol_blocked.cuh)
    L[j][j] = dinv;  // reciprocal diagonal
#pragma unroll
    for (int i = j + 1; i < 7; i++) {
      double s = L[i][j];
#pragma unroll
      for (int k = 0; k < j; k++) s -= L[i][k] * L[j][k];
      L[i][j] = s * dinv;
    }
  }
  double y[7];
#pragma unroll
  for (int i = 0; i < 7; i++) {
    double s = g[i];
#pragma unroll
    for (int k = 0; k < i; k++) s -= L[i][k] * y[k];
    y[i] = s * L[i][i];
  }
#pragma unroll
  for (int i = 6; i >= 0; i--) {
    double s = y[i];
#pragma unroll
    for (int k = i + 1; k < 7; k++) s -= L[k][i] * x[k];
    x[i] = s * L[i][i];
  }
  return ok;
}

template <int NT>
__device__ inline void small_tail(float *Twc, const int64_t *ii_row, const int64_t *jj_row,
                                  const double *partials, int S, int E, int N, float delta_thresh, float *dx,
                                  double *edge_blocks, State *state) {
  __shared__ double sc[SMALL_FIN_E][FIN_SCRATCH];
  __shared__ double Ap[SMALL_MAX_D * (SMALL_MAX_D + 1) / 2];
  __shared__ double xs[64], cols[64];
  __shared__ int fail;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int D = 7 * (N - NUM_FIX);
  // finalize_small writes items * NC <= max(NT, SMALL_FIN_E * NACC) doubles (E = 8: 280 items > 256 threads)
  __shared__ double ebs[SMALL_FIN_E * 56], fin_tmp[NT > SMALL_FIN_E * NACC ? NT : SMALL_FIN_E * NACC];
  if (E <= SMALL_FIN_E) {  // few edges (the tracker): whole-CTA stage 2, blocks stay in shared memory
    finalize_small<NT>(Twc, ii_row, partials, S, E, ebs, sc, fin_tmp);
    edge_blocks = ebs;
  } else {
    if (warp < SMALL_FIN_E)
      for (int e = warp; e < E; e += SMALL_FIN_E) finalize_edge(Twc, ii_row, partials, S, e, e, edge_blocks + (size_t)e * 56, sc[warp], lane);
    __syncthreads();
  }
  if (tid == 0) state->t[1] = gtimer();
  if (D == 7) {
    // one free pose (the tracker's N = 2 graph): 28 + 7 sums over the edges, then the unrolled
    // single-thread 7x7 solve — the generic path below costs 4.4 us per iteration in barriers
    __shared__ double Hu7[28], g7[7];
    if (tid < 35) {
      int a = 0, b = tid;
      if (tid < 28) {
        while (b >= 7 - a) { b -= 7 - a; a++; }
        b += a;
      }
      double s = 0.0;
      for (int e = 0; e < E; e++) {
        const int i = (int)ii_row[e] - NUM_FIX, j = (int)jj_row[e] - NUM_FIX;
        if (tid < 28) {
          const double v = edge_blocks[(size_t)e * 56 + a * 7 + b];
          if (i == 0) s += v;
          if (i == 0 && j == 0) s -= 2.0 * v;
          if (j == 0) s += v;
        } else {
          const double v = edge_blocks[(size_t)e * 56 + 49 + (tid - 28)];
          if (i == 0) s -= v;
          if (j == 0) s += v;
        }
      }
      if (tid < 28) Hu7[tid] = s;
      else g7[tid - 28] = s;
    }
    __syncthreads();
    if (tid == 0) {
      double x[7];
      fail = solve7(Hu7, g7, x) ? 0 : 1;
#pragma unroll
      for (int k = 0; k < 7; k++) xs[k] = x[k];
      state->t[2] = gtimer();
    }
    __syncthreads();
    finish_update(fail == 0, xs, cols, D, N, Twc, dx, delta_thresh, state);
    if (tid == 0) state->t[3] = gtimer();
    return;
  }
  PackedLower L{Ap};
  for (int t = tid; t < D * D; t += NT) {
    const int r = t / D, c = t % D;
    if (c > r) continue;
    const int br = r / 7, kr = r % 7, bc = c / 7, kc = c % 7;
    double s = 0.0;
    for (int e = 0; e < E; e++) {
      const int i = (int)ii_row[e] - NUM_FIX, j = (int)jj_row[e] - NUM_FIX;
      const double v = edge_blocks[(size_t)e * 56 + kr * 7 + kc];
      if (i == br && i == bc) s += v;
      if (i == br && j == bc) s -= v;
      if (j == br && j == bc) s += v;
      if (j == br && i == bc) s -= v;
    }
    L.at(r, c) = s;
  }
  for (int r = tid; r < D; r += NT) {
    const int br = r / 7, kr = r % 7;
    double s = 0.0;
    for (int e = 0; e < E; e++) {
      const int i = (int)ii_row[e] - NUM_FIX, j = (int)jj_row[e] - NUM_FIX;
      const double v = edge_blocks[(size_t)e * 56 + 49 + kr];
      if (i == br) s -= v;
      if (j == br) s += v;
    }
    xs[r] = s;
  }
  __syncthreads();
  bool ok;
  if (D <= 32) {  // tiny system (tracker: 7 unknowns): one warp, no block barriers
    if (warp == 0) cholesky_solve<PackedLower, true>(L, D, xs, cols, &fail);
    __syncthreads();
    ok = fail == 0;
  } else {
    ok = cholesky_solve(L, D, xs, cols, &fail);
  }
  if (tid == 0) state->t[2] = gtimer();
  finish_update(ok, xs, cols, D, N, Twc, dx, delta_thresh, state);
  if (tid == 0) state->t[3] = gtimer();
}

}  // namespace b200gn
