// b200slam — Gauss-Newton driver for sm_100a: stage-1 build kernel, stage-2 per-edge
// finalisation (M^T H0 M in fp64), assembly, fp64 Cholesky, Sim3 retraction, convergence flag.
// Replaces gauss_newton_{rays,calib,points}_cuda + SparseBlock + pose_retr_kernel of the
// reference (gn_kernels.cu:56-158, :414-452, :724-810, :1139-1227, :1545-1637).
// The whole max_iter loop is enqueued on one stream with NO host round trip per iteration
// (the reference does D2H -> Eigen SimplicialLLT on the CPU -> H2D -> .item() each iteration).
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "../../include/b200slam.h"
#include "gn_build.cuh"

namespace b200gn {

// ---- workspace layout -------------------------------------------------------------------------
struct State {  // head of the workspace
  int iters;      // iterations executed so far
  int done;       // set when ||dx|| < delta_thresh: later launches are no-ops
  int chol_fail;  // number of failed factorisations (dx = 0)
  int pad;
};
constexpr size_t STATE_BYTES = 256;
constexpr int MAX_SPLITS = 296;

__host__ __device__ inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

struct Layout {
  size_t partials, edge_blocks, A, rhs, total;
  int S;
};

inline int choose_splits(int E, int n) {
  // aim for >= 2 CTAs per SM (148 SMs) but never fewer than 1024 points per CTA
  int S = (2 * 148 + E - 1) / E;
  const int maxS = (n + 1023) / 1024;
  if (S > maxS) S = maxS;
  if (S > MAX_SPLITS) S = MAX_SPLITS;
  if (S < 1) S = 1;
  return S;
}

inline Layout make_layout(int N, int E, int n) {
  Layout L;
  L.S = choose_splits(E > 0 ? E : 1, n);
  const size_t D = (size_t)7 * (N > 1 ? N - 1 : 1);
  size_t off = STATE_BYTES;
  L.partials = off; off = align256(off + (size_t)E * L.S * NACC * sizeof(double));
  L.edge_blocks = off; off = align256(off + (size_t)E * 56 * sizeof(double));
  L.A = off; off = align256(off + D * D * sizeof(double));
  L.rhs = off; off = align256(off + D * sizeof(double));
  L.total = off;
  return L;
}

// ---- stage 1 ----------------------------------------------------------------------------------
template <int KIND>
__global__ void __launch_bounds__(BUILD_THREADS)
build_kernel(const float *__restrict__ Twc, const float *__restrict__ Xs, const float *__restrict__ Cs,
             const int64_t *__restrict__ ii_row, const int64_t *__restrict__ jj_row,
             const int64_t *__restrict__ idx, const uint8_t *__restrict__ valid,
             const float *__restrict__ Q, int n, int e0, BuildConsts C, double *__restrict__ partials,
             const State *__restrict__ state) {
  if (state->done) return;
  const int e = e0 + blockIdx.y;
  const int S = gridDim.x, s = blockIdx.x;
  const int ix = (int)ii_row[e], jx = (int)jj_row[e];
  const EdgePose P = rel_pose(Twc + 8 * ix, Twc + 8 * jx);

  const float *Xi_base = Xs + (size_t)ix * n * 3, *Xj_base = Xs + (size_t)jx * n * 3;
  const float *Ci_base = Cs + (size_t)ix * n, *Cj_base = Cs + (size_t)jx * n;
  const int64_t *idx_e = idx + (size_t)e * n;
  const uint8_t *val_e = valid + (size_t)e * n;
  const float *Q_e = Q + (size_t)e * n;

  float acc[NACC];
#pragma unroll
  for (int l = 0; l < NACC; l++) acc[l] = 0.0f;

  const int groups = n >> 2;  // groups of 4 points
  const int gper = (groups + S - 1) / S;
  const int g_begin = s * gper, g_end = min(groups, g_begin + gper);
  for (int g = g_begin + threadIdx.x; g < g_end; g += BUILD_THREADS) {
    const uint32_t vm4 = __ldg(reinterpret_cast<const uint32_t *>(val_e) + g);
    if (vm4 == 0) continue;
    const longlong2 ia = __ldg(reinterpret_cast<const longlong2 *>(idx_e) + 2 * g);
    const longlong2 ib = __ldg(reinterpret_cast<const longlong2 *>(idx_e) + 2 * g + 1);
    const float4 q4 = __ldg(reinterpret_cast<const float4 *>(Q_e) + g);
    const float4 c4 = __ldg(reinterpret_cast<const float4 *>(Cj_base) + g);
    const float4 xa = __ldg(reinterpret_cast<const float4 *>(Xj_base) + 3 * g);
    const float4 xb = __ldg(reinterpret_cast<const float4 *>(Xj_base) + 3 * g + 1);
    const float4 xc = __ldg(reinterpret_cast<const float4 *>(Xj_base) + 3 * g + 2);
    const long long ind[4] = {ia.x, ia.y, ib.x, ib.y};
    const float qq[4] = {q4.x, q4.y, q4.z, q4.w};
    const float cj[4] = {c4.x, c4.y, c4.z, c4.w};
    const float Xj[4][3] = {{xa.x, xa.y, xa.z}, {xa.w, xb.x, xb.y}, {xb.z, xb.w, xc.x}, {xc.y, xc.z, xc.w}};
    float Xi[4][3], ci[4];
#pragma unroll
    for (int t = 0; t < 4; t++) {  // issue all gathers first (memory-level parallelism)
      const bool vm = (vm4 >> (8 * t)) & 0xffu;
      const long long id = vm ? ind[t] : 0;
      Xi[t][0] = __ldg(Xi_base + 3 * id);
      Xi[t][1] = __ldg(Xi_base + 3 * id + 1);
      Xi[t][2] = __ldg(Xi_base + 3 * id + 2);
      ci[t] = __ldg(Ci_base + id);
    }
#pragma unroll
    for (int t = 0; t < 4; t++)
      if ((vm4 >> (8 * t)) & 0xffu) point_contrib<KIND>(acc, P, C, Xi[t], ci[t], Xj[t], cj[t], qq[t], ind[t]);
  }
  // tail points (n % 4) handled by the last split, one point per thread
  if (s == S - 1) {
    for (int k = (groups << 2) + threadIdx.x; k < n; k += BUILD_THREADS) {
      if (!val_e[k]) continue;
      const long long id = idx_e[k];
      const float Xi[3] = {Xi_base[3 * id], Xi_base[3 * id + 1], Xi_base[3 * id + 2]};
      const float Xj[3] = {Xj_base[3 * (size_t)k], Xj_base[3 * (size_t)k + 1], Xj_base[3 * (size_t)k + 2]};
      point_contrib<KIND>(acc, P, C, Xi, Ci_base[id], Xj, Cj_base[k], Q_e[k], id);
    }
  }
  // warp tree (fp32) -> cross-warp sum (fp64, fixed order) -> per-CTA partial
  __shared__ float red[BUILD_THREADS / 32][NACC];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int l = 0; l < NACC; l++) {
    float v = acc[l];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) red[warp][l] = v;
  }
  __syncthreads();
  if (threadIdx.x < NACC) {
    double sum = 0.0;
#pragma unroll
    for (int wq = 0; wq < BUILD_THREADS / 32; wq++) sum += (double)red[wq][threadIdx.x];
    partials[((size_t)e * S + s) * NACC + threadIdx.x] = sum;
  }
}

// ---- stage 2: per-edge reduce over splits + M^T H0 M in fp64 -----------------------------------
// M rows = apply_Sim3_adj_inv(T_i) applied to unit vectors (gn_kernels.cu:276-296), in fp64.
__device__ void adj_inv_matrix(const float *Ti, double (&M)[7][7]) {
  const double t[3] = {Ti[0], Ti[1], Ti[2]}, q[4] = {Ti[3], Ti[4], Ti[5], Ti[6]};
  const double s_inv = 1.0 / (double)Ti[7];
  // rotation matrix of q (same formula as actSO3 expanded)
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
  for (int k = 0; k < 7; k++)
    for (int c = 0; c < 7; c++) M[k][c] = 0.0;
  for (int k = 0; k < 3; k++) {  // input component a_k: Ra = R[:,k]
    const double Ra[3] = {R[0][k], R[1][k], R[2][k]};
    M[k][0] = s_inv * Ra[0]; M[k][1] = s_inv * Ra[1]; M[k][2] = s_inv * Ra[2];
    M[k][3] = s_inv * (t[1] * Ra[2] - t[2] * Ra[1]);
    M[k][4] = s_inv * (t[2] * Ra[0] - t[0] * Ra[2]);
    M[k][5] = s_inv * (t[0] * Ra[1] - t[1] * Ra[0]);
    M[k][6] = s_inv * (t[0] * Ra[0] + t[1] * Ra[1] + t[2] * Ra[2]);
  }
  for (int k = 0; k < 3; k++) {  // input component b_k: Rb = R[:,k]
    M[3 + k][3] = R[0][k]; M[3 + k][4] = R[1][k]; M[3 + k][5] = R[2][k];
  }
  M[6][6] = 1.0;
}

// one warp per edge; lanes cooperate on the 35 sums, lane 0 does the 7x7x7 transform
__global__ void __launch_bounds__(128)
finalize_kernel(const float *__restrict__ Twc, const int64_t *__restrict__ ii_row,
                const double *__restrict__ partials, int S, int e0, int Ecount,
                double *__restrict__ edge_blocks, const State *__restrict__ state) {
  if (state->done) return;
  const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (wid >= Ecount) return;
  const int e = e0 + wid;
  __shared__ double sh[4][2][NACC + 1];
  double(&acc)[NACC + 1] = sh[(threadIdx.x >> 5)][0];
  for (int l = lane; l < NACC; l += 32) {
    double s = 0.0;
    for (int k = 0; k < S; k++) s += partials[((size_t)e * S + k) * NACC + l];
    acc[l] = s;
  }
  __syncwarp();
  if (lane == 0) {
    double M[7][7], H0[7][7], T[7][7];
    adj_inv_matrix(Twc + 8 * (int)ii_row[e], M);
    int l = 0;
    for (int a = 0; a < 7; a++)
      for (int b = a; b < 7; b++, l++) H0[a][b] = H0[b][a] = acc[l];
    for (int a = 0; a < 7; a++)
      for (int c = 0; c < 7; c++) {
        double s = 0.0;
        for (int k = 0; k < 7; k++) s += H0[a][k] * M[k][c];
        T[a][c] = s;
      }
    double *out = edge_blocks + (size_t)e * 56;
    for (int r = 0; r < 7; r++)
      for (int c = r; c < 7; c++) {
        double s = 0.0;
        for (int k = 0; k < 7; k++) s += M[k][r] * T[k][c];
        out[r * 7 + c] = s;
        out[c * 7 + r] = s;
      }
    for (int r = 0; r < 7; r++) {
      double s = 0.0;
      for (int k = 0; k < 7; k++) s += M[k][r] * acc[28 + k];
      out[49 + r] = s;
    }
  }
}

// reference-layout export: Hs (4,E,7,7) f32, gs (2,E,7) f32 (gn_kernels.cu:1119-1132)
__global__ void export_blocks_kernel(const double *__restrict__ edge_blocks, int E, int e0, int Ecount,
                                     float *__restrict__ Hs, float *__restrict__ gs) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= Ecount * 56) return;
  const int e = e0 + t / 56, c = t % 56;
  const double v = edge_blocks[(size_t)e * 56 + c];
  if (c < 49) {
    Hs[((size_t)0 * E + e) * 49 + c] = (float)v;
    Hs[((size_t)1 * E + e) * 49 + c] = (float)-v;
    Hs[((size_t)2 * E + e) * 49 + c] = (float)-v;
    Hs[((size_t)3 * E + e) * 49 + c] = (float)v;
  } else {
    gs[((size_t)0 * E + e) * 7 + (c - 49)] = (float)-v;
    gs[((size_t)1 * E + e) * 7 + (c - 49)] = (float)v;
  }
}

}  // namespace b200gn

#include "gn_solve.cuh"
#include "gn_api.cuh"
