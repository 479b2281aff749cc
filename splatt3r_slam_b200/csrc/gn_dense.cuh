// b200slam — stage 3 for graphs beyond the single-CTA solver (7(N-1) > 217): exact dense fp64
// Cholesky of the assembled normal equations by ALL SMs — one cooperative kernel per GN iteration,
// 64x64 tiles, right-looking:
//   per tile column kk   (1) CTA 0 factors the diagonal tile in shared memory (cholesky_blocked, the
//                            right-hand side rides along) and inverts it: W = L_kk^-1
//                        (2) every CTA: panel tiles L_ik = A_ik W^T (a tile GEMM) and b_i -= L_ik y_kk
//                        (3) every CTA: trailing tiles A_ij -= L_ik L_jk^T
//   then the backward substitution tile row by tile row (x_kk = W_kk^T y_kk; y_i -= L_ki^T x_kk),
//   dx = -x, Sim3 retraction and the ||dx|| test (finish_update).  Grid barriers between phases.
// Same semantics as the reference's host-side SimplicialLLT (gn_kernels.cu:131-152): exact
// factorisation, dx = 0 when a pivot is not positive.  The tile GEMM is fp64-pipe-bound (64 DFMA
// per clock per SM): ~2 us per tile per SM; D = 1785 (256 keyframes) ~ 1 ms, D = 7161 ~ 10 ms.
#pragma once

namespace b200gn {

constexpr int DT = 64;               // tile edge
constexpr int DLD = DT + 2;          // leading dimension of the transposed staging tiles (16-byte aligned rows)
constexpr int DENSE_THREADS = 256;   // 16 x 16 threads, 4 x 4 outputs each
constexpr int DENSE_SMEM_DOUBLES = 2 * DT * DLD + 2 * DT;  // As, Bs, two 64-vectors

struct DenseBufs {
  double *A;       // (D,D) row-major, lower triangle valid on entry, overwritten by L (off-diagonal tiles)
  double *b;       // (D) rhs in, solution out
  double *W;       // (nt, 64, 64) inverses of the diagonal tiles, row-major, zero above the diagonal
  unsigned *bar;   // grid barrier [counter, generation]
  int *flag;       // non-positive pivot
};

// S[k * DLD + r] = G[r * ld + k]  (transposed staging; zero outside rows x cols).  All 16 loads of a
// thread are issued before the first store: one L2 round trip per tile instead of sixteen (the
// one-at-a-time version made a tile cost ~44 k cycles against 4 k for its 64^3 DFMAs).
__device__ __forceinline__ void stage_tile_T(double *__restrict__ S, const double *__restrict__ G, size_t ld, int rows,
                                             int cols) {
  double v[DT * DT / DENSE_THREADS];
#pragma unroll
  for (int u = 0; u < DT * DT / DENSE_THREADS; u++) {
    const int t = threadIdx.x + u * DENSE_THREADS, r = t >> 6, k = t & 63;
    v[u] = (r < rows && k < cols) ? __ldcg(G + (size_t)r * ld + k) : 0.0;
  }
#pragma unroll
  for (int u = 0; u < DT * DT / DENSE_THREADS; u++) {
    const int t = threadIdx.x + u * DENSE_THREADS, r = t >> 6, k = t & 63;
    S[k * DLD + r] = v[u];
  }
}

// acc[a][c] = sum_k As[k][4*ty + a] * Bs[k][4*tx + c]   (C = A B^T on the staged tiles)
__device__ __forceinline__ void tile_gemm(const double *__restrict__ As, const double *__restrict__ Bs, int kmax,
                                          double (&acc)[4][4]) {
  const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
#pragma unroll
  for (int a = 0; a < 4; a++)
#pragma unroll
    for (int c = 0; c < 4; c++) acc[a][c] = 0.0;
#pragma unroll 4
  for (int k = 0; k < kmax; k++) {
    const double2 a01 = *reinterpret_cast<const double2 *>(As + k * DLD + 4 * ty);
    const double2 a23 = *reinterpret_cast<const double2 *>(As + k * DLD + 4 * ty + 2);
    const double2 b01 = *reinterpret_cast<const double2 *>(Bs + k * DLD + 4 * tx);
    const double2 b23 = *reinterpret_cast<const double2 *>(Bs + k * DLD + 4 * tx + 2);
    const double av[4] = {a01.x, a01.y, a23.x, a23.y}, bv[4] = {b01.x, b01.y, b23.x, b23.y};
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
      for (int c = 0; c < 4; c++) acc[a][c] += av[a] * bv[c];
  }
}

__global__ void __launch_bounds__(DENSE_THREADS)
dense_solve_retract_kernel(float *Twc, DenseBufs B, int N, int D, float delta_thresh, float *dx, State *state) {
  if (state->done) return;  // written only at the end of an earlier launch: uniform over the grid
  extern __shared__ __align__(16) double dsm[];
  double *As = dsm, *Bs = dsm + DT * DLD, *v0 = dsm + 2 * DT * DLD;
  __shared__ int fail;
  __shared__ double fin_scratch[64];
  const int tid = threadIdx.x, G = gridDim.x, cta = blockIdx.x;
  const int ty = tid >> 4, tx = tid & 15;
  const int nt = (D + DT - 1) / DT;
  const size_t ld = (size_t)D;
  unsigned gen = 0;
  if (tid == 0) gen = ld_acquire_u32(&B.bar[1]);
  bool ok = true;
  double acc[4][4];

  for (int kk = 0; kk < nt; kk++) {
    const int k0 = kk * DT, ts = min(DT, D - k0);
    // ---- (1) diagonal tile: factor (+ forward substitution of its rhs segment), invert -----------------
    if (cta == 0) {
      double *Lp = As, *Pt = As + 2112, *Dg = Pt + CB * 72;  // 2080 + 576 + 72 doubles inside As
      for (int t = tid; t < ts * ts; t += DENSE_THREADS) {
        const int i = t / ts, k = t - i * ts;
        if (k <= i) Lp[tri(i) + k] = __ldcg(B.A + (size_t)(k0 + i) * ld + k0 + k);
      }
      if (tid < ts) v0[tid] = __ldcg(B.b + k0 + tid);
      for (int t = tid; t < DT * DLD; t += DENSE_THREADS) Bs[t] = 0.0;
      __syncthreads();
      if (!cholesky_blocked(Lp, ts, v0, Pt, 72, Dg, &fail, false)) {
        if (tid == 0) *B.flag = 1;
      } else {
        // W = L^-1 row by row: thread (c, q) = (tid >> 2, tid & 3) sums the k = q (mod 4) slice of
        // W[i][c] = -(1/L_ii) sum_{k=c}^{i-1} L[i][k] W[k][c]; the 4 slices are combined by shuffles
        {
          const int c = tid >> 2, q = tid & 3;
          if (q == 0 && c < ts) Bs[c * DLD + c] = Lp[tri(c) + c];
          __syncthreads();
          for (int i = 1; i < ts; i++) {
            double s = 0.0;
            if (c < i)
              for (int k = c + q; k < i; k += 4) s -= Lp[tri(i) + k] * Bs[k * DLD + c];
            s += __shfl_xor_sync(0xffffffffu, s, 1);
            s += __shfl_xor_sync(0xffffffffu, s, 2);
            if (q == 0 && c < i) Bs[i * DLD + c] = s * Lp[tri(i) + i];
            __syncthreads();
          }
        }
        double *Wg = B.W + (size_t)kk * DT * DT;
        for (int t = tid; t < DT * DT; t += DENSE_THREADS) Wg[t] = Bs[(t >> 6) * DLD + (t & 63)];
        if (tid < ts) B.b[k0 + tid] = v0[tid];
      }
    }
    grid_barrier(B.bar, gen);
    if (__ldcg(B.flag)) { ok = false; break; }
    // ---- (2) panel tiles L_ik = A_ik W^T, b_i -= L_ik y_kk ----------------------------------------------
    for (int i = kk + 1 + cta; i < nt; i += G) {
      const int rows = min(DT, D - i * DT);
      stage_tile_T(As, B.A + (size_t)(i * DT) * ld + k0, ld, rows, ts);
      stage_tile_T(Bs, B.W + (size_t)kk * DT * DT, DT, DT, DT);
      if (tid < DT) v0[tid] = tid < ts ? __ldcg(B.b + k0 + tid) : 0.0;
      __syncthreads();
      tile_gemm(As, Bs, ts, acc);
      __syncthreads();
#pragma unroll
      for (int a = 0; a < 4; a++)
#pragma unroll
        for (int c = 0; c < 4; c++) {
          const int r = 4 * ty + a, cc = 4 * tx + c;
          As[r * DLD + cc] = acc[a][c];
          if (r < rows && cc < ts) B.A[(size_t)(i * DT + r) * ld + k0 + cc] = acc[a][c];
        }
      __syncthreads();
      if (tid < rows) {
        double s = 0.0;
        for (int c = 0; c < ts; c++) s += As[tid * DLD + c] * v0[c];
        B.b[i * DT + tid] = __ldcg(B.b + i * DT + tid) - s;
      }
      __syncthreads();
    }
    grid_barrier(B.bar, gen);
    // ---- (3) trailing tiles A_ij -= L_ik L_jk^T, kk < j <= i ----------------------------------------------
    const int m = nt - kk - 1;
    const long long total = (long long)m * (m + 1) / 2;
    const long long per = (total + G - 1) / G, t_begin = per * cta, t_end = min(total, t_begin + per);
    int staged_i = -1;  // consecutive items share the row tile i: L_ik stays staged in As
    for (long long t = t_begin; t < t_end; t++) {
      int ip = (int)((sqrt(8.0 * (double)t + 1.0) - 1.0) * 0.5);
      while ((long long)(ip + 1) * (ip + 2) / 2 <= t) ip++;
      while ((long long)ip * (ip + 1) / 2 > t) ip--;
      const int jp = (int)(t - (long long)ip * (ip + 1) / 2);
      const int i = kk + 1 + ip, j = kk + 1 + jp;
      const int rows_i = min(DT, D - i * DT), rows_j = min(DT, D - j * DT);
      if (i != staged_i) {
        stage_tile_T(As, B.A + (size_t)(i * DT) * ld + k0, ld, rows_i, ts);
        staged_i = i;
      }
      stage_tile_T(Bs, B.A + (size_t)(j * DT) * ld + k0, ld, rows_j, ts);
      __syncthreads();
      // the old values of the target tile are fetched BEFORE the product so their L2 latency hides under it
      double *ct = B.A + (size_t)(i * DT + 4 * ty) * ld + j * DT + 4 * tx;
      double cold[4][4];
#pragma unroll
      for (int a = 0; a < 4; a++)
#pragma unroll
        for (int c = 0; c < 4; c++)
          cold[a][c] = (4 * ty + a < rows_i && 4 * tx + c < rows_j) ? __ldcg(ct + (size_t)a * ld + c) : 0.0;
      tile_gemm(As, Bs, ts, acc);
#pragma unroll
      for (int a = 0; a < 4; a++)
#pragma unroll
        for (int c = 0; c < 4; c++)
          if (4 * ty + a < rows_i && 4 * tx + c < rows_j) ct[(size_t)a * ld + c] = cold[a][c] - acc[a][c];
      __syncthreads();
    }
    grid_barrier(B.bar, gen);
  }
  // ---- backward substitution L^T x = y, one tile row at a time from the bottom --------------------------
  if (ok) {
    for (int kk = nt - 1; kk >= 0; kk--) {
      const int k0 = kk * DT, ts = min(DT, D - k0);
      if (cta == 0) {  // x_kk = W_kk^T y_kk
        if (tid < DT) v0[tid] = tid < ts ? __ldcg(B.b + k0 + tid) : 0.0;
        __syncthreads();
        if (tid < ts) {
          const double *Wg = B.W + (size_t)kk * DT * DT;
          double s = 0.0;
          for (int i = tid; i < ts; i++) s += __ldcg(Wg + i * DT + tid) * v0[i];
          B.b[k0 + tid] = s;
        }
      }
      grid_barrier(B.bar, gen);
      for (int i = cta; i < kk; i += G) {  // y_i -= L_{kk,i}^T x_kk
        if (tid < DT) v0[tid] = tid < ts ? __ldcg(B.b + k0 + tid) : 0.0;
        __syncthreads();
        if (tid < DT) {
          double s = 0.0;
          for (int r = 0; r < ts; r++) s += __ldcg(B.A + (size_t)(k0 + r) * ld + i * DT + tid) * v0[r];
          B.b[i * DT + tid] = __ldcg(B.b + i * DT + tid) - s;
        }
        __syncthreads();
      }
      grid_barrier(B.bar, gen);
    }
  }
  if (cta == 0) finish_update(ok, B.b, fin_scratch, D, N, Twc, dx, delta_thresh, state);
}

}  // namespace b200gn
