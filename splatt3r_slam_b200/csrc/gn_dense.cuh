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
  double *b;       // (D) rhs in, y = L^-1 b out
  double *x;       // (D) solution of L^T x = y
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

// Column owned by accumulator slot c of thread column tx: {2tx, 2tx+1, 32+2tx, 33+2tx}.  With 4tx + c a
// half-warp's 16-byte B reads sat at a 32-byte stride (2-way bank conflict: 4 wavefronts per load);
// this way they are contiguous (2 wavefronts), and the C-tile rows are read/written as 16-byte pairs.
__device__ __forceinline__ int dcol(int tx, int c) { return (c < 2 ? 0 : 30) + 2 * tx + c; }

// acc[a][c] = sum_k As[k][4*ty + a] * Bs[k][dcol(tx, c)]   (C = A B^T on the staged tiles)
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
    const double2 b01 = *reinterpret_cast<const double2 *>(Bs + k * DLD + 2 * tx);
    const double2 b23 = *reinterpret_cast<const double2 *>(Bs + k * DLD + 32 + 2 * tx);
    const double av[4] = {a01.x, a01.y, a23.x, a23.y}, bv[4] = {b01.x, b01.y, b23.x, b23.y};
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
      for (int c = 0; c < 4; c++) acc[a][c] += av[a] * bv[c];
  }
}

// Diagonal tile kk by one CTA (256 threads): Cholesky L, its inverse W = L^-1 and y = W b, in shared
// memory, blocked by 16 so that every step is a 256-thread data-parallel sweep with ONE barrier:
//   per 16-column block  (a) right-looking elimination of the 16x16 diagonal block, one thread per
//                            element, 16 rank-1 steps, normalised at the end (L_rc = a_rc / sqrt(a_cc))
//                        (b) its inverse, one thread per column (registers, 136 FMA)
//                        (c) rows below: L_rb = A_rb W_bb^T   (d) trailing A_rc -= L_rb L_cb^T
//   then the off-diagonal blocks of W by distance from the diagonal:
//                        W_ab = -W_aa (sum_{b<=c<a} L_ac W_cb), temporaries in T's unused upper triangle.
// ~25 k cycles per tile; the column-panel routine of the single-CTA solver (built for D = 217, not 64)
// plus a row-by-row inversion took 107-143 k and bounded every graph below ~500 keyframes.
constexpr int DB = 16;  // block edge inside the diagonal tile
__device__ __noinline__ void dense_factor_diag(const DenseBufs &B, int kk, int D, double *T, double *W, double *v0,
                                               int *fail) {
  const int tid = threadIdx.x, k0 = kk * DT, ts = min(DT, D - k0);
  const int r16 = tid >> 4, c16 = tid & 15;
  const size_t ld = (size_t)D;
  __shared__ double dinv[DT];
  // tile (lower triangle; identity padding beyond ts), W = 0, rhs segment
  for (int t = tid; t < DT * DT; t += DENSE_THREADS) {
    const int i = t >> 6, k = t & 63;
    double v = i == k ? 1.0 : 0.0;
    if (i < ts && k <= i) v = __ldcg(B.A + (size_t)(k0 + i) * ld + k0 + k);
    T[i * DLD + k] = k <= i ? v : 0.0;
    W[i * DLD + k] = 0.0;
  }
  if (tid < DT) v0[tid] = tid < ts ? __ldcg(B.b + k0 + tid) : 0.0;
  if (tid == 0) *fail = 0;
  __syncthreads();
  for (int b = 0; b < DT / DB; b++) {
    const int o = b * DB;
    // (a) eliminate the diagonal block
    for (int j = 0; j < DB - 1; j++) {
      if (c16 > j && r16 >= c16) {
        const double piv = T[(o + j) * DLD + o + j];
        T[(o + r16) * DLD + o + c16] -= T[(o + r16) * DLD + o + j] * T[(o + c16) * DLD + o + j] / piv;
      }
      __syncthreads();
    }
    if (tid < DB) {
      const double piv = T[(o + tid) * DLD + o + tid];
      if (!(piv > 0.0)) *fail = 1;
      dinv[o + tid] = rsqrt(piv);
    }
    __syncthreads();
    if (r16 >= c16) T[(o + r16) * DLD + o + c16] *= dinv[o + c16];  // diagonal: a / sqrt(a) = L_cc
    __syncthreads();
    // (b) W_bb: thread c owns column c (uniform loops: w[k] = 0 above the diagonal)
    if (tid < DB) {
      double w[DB];
#pragma unroll
      for (int i = 0; i < DB; i++) {
        double s = i == tid ? 1.0 : 0.0;
#pragma unroll
        for (int k = 0; k < i; k++) s -= T[(o + i) * DLD + o + k] * w[k];
        w[i] = i < tid ? 0.0 : s * dinv[o + i];
      }
#pragma unroll
      for (int i = 0; i < DB; i++) W[(o + i) * DLD + o + tid] = w[i];
    }
    __syncthreads();
    // (c) rows below the block: L_rb = A_rb W_bb^T (results held in registers until every input is read)
    const int below = DT - o - DB;
    double out[3];
#pragma unroll
    for (int u = 0; u < 3; u++) {
      const int idx = tid + u * DENSE_THREADS;
      out[u] = 0.0;
      if (idx < below * DB) {
        const int r = o + DB + (idx >> 4), c = idx & 15;
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < DB; k++) s += T[r * DLD + o + k] * W[(o + c) * DLD + o + k];
        out[u] = s;
      }
    }
    __syncthreads();
#pragma unroll
    for (int u = 0; u < 3; u++) {
      const int idx = tid + u * DENSE_THREADS;
      if (idx < below * DB) T[(o + DB + (idx >> 4)) * DLD + o + (idx & 15)] = out[u];
    }
    __syncthreads();
    // (d) trailing blocks: A_rc -= L_rb L_cb^T, r >= c > block b
    for (int idx = tid; idx < below * below; idx += DENSE_THREADS) {
      const int r = o + DB + idx / below, c = o + DB + idx % below;
      if (c > r) continue;
      double s = 0.0;
#pragma unroll
      for (int k = 0; k < DB; k++) s += T[r * DLD + o + k] * T[c * DLD + o + k];
      T[r * DLD + c] -= s;
    }
    __syncthreads();
  }
  // off-diagonal blocks of W, by distance from the diagonal
  for (int d = 1; d < DT / DB; d++) {
    for (int bb = 0; bb + d < DT / DB; bb++) {  // tmp_ab = sum_c L_ac W_cb  ->  T[upper block (b,a)]
      const int a = bb + d, oa = a * DB, ob = bb * DB;
      double s = 0.0;
      for (int c = bb; c < a; c++) {
        const int oc = c * DB;
#pragma unroll
        for (int k = 0; k < DB; k++) s += T[(oa + r16) * DLD + oc + k] * W[(oc + k) * DLD + ob + c16];
      }
      T[(ob + r16) * DLD + oa + c16] = s;
    }
    __syncthreads();
    for (int bb = 0; bb + d < DT / DB; bb++) {  // W_ab = -W_aa tmp_ab
      const int a = bb + d, oa = a * DB, ob = bb * DB;
      double s = 0.0;
#pragma unroll
      for (int k = 0; k < DB; k++) s -= W[(oa + r16) * DLD + oa + k] * T[(ob + k) * DLD + oa + c16];
      W[(oa + r16) * DLD + ob + c16] = s;
    }
    __syncthreads();
  }
  if (*fail) {
    if (tid == 0) *B.flag = 1;
  } else {
    double *Wg = B.W + (size_t)kk * DT * DT;
    for (int t = tid; t < DT * DT; t += DENSE_THREADS) Wg[t] = W[(t >> 6) * DLD + (t & 63)];
    if (tid < ts) {  // y_kk = W b_kk
      double s = 0.0;
      for (int k = 0; k <= tid; k++) s += W[tid * DLD + k] * v0[k];
      B.b[k0 + tid] = s;
    }
  }
  __syncthreads();
}

// Trailing tile A_ij -= L_ik L_jk^T (column kk = k0 / DT).  staged_i: row tile currently held in As.
__device__ __forceinline__ void dense_trail_tile(const DenseBufs &B, int D, int k0, int ts, int i, int j, int &staged_i,
                                                 double *As, double *Bs) {
  const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
  const size_t ld = (size_t)D;
  const int rows_i = min(DT, D - i * DT), rows_j = min(DT, D - j * DT);
  if (i != staged_i) {
    stage_tile_T(As, B.A + (size_t)(i * DT) * ld + k0, ld, rows_i, ts);
    staged_i = i;
  }
  stage_tile_T(Bs, B.A + (size_t)(j * DT) * ld + k0, ld, rows_j, ts);
  __syncthreads();
  // the old values of the target tile are fetched BEFORE the product so their L2 latency hides under it
  double *ct = B.A + (size_t)(i * DT + 4 * ty) * ld + j * DT;
  double cold[4][4], acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; a++)
#pragma unroll
    for (int c = 0; c < 4; c++)
      cold[a][c] = (4 * ty + a < rows_i && dcol(tx, c) < rows_j) ? __ldcg(ct + (size_t)a * ld + dcol(tx, c)) : 0.0;
  tile_gemm(As, Bs, ts, acc);
#pragma unroll
  for (int a = 0; a < 4; a++)
#pragma unroll
    for (int c = 0; c < 4; c++)
      if (4 * ty + a < rows_i && dcol(tx, c) < rows_j) ct[(size_t)a * ld + dcol(tx, c)] = cold[a][c] - acc[a][c];
  __syncthreads();
}

__global__ void __launch_bounds__(DENSE_THREADS, 2)
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
  // cycle counts for tools/dense_probe.py: [0] diagonal factorisations (CTA 0), [1] panel work, [2] trailing
  // work, [3] whole kernel (CTA 1, or CTA 0 on a one-CTA grid)
  long long c_diag = 0, c_panel = 0, c_trail = 0, c_t;
  const long long c_start = clock64();
  const bool probe = cta == (G > 1 ? 1 : 0);

  // Schedule per tile column kk (the diagonal tile kk is already factored when the column starts):
  //   (2)  every CTA: panel tiles L_ik = A_ik W_kk^T and b_i -= L_ik y_kk
  //   (3a) every CTA: trailing tiles of column kk+1 only
  //   (3b) CTA 0 factors the diagonal tile kk+1 (serial: ~100 k cycles) WHILE the other CTAs do the
  //        rest of the trailing update — the look-ahead hides the serial step whenever a CTA's share
  //        of the trailing tiles takes longer than one diagonal factorisation.
  if (cta == 0) {
    c_t = clock64();
    dense_factor_diag(B, 0, D, As, Bs, v0, &fail);
    c_diag += clock64() - c_t;
  }
  grid_barrier(B.bar, gen);
  if (__ldcg(B.flag)) ok = false;
  for (int kk = 0; ok && kk < nt; kk++) {
    const int k0 = kk * DT, ts = min(DT, D - k0);
    // ---- (2) panel tiles L_ik = A_ik W^T, b_i -= L_ik y_kk ----------------------------------------------
    c_t = clock64();
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
          const int r = 4 * ty + a, cc = dcol(tx, c);
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
    c_panel += clock64() - c_t;
    grid_barrier(B.bar, gen);
    const int m = nt - kk - 1;
    if (m == 0) break;
    c_t = clock64();
    // ---- (3a) trailing tiles of column kk+1: A_{i,kk+1} -= L_ik L_{kk+1,k}^T ------------------------------
    {
      int staged_i = -1;
      for (int i = kk + 1 + cta; i < nt; i += G) dense_trail_tile(B, D, k0, ts, i, kk + 1, staged_i, As, Bs);
    }
    c_trail += clock64() - c_t;
    grid_barrier(B.bar, gen);
    c_t = clock64();
    // ---- (3b) look-ahead: CTA 0 factors diagonal tile kk+1, the others finish the trailing update ----------
    if (cta == 0) {
      dense_factor_diag(B, kk + 1, D, As, Bs, v0, &fail);
      c_diag += clock64() - c_t;
      c_t = clock64();
    }
    if (cta > 0 || G == 1) {
      const int WG = G == 1 ? 1 : G - 1, wc = G == 1 ? 0 : cta - 1, m2 = m - 1;
      const long long total = (long long)m2 * (m2 + 1) / 2;
      const long long per = (total + WG - 1) / WG, t_begin = per * wc, t_end = min(total, t_begin + per);
      int staged_i = -1;  // consecutive items share the row tile i: L_ik stays staged in As
      for (long long t = t_begin; t < t_end; t++) {
        int ip = (int)((sqrt(8.0 * (double)t + 1.0) - 1.0) * 0.5);
        while ((long long)(ip + 1) * (ip + 2) / 2 <= t) ip++;
        while ((long long)ip * (ip + 1) / 2 > t) ip--;
        const int jp = (int)(t - (long long)ip * (ip + 1) / 2);
        dense_trail_tile(B, D, k0, ts, kk + 2 + ip, kk + 2 + jp, staged_i, As, Bs);
      }
      c_trail += clock64() - c_t;
    }
    grid_barrier(B.bar, gen);
    if (__ldcg(B.flag)) ok = false;
  }
  if (tid == 0) {
    if (cta == 0) state->t[0] = (unsigned long long)c_diag;
    if (probe) {
      state->t[1] = (unsigned long long)c_panel;
      state->t[2] = (unsigned long long)c_trail;
      state->t[3] = (unsigned long long)(clock64() - c_start);
    }
  }
  // ---- backward substitution L^T x = y, one tile row per step from the bottom, ONE barrier per step:
  // every CTA with a tile in the row derives x_kk = W_kk^T y_kk itself (32 KB of W from L2; x goes to its own
  // array so nobody reads a half-written y_kk), then updates its y_i -= L_{kk,i}^T x_kk.  Loads are issued
  // 16 at a time per thread (the one-at-a-time version spent 45 k cycles per tile on L2 round trips).
  if (ok) {
    double *xs = v0, *red = As;  // xs[64]; red[4][64] partial sums
    const int c = tid & 63, part = tid >> 6;
    for (int kk = nt - 1; kk >= 0; kk--) {
      const int k0 = kk * DT, ts = min(DT, D - k0);
      if (cta < kk || cta == 0) {
        {
          const double *Wg = B.W + (size_t)kk * DT * DT;
          double w[16], y[16];
#pragma unroll
          for (int u = 0; u < 16; u++) {
            const int i = part * 16 + u;
            w[u] = __ldcg(Wg + i * DT + c);
            y[u] = i < ts ? __ldcg(B.b + k0 + i) : 0.0;
          }
          double sacc = 0.0;
#pragma unroll
          for (int u = 0; u < 16; u++) sacc += w[u] * y[u];
          red[part * DT + c] = sacc;
        }
        __syncthreads();
        if (tid < DT) {
          const double xv = (red[tid] + red[DT + tid]) + (red[2 * DT + tid] + red[3 * DT + tid]);
          xs[tid] = tid < ts ? xv : 0.0;
          if (cta == 0 && tid < ts) B.x[k0 + tid] = xv;
        }
        __syncthreads();
        for (int i = cta; i < kk; i += G) {  // y_i -= L_{kk,i}^T x_kk
          double lv[16];
#pragma unroll
          for (int u = 0; u < 16; u++) {
            const int r = part * 16 + u;
            lv[u] = r < ts ? __ldcg(B.A + (size_t)(k0 + r) * ld + i * DT + c) : 0.0;
          }
          double sacc = 0.0;
#pragma unroll
          for (int u = 0; u < 16; u++) sacc += lv[u] * xs[part * 16 + u];
          red[part * DT + c] = sacc;
          __syncthreads();
          if (tid < DT)
            B.b[i * DT + tid] = __ldcg(B.b + i * DT + tid) -
                                ((red[tid] + red[DT + tid]) + (red[2 * DT + tid] + red[3 * DT + tid]));
          __syncthreads();
        }
      }
      grid_barrier(B.bar, gen);
    }
  }
  if (cta == 0) finish_update(ok, B.x, fin_scratch, D, N, Twc, dx, delta_thresh, state);
}

}  // namespace b200gn
