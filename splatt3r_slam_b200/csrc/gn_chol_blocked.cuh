// b200slam — blocked fp64 Cholesky solve for one CTA, packed lower triangle in shared memory
// (7 < D <= 217: up to 32 keyframes).  Included by gn_solve.cuh.
//
// Why: the column-at-a-time factorisation ran 375 us at D = 217 on B200.  clock64 probes showed
// it is ISSUE-bound, not latency- or barrier-bound: with one warp per row of the trailing matrix
// every row costs ~70 instructions (address arithmetic, predication, LDS/STS) for <= 7 useful DFMA
// per lane: ~1 instruction per lane-FMA.  Here the trailing update is done per PANEL of CB = 8
// columns with 4 x 32 register tiles per warp (8 x (5 LDS + 4 DFMA) + 4 LDS + 4 STS per 1024
// lane-FMAs: 13x fewer instructions), and barriers drop from 2 per column to 3 per panel.
//
//   panel step 1  warp 0 factors the CB x CB diagonal block with register rows + shuffles (lane r owns
//                 row r, lane CB carries the right-hand side: forward substitution rides along)
//   panel step 2  one thread per row below the block: CB-step triangular solve -> L and the
//                 transposed panel buffer Pt[c][i]; x[i] -= sum_c L(i,c) y_c
//   panel step 3  trailing update A(i,k) -= sum_c Pt[c][i] Pt[c][k], 4 rows x 32 columns per warp tile
// Diagonal entries hold 1/L(j,j) (the solves multiply by it).  Returns false on a non-positive pivot.
#pragma once

namespace b200gn {

constexpr int CB = 8;

__device__ __forceinline__ double shfl_d(double v, int src) { return __shfl_sync(0xffffffffu, v, src); }

__device__ __forceinline__ size_t tri(int i) { return (size_t)i * (i + 1) / 2; }

// One panel of the factorisation (columns j0 .. j0+nb-1).  Returns false on a non-positive pivot.
template <bool FULL>
__device__ __forceinline__ bool chol_panel(double *__restrict__ Lp, int D, double *__restrict__ x,
                                           double *__restrict__ Pt, int ldp, double *__restrict__ Dg,
                                           double *__restrict__ yv, int *fail, int j0, int nb_rt) {
  const int tid = threadIdx.x, T = blockDim.x, lane = tid & 31, warp = tid >> 5, nwarp = T >> 5;
  const int nb = FULL ? CB : nb_rt, R0 = j0 + nb;  // FULL: compile-time width => straight-line, batched loads
  // ---- step 1: diagonal block (+ rhs) by ONE thread, fully unrolled in registers -------------------
  // (a shuffle-based warp version measured ~5 k cycles per panel: 64-bit shuffles + rsqrt chains)
  if (tid == 0) {
    double a[CB][CB], y[CB];
#pragma unroll
    for (int r = 0; r < CB; r++) {
#pragma unroll
      for (int c = 0; c <= r; c++) a[r][c] = r < nb ? Lp[tri(j0 + r) + j0 + c] : (r == c ? 1.0 : 0.0);
      y[r] = r < nb ? x[j0 + r] : 0.0;
    }
    bool bad = false;
#pragma unroll
    for (int c = 0; c < CB; c++) {
      const double d = a[c][c];
      if (!(d > 0.0)) bad = true;
      const double dinv = rsqrt(d);
      a[c][c] = dinv;
      y[c] *= dinv;
#pragma unroll
      for (int r = c + 1; r < CB; r++) a[r][c] *= dinv;
#pragma unroll
      for (int r = c + 1; r < CB; r++) {
#pragma unroll
        for (int c2 = c + 1; c2 <= r; c2++) a[r][c2] -= a[r][c] * a[c2][c];
        y[r] -= a[r][c] * y[c];
      }
    }
    if (bad) *fail = 1;
#pragma unroll
    for (int r = 0; r < CB; r++) {
      if (r < nb) {
#pragma unroll
        for (int c = 0; c <= r; c++) {
          Lp[tri(j0 + r) + j0 + c] = a[r][c];
          Dg[r * CB + c] = a[r][c];
        }
        x[j0 + r] = y[r];
        yv[r] = y[r];
      }
    }
  }
  __syncthreads();
  if (*fail) return false;
  // ---- step 2: rows below the block ---------------------------------------------------------------
  for (int i = R0 + tid; i < D; i += T) {
    double *rp = Lp + tri(i) + j0;
    double a[CB];
#pragma unroll
    for (int c = 0; c < CB; c++) a[c] = c < nb ? rp[c] : 0.0;
#pragma unroll
    for (int c = 0; c < CB; c++) {
      if (c < nb) {
        double s = a[c];
#pragma unroll
        for (int c1 = 0; c1 < c; c1++) s -= a[c1] * Dg[c * CB + c1];
        a[c] = s * Dg[c * CB + c];
      }
    }
    double xi = x[i];
#pragma unroll
    for (int c = 0; c < CB; c++) {
      if (c < nb) {
        rp[c] = a[c];
        Pt[c * ldp + i] = a[c];
        xi -= a[c] * yv[c];
      }
    }
    x[i] = xi;
  }
  __syncthreads();
  // ---- step 3: trailing update, warp tiles of 4 rows x 32 columns ---------------------------------
  const int rem = D - R0;
  const int G = (rem + 3) >> 2;
  for (int pass = 0; pass * nwarp < G; pass++) {
    // snake order: a warp that took a short row group in one pass takes a long one in the next
    const int gi = pass * nwarp + ((pass & 1) ? nwarp - 1 - warp : warp);
    if (gi >= G) continue;
    const int i0 = R0 + 4 * gi;
    const int nblk = ((4 * gi + 3) >> 5) + 1;  // column blocks with k0 <= i0 + 3
    for (int gk = 0; gk < nblk; gk++) {
      const int k = R0 + 32 * gk + lane;
      const bool kin = k < D;
      double acc0 = 0.0, acc1 = 0.0, acc2 = 0.0, acc3 = 0.0;
#pragma unroll
      for (int c = 0; c < CB; c++) {
        if (c < nb) {
          const double *pc = Pt + c * ldp;
          const double pk = kin ? pc[k] : 0.0;
          acc0 += pc[i0] * pk;      // Pt is padded: i0 + 3 < ldp (values beyond D are never stored)
          acc1 += pc[i0 + 1] * pk;
          acc2 += pc[i0 + 2] * pk;
          acc3 += pc[i0 + 3] * pk;
        }
      }
      if (kin) {
        if (i0 < D && k <= i0) Lp[tri(i0) + k] -= acc0;
        if (i0 + 1 < D && k <= i0 + 1) Lp[tri(i0 + 1) + k] -= acc1;
        if (i0 + 2 < D && k <= i0 + 2) Lp[tri(i0 + 2) + k] -= acc2;
        if (i0 + 3 < D && k <= i0 + 3) Lp[tri(i0 + 3) + k] -= acc3;
      }
    }
  }
  __syncthreads();
  return true;
}

template <bool FULL>
__device__ __forceinline__ void chol_back_panel(const double *__restrict__ Lp, double *__restrict__ x, int j0, int nb_rt) {
  const int tid = threadIdx.x, T = blockDim.x;
  const int nb = FULL ? CB : nb_rt;
  if (tid == 0) {  // CB x CB upper-triangular solve, serial (28 FMA)
    double xs[CB];
#pragma unroll
    for (int c = CB - 1; c >= 0; c--) {
      if (c < nb) {
        double s = x[j0 + c];
#pragma unroll
        for (int c1 = CB - 1; c1 > c; c1--)
          if (c1 < nb) s -= Lp[tri(j0 + c1) + j0 + c] * xs[c1];
        xs[c] = s * Lp[tri(j0 + c) + j0 + c];
        x[j0 + c] = xs[c];
      }
    }
  }
  __syncthreads();
  for (int i = tid; i < j0; i += T) {
    double s = x[i];
#pragma unroll
    for (int c = 0; c < CB; c++)
      if (c < nb) s -= Lp[tri(j0 + c) + i] * x[j0 + c];
    x[i] = s;
  }
  __syncthreads();
}

// Lp: packed lower triangle; x: rhs in / solution out; Pt: CB * ldp doubles; Dg: CB*CB + CB doubles.
__device__ inline bool cholesky_blocked(double *__restrict__ Lp, int D, double *__restrict__ x,
                                        double *__restrict__ Pt, int ldp, double *__restrict__ Dg, int *fail,
                                        bool do_back = true) {
  double *yv = Dg + CB * CB;
  if (threadIdx.x == 0) *fail = 0;
  __syncthreads();
  for (int j0 = 0; j0 < D; j0 += CB) {
    const int nb = min(CB, D - j0);
    const bool ok = nb == CB ? chol_panel<true>(Lp, D, x, Pt, ldp, Dg, yv, fail, j0, CB)
                             : chol_panel<false>(Lp, D, x, Pt, ldp, Dg, yv, fail, j0, nb);
    if (!ok) return false;
  }
  if (!do_back) return true;  // caller wants L (reciprocal diagonal) and y = L^-1 b only
  // backward substitution L^T x = y, panel by panel from the bottom
  for (int j0 = ((D - 1) / CB) * CB; j0 >= 0; j0 -= CB) {
    const int nb = min(CB, D - j0);
    if (nb == CB) chol_back_panel<true>(Lp, x, j0, CB);
    else chol_back_panel<false>(Lp, x, j0, nb);
  }
  return true;
}

}  // namespace b200gn
