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
constexpr int GLD = DT + 4;          // leading dimension of the row-major GEMM operand tiles: 8 rows x 4 k-columns of an
                                     // mma fragment land on 32 distinct banks (row stride 136 words = 8 mod 32)
constexpr int DENSE_SMEM_DOUBLES = 2 * DT * GLD + 2 * DT;  // As, Bs, two 64-vectors

struct DenseBufs {
  double *A;       // (D,D) row-major, lower triangle valid on entry, overwritten by L (off-diagonal tiles)
  double *b;       // (D) rhs in, y = L^-1 b out
  double *x;       // (D) solution of L^T x = y
  double *W;       // (nt, 64, 64) inverses of the diagonal tiles, row-major, zero above the diagonal
  unsigned *bar;   // grid barrier [counter, generation]
  int *flag;       // non-positive pivot
};

// S[r * GLD + k] = G[r * ld + k]  (zero outside rows x cols).  All 16 loads of a thread are issued before
// the first store: one L2 round trip per tile instead of sixteen.
__device__ __forceinline__ void stage_tile(double *__restrict__ S, const double *__restrict__ G, size_t ld, int rows,
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
    S[r * GLD + k] = v[u];
  }
}

// fp64 tensor-core tile product C = A B^T on staged row-major operands As[r][k], Bs[c][k] (64 x 64 x kmax):
// mma.sync.m8n8k4 (DMMA).  Warp w owns rows 16*(w>>1)..+16 and columns 32*(w&1)..+32 = 2 x 4 mma tiles; per
// 4-wide k step it loads 2 + 4 fragments (one double per lane each, conflict-free) for 8 DMMAs = 2048 FMAs.
// The scalar version (4 x 4 register tiles) needed four 128-bit shared loads per 16 DFMAs and was bound by
// shared-memory wavefronts at a third of the fp64 peak.
// acc[mi][ni][0..1] = C[16*(w>>1) + 8*mi + (lane>>2)][32*(w&1) + 8*ni + 2*(lane&3) + {0,1}]
__device__ __forceinline__ void tile_gemm(const double *__restrict__ As, const double *__restrict__ Bs, int kmax,
                                          double (&acc)[2][4][2]) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, g = lane >> 2, q = lane & 3;
  const double *ap = As + (16 * (w >> 1) + g) * GLD + q;
  const double *bp = Bs + (32 * (w & 1) + g) * GLD + q;
#pragma unroll
  for (int mi = 0; mi < 2; mi++)
#pragma unroll
    for (int ni = 0; ni < 4; ni++) acc[mi][ni][0] = acc[mi][ni][1] = 0.0;
#pragma unroll 4
  for (int k4 = 0; k4 < kmax; k4 += 4) {  // operands are zero-padded beyond kmax
    double a[2], b[4];
#pragma unroll
    for (int mi = 0; mi < 2; mi++) a[mi] = ap[mi * 8 * GLD + k4];
#pragma unroll
    for (int ni = 0; ni < 4; ni++) b[ni] = bp[ni * 8 * GLD + k4];
#pragma unroll
    for (int mi = 0; mi < 2; mi++)
#pragma unroll
      for (int ni = 0; ni < 4; ni++)
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                     : "+d"(acc[mi][ni][0]), "+d"(acc[mi][ni][1])
                     : "d"(a[mi]), "d"(b[ni]));
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

// Trailing tile A_ij -= L_ik L_jk^T on staged operands (As = L_ik rows, Bs = L_jk rows).  `wait` is called after
// the old values of the target tile have been requested (their L2 latency hides under the product) and must
// make the staged operands visible to the CTA.
template <class Wait>
__device__ __forceinline__ void dense_trail_compute(const DenseBufs &B, int D, int ts, int i, int j, const double *As,
                                                    const double *Bs, Wait wait) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, g = lane >> 2, q = lane & 3;
  const size_t ld = (size_t)D;
  const int rows_i = min(DT, D - i * DT), rows_j = min(DT, D - j * DT);
  const int r0 = 16 * (w >> 1) + g, c0 = 32 * (w & 1) + 2 * q;
  double *ct = B.A + (size_t)(i * DT + r0) * ld + j * DT + c0;
  double cold[2][4][2], acc[2][4][2];
#pragma unroll
  for (int mi = 0; mi < 2; mi++)
#pragma unroll
    for (int ni = 0; ni < 4; ni++)
#pragma unroll
      for (int e = 0; e < 2; e++)
        cold[mi][ni][e] = (r0 + 8 * mi < rows_i && c0 + 8 * ni + e < rows_j)
                              ? __ldcg(ct + (size_t)(8 * mi) * ld + 8 * ni + e) : 0.0;
  wait();
  tile_gemm(As, Bs, ts, acc);
#pragma unroll
  for (int mi = 0; mi < 2; mi++)
#pragma unroll
    for (int ni = 0; ni < 4; ni++)
#pragma unroll
      for (int e = 0; e < 2; e++)
        if (r0 + 8 * mi < rows_i && c0 + 8 * ni + e < rows_j)
          ct[(size_t)(8 * mi) * ld + 8 * ni + e] = cold[mi][ni][e] - acc[mi][ni][e];
  __syncthreads();
}

__global__ void __launch_bounds__(DENSE_THREADS, 2)
dense_solve_retract_kernel(float *Twc, DenseBufs B, int N, int D, float delta_thresh, float *dx, State *state,
                           int gated) {
  // both flags are written only by earlier launches: uniform over the grid
  if (state->done || (gated && !state->pcg_bad)) return;
  extern __shared__ __align__(16) double dsm[];
  double *As = dsm, *Bs = dsm + DT * GLD, *v0 = dsm + 2 * DT * GLD;
  __shared__ int fail;
  __shared__ double fin_scratch[64];
  const int tid = threadIdx.x, G = gridDim.x, cta = blockIdx.x;
  const int nt = (D + DT - 1) / DT;
  const size_t ld = (size_t)D;
  unsigned gen = 0;
  if (tid == 0) gen = ld_acquire_u32(&B.bar[1]);
  bool ok = true;
  double acc[2][4][2];
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
  // Workers of phase (3b): every CTA that does NOT share an SM with CTA 0 — a co-resident CTA running DMMAs
  // slowed the serial diagonal factorisation from ~61 k to ~114 k cycles per tile.  SM ids are exchanged
  // through the head of the (not yet used) solution vector.
  __shared__ int s_wc, s_WG;
  {
    unsigned smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    int *smids = reinterpret_cast<int *>(B.x);
    if (tid == 0) smids[cta] = (int)smid;
  }
  if (cta == 0) {
    c_t = clock64();
    dense_factor_diag(B, 0, D, As, Bs, v0, &fail);
    c_diag += clock64() - c_t;
  }
  grid_barrier(B.bar, gen);
  if (tid == 0) {
    const int *smids = reinterpret_cast<const int *>(B.x);
    const int sm0 = __ldcg(smids), mine = __ldcg(smids + cta);
    int wc = 0, WG = 0;
    for (int c = 1; c < G; c++) {
      const bool worker = __ldcg(smids + c) != sm0;
      if (worker && c < cta) wc++;
      if (worker) WG++;
    }
    s_WG = WG;
    s_wc = (cta > 0 && mine != sm0) ? wc : -1;  // -1: idles in (3b)
    if (G == 1 || WG == 0) { s_WG = 1; s_wc = cta == 0 ? 0 : -1; }  // CTA 0 does everything itself
  }
  __syncthreads();
  const int wc3b = s_wc, WG3b = s_WG;
  const bool diag_also_works = (G == 1) || (WG3b == 1 && wc3b == 0 && cta == 0);
  if (__ldcg(B.flag)) ok = false;
  for (int kk = 0; ok && kk < nt; kk++) {
    const int k0 = kk * DT, ts = min(DT, D - k0);
    // ---- (2) panel tiles L_ik = A_ik W^T, b_i -= L_ik y_kk ----------------------------------------------
    c_t = clock64();
    for (int i = kk + 1 + cta; i < nt; i += G) {
      const int rows = min(DT, D - i * DT);
      stage_tile(As, B.A + (size_t)(i * DT) * ld + k0, ld, rows, ts);
      stage_tile(Bs, B.W + (size_t)kk * DT * DT, DT, DT, DT);
      if (tid < DT) v0[tid] = tid < ts ? __ldcg(B.b + k0 + tid) : 0.0;
      __syncthreads();
      tile_gemm(As, Bs, ts, acc);
      __syncthreads();
      {
        const int lane = tid & 31, w = tid >> 5, r0 = 16 * (w >> 1) + (lane >> 2), c0 = 32 * (w & 1) + 2 * (lane & 3);
#pragma unroll
        for (int mi = 0; mi < 2; mi++)
#pragma unroll
          for (int ni = 0; ni < 4; ni++)
#pragma unroll
            for (int e = 0; e < 2; e++) {
              const int r = r0 + 8 * mi, cc = c0 + 8 * ni + e;
              As[r * GLD + cc] = acc[mi][ni][e];
              if (r < rows && cc < ts) B.A[(size_t)(i * DT + r) * ld + k0 + cc] = acc[mi][ni][e];
            }
      }
      __syncthreads();
      if (tid < rows) {
        double s = 0.0;
        for (int c = 0; c < ts; c++) s += As[tid * GLD + c] * v0[c];
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
    for (int i = kk + 1 + cta; i < nt; i += G) {
      stage_tile(As, B.A + (size_t)(i * DT) * ld + k0, ld, min(DT, D - i * DT), ts);
      stage_tile(Bs, B.A + (size_t)((kk + 1) * DT) * ld + k0, ld, min(DT, D - (kk + 1) * DT), ts);
      dense_trail_compute(B, D, ts, i, kk + 1, As, Bs, [] { __syncthreads(); });
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
    if (wc3b >= 0 && (cta > 0 || diag_also_works)) {
      const int WG = WG3b, wc = wc3b, m2 = m - 1;
      const long long total = (long long)m2 * (m2 + 1) / 2;
      const long long per = (total + WG - 1) / WG, t_begin = per * wc, t_end = min(total, t_begin + per);
      // consecutive items share the row tile i: L_ik stays staged in As.  (Streaming the next L_jk with 8-byte
      // cp.async into a second buffer was measured 25 % SLOWER: 4096 tiny async copies per tile; rows of an
      // odd-D matrix are not 16-byte aligned.)
      if (t_begin < t_end) {
        int ip = (int)((sqrt(8.0 * (double)t_begin + 1.0) - 1.0) * 0.5);
        while ((long long)(ip + 1) * (ip + 2) / 2 <= t_begin) ip++;
        while ((long long)ip * (ip + 1) / 2 > t_begin) ip--;
        int jp = (int)(t_begin - (long long)ip * (ip + 1) / 2);
        int staged_i = -1;
        for (long long t = t_begin; t < t_end; t++) {
          const int i = kk + 2 + ip, j = kk + 2 + jp;
          if (i != staged_i) {
            stage_tile(As, B.A + (size_t)(i * DT) * ld + k0, ld, min(DT, D - i * DT), ts);
            staged_i = i;
          }
          stage_tile(Bs, B.A + (size_t)(j * DT) * ld + k0, ld, min(DT, D - j * DT), ts);
          dense_trail_compute(B, D, ts, i, j, As, Bs, [] { __syncthreads(); });
          if (jp < ip) jp++;
          else { ip++; jp = 0; }
        }
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
  if (cta == 0) {
    if (tid == 0) state->dense_runs += 1;
    finish_update(ok, B.x, fin_scratch, D, N, Twc, dx, delta_thresh, state);
  }
}

}  // namespace b200gn
