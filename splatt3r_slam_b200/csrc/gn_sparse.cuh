// b200slam — stage 3 for LARGE, SPARSE graphs: numeric block-sparse Cholesky on the device, executing the
// static plan of gn_sparse_plan.h (ordering + symbolic factorisation done once per topology on the host).
//
// What it replaces: the numeric half of Eigen's SimplicialLLT in SparseBlock::solve (gn_kernels.cu:131-152).
// When it runs: graphs whose factor stays sparse under a minimum-degree ordering — odometry chains with few loop
// closures, the typical SLAM graph — which are exactly the graphs where the pipelined PCG (gn_pcg2.cuh) needs
// ~N iterations.  Graphs that fill in (random chords: configs[3]) keep the PCG.  An exact solve: a non-positive
// pivot sets state->pcg_bad and the gated dense Cholesky behind reproduces the reference's "dx = 0" rule.
//
// Shape: ONE CTA of 1024 threads.  The work is a few thousand 7x7 block products along an elimination tree of
// 10-50 levels; what matters is the latency of a level, and a block barrier costs ~50 ns where a grid barrier
// costs 1-2 us.  Phase 0 gathers all blocks of H from the edge blocks (one warp per block, fixed edge order);
// then per tree level: left-looking updates L_ij -= L_ik L_jk^T from the finished descendant columns, 7x7 Cholesky
// of the diagonal block (its inverse W is kept), forward substitution of the column's right-hand side, and
// L_ij = A_ij W^T — three block-synchronised phases so that levels with few, heavy columns still use all warps;
// the backward substitution walks the levels in reverse.
// All sums run in plan order => deterministic, bit-identical on every rank.
#pragma once

namespace b200gn {

constexpr int SP_THREADS = 1024;
constexpr int SP_WARPS = SP_THREADS / 32;

struct SparseDev {  // device view of b200sparse::Plan + value arrays
  int M, nL, levels;
  const int *perm, *colptr, *rowidx, *level_ptr, *level_cols, *upd_ptr, *upd, *row_ptr, *row_ent, *asm_ptr, *asm_ent,
      *rhs_ptr, *rhs_ent;
  double *Dg;  // (M,49) diagonal blocks -> W = C^-1
  double *Lv;  // (nL,49) off-diagonal blocks of L
  double *y;   // (M,7)
  double *xs;  // (M,7) solution in elimination order
  double *x;   // (7M) solution in pose order
};

// C C^T = A (7x7 SPD, row-major in A); writes W = C^-1 (lower triangular, row-major, zeros above) over A.
// Serial (one lane); C lives in the caller's shared scratch (49 doubles), W is formed in place of A.
__device__ inline bool chol7_winv(double *A, double *scratch) {
  double *C = scratch, *W = A;
  for (int t = 0; t < 49; t++) scratch[t] = 0.0;
  for (int j = 0; j < 7; j++) {
    double s = A[j * 7 + j];
    for (int k = 0; k < j; k++) s -= C[j * 7 + k] * C[j * 7 + k];
    if (!(s > 0.0)) return false;
    const double d = sqrt(s), dinv = 1.0 / d;
    C[j * 7 + j] = d;
    for (int i = j + 1; i < 7; i++) {
      double t = A[i * 7 + j];
      for (int k = 0; k < j; k++) t -= C[i * 7 + k] * C[j * 7 + k];
      C[i * 7 + j] = t * dinv;
    }
  }
  for (int t = 0; t < 49; t++) W[t] = 0.0;  // A is consumed: C holds the factor
  for (int j = 0; j < 7; j++) {
    W[j * 7 + j] = 1.0 / C[j * 7 + j];
    for (int i = j + 1; i < 7; i++) {
      double t = 0.0;
      for (int k = j; k < i; k++) t -= C[i * 7 + k] * W[k * 7 + j];
      W[i * 7 + j] = t / C[i * 7 + i];
    }
  }
  return true;
}

__global__ void __launch_bounds__(SP_THREADS, 1)
sparse_solve_retract_kernel(float *Twc, const double *__restrict__ eb, int N, SparseDev P, float delta_thresh, float *dx,
                            State *state) {
  // all three flags are written only by earlier launches / copies on the stream
  if (state->done || state->pcg_bad || state->solver_sel != 1) return;
  __shared__ double blkA[SP_WARPS][49], blkB[SP_WARPS][49];
  __shared__ double fin_scratch[64];
  __shared__ int s_fail;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int M = P.M;
  if (tid == 0) { s_fail = 0; state->t[0] = gtimer(); }
  // ---- phase 0: gather H (lower block triangle) and b from the edge blocks --------------------------------------
  for (int b = warp; b < M + P.nL; b += SP_WARPS) {
    double a0 = 0.0, a1 = 0.0;
    for (int t = P.asm_ptr[b]; t < P.asm_ptr[b + 1]; t++) {
      const int code = P.asm_ent[t];
      const double *B = eb + (size_t)(code >> 1) * 56;
      const double v0 = __ldcg(B + lane), v1 = lane + 32 < 49 ? __ldcg(B + lane + 32) : 0.0;
      a0 += (code & 1) ? -v0 : v0;
      a1 += (code & 1) ? -v1 : v1;
    }
    double *dst = b < M ? P.Dg + (size_t)b * 49 : P.Lv + (size_t)(b - M) * 49;
    dst[lane] = a0;
    if (lane + 32 < 49) dst[lane + 32] = a1;
  }
  for (int k = warp; k < M; k += SP_WARPS) {
    double a = 0.0;
    if (lane < 7)
      for (int t = P.rhs_ptr[k]; t < P.rhs_ptr[k + 1]; t++) {
        const int code = P.rhs_ent[t];
        const double g = __ldcg(eb + (size_t)(code >> 1) * 56 + 49 + lane);
        a += (code & 1) ? g : -g;
      }
    if (lane < 7) P.y[(size_t)k * 7 + lane] = a;
  }
  __syncthreads();
  if (tid == 0) state->t[1] = gtimer();
  // ---- factorisation + forward substitution, level by level --------------------------------------------------
  // Each level runs in three block-synchronised phases so that the levels near the root — few columns, each with
  // hundreds of update triples and dozens of blocks — still use all 32 warps:
  //   A  left-looking updates: g = 32 / (columns of the level) warps share a column; warp `sub` of the group takes
  //      the triples whose DESTINATION block is congruent to sub mod g (no two warps touch one block, and every
  //      block still receives its updates in plan order);
  //   B  one warp per column: 7x7 Cholesky of the diagonal block (W = C^-1 kept), forward substitution;
  //   C  L_ij = A_ij W^T, the column's blocks dealt to the same warp groups.
  const int r0 = lane / 7, c0 = lane % 7, r1 = (lane + 32) / 7, c1 = (lane + 32) % 7;
  for (int l = 0; l < P.levels; l++) {
    const int lbeg = P.level_ptr[l], ncol = P.level_ptr[l + 1] - lbeg;
    const int g = ncol >= SP_WARPS ? 1 : SP_WARPS / ncol;          // warps per column
    const int stride = ncol >= SP_WARPS ? SP_WARPS : ncol;          // columns in flight per round
    const int wcol = g == 1 ? warp : warp / g, sub = g == 1 ? 0 : warp % g;
    // ---- A: updates ------------------------------------------------------------------------------------------
    if (wcol < stride) {
      for (int c = lbeg + wcol; c < lbeg + ncol; c += stride) {
        const int j = P.level_cols[c];
        double *D = P.Dg + (size_t)j * 49;
        const int sbase = P.colptr[j];
        // The triples of a column are grouped by destination block (plan): the warp reads their indices 32 at a time
        // (one coalesced round trip instead of one per triple), accumulates the products of one destination in
        // registers and subtracts them from the block ONCE (the first version paid three dependent L2 round trips per
        // triple: indices, source blocks, read-modify-write of the destination).
        const int t_end = P.upd_ptr[j + 1];
        double acc0 = 0.0, acc1 = 0.0;
        int cur_ds = -2;  // -2: nothing accumulated yet (-1 is the diagonal block)
        bool cur_mine = false;
        auto flush = [&]() {
          if (cur_ds != -2 && cur_mine) {
            double *dst = cur_ds < 0 ? D : P.Lv + (size_t)cur_ds * 49;
            dst[lane] -= acc0;
            if (lane + 32 < 49) dst[lane + 32] -= acc1;
          }
        };
        for (int tb = P.upd_ptr[j]; tb < t_end; tb += 32) {
          const int tt = tb + lane;
          int ia = 0, ib = 0, ids = -2;
          if (tt < t_end) { ia = P.upd[3 * tt]; ib = P.upd[3 * tt + 1]; ids = P.upd[3 * tt + 2]; }
          const int cnt = min(32, t_end - tb);
          for (int q = 0; q < cnt; q++) {
            const int ds = __shfl_sync(0xffffffffu, ids, q);
            if (ds != cur_ds) {  // warp-uniform
              flush();
              cur_ds = ds;
              cur_mine = g == 1 || ((ds < 0 ? 0 : 1 + ds - sbase) % g) == sub;
              acc0 = acc1 = 0.0;
            }
            if (!cur_mine) continue;
            const double *A = P.Lv + (size_t)__shfl_sync(0xffffffffu, ia, q) * 49;
            const double *Bk = P.Lv + (size_t)__shfl_sync(0xffffffffu, ib, q) * 49;
            blkA[warp][lane] = A[lane];
            blkB[warp][lane] = Bk[lane];
            if (lane + 32 < 49) { blkA[warp][lane + 32] = A[lane + 32]; blkB[warp][lane + 32] = Bk[lane + 32]; }
            __syncwarp();
#pragma unroll
            for (int qq = 0; qq < 7; qq++) {
              acc0 = fma(blkA[warp][r0 * 7 + qq], blkB[warp][c0 * 7 + qq], acc0);
              if (lane + 32 < 49) acc1 = fma(blkA[warp][r1 * 7 + qq], blkB[warp][c1 * 7 + qq], acc1);
            }
            __syncwarp();
          }
        }
        flush();
      }
    }
    __syncthreads();
    // ---- B: diagonal block + forward substitution, one warp per column -------------------------------------
    for (int c = lbeg + warp; c < lbeg + ncol; c += SP_WARPS) {
      const int j = P.level_cols[c];
      double *D = P.Dg + (size_t)j * 49;
      blkA[warp][lane] = D[lane];
      if (lane + 32 < 49) blkA[warp][lane + 32] = D[lane + 32];
      __syncwarp();
      int okw = 1;
      if (lane == 0) okw = chol7_winv(blkA[warp], blkB[warp]) ? 1 : 0;
      okw = __shfl_sync(0xffffffffu, okw, 0);
      if (!okw) { if (lane == 0) s_fail = 1; continue; }  // warp-uniform
      __syncwarp();
      D[lane] = blkA[warp][lane];
      if (lane + 32 < 49) D[lane + 32] = blkA[warp][lane + 32];
      // forward substitution: y_j = W (b_j - sum_k L_jk y_k)   (row j's blocks belong to finished columns)
      double rhs = lane < 7 ? P.y[(size_t)j * 7 + lane] : 0.0;
      for (int t = P.row_ptr[j]; t < P.row_ptr[j + 1]; t++) {
        const int k = P.row_ent[2 * t];
        const double *Lk = P.Lv + (size_t)P.row_ent[2 * t + 1] * 49;
        if (lane < 7) {
#pragma unroll
          for (int q = 0; q < 7; q++) rhs = fma(-Lk[lane * 7 + q], P.y[(size_t)k * 7 + q], rhs);
        }
      }
      double yv = 0.0;
#pragma unroll
      for (int q = 0; q < 7; q++) {
        const double rq = __shfl_sync(0xffffffffu, rhs, q);
        if (lane < 7 && q <= lane) yv = fma(blkA[warp][lane * 7 + q], rq, yv);
      }
      if (lane < 7) P.y[(size_t)j * 7 + lane] = yv;
      __syncwarp();
    }
    __syncthreads();
    if (s_fail) break;  // CTA-uniform after the barrier
    // ---- C: column blocks L_ij = A_ij W^T (W lower triangular), dealt to the warp groups -------------------
    if (wcol < stride) {
      for (int c = lbeg + wcol; c < lbeg + ncol; c += stride) {
        const int j = P.level_cols[c];
        const double *Wj = P.Dg + (size_t)j * 49;
        blkA[warp][lane] = Wj[lane];
        if (lane + 32 < 49) blkA[warp][lane + 32] = Wj[lane + 32];
        for (int s = P.colptr[j] + sub; s < P.colptr[j + 1]; s += g) {
          double *Ls = P.Lv + (size_t)s * 49;
          __syncwarp();
          blkB[warp][lane] = Ls[lane];
          if (lane + 32 < 49) blkB[warp][lane + 32] = Ls[lane + 32];
          __syncwarp();
          double v0 = 0.0, v1 = 0.0;
          for (int q = 0; q <= c0; q++) v0 = fma(blkB[warp][r0 * 7 + q], blkA[warp][c0 * 7 + q], v0);
          if (lane + 32 < 49)
            for (int q = 0; q <= c1; q++) v1 = fma(blkB[warp][r1 * 7 + q], blkA[warp][c1 * 7 + q], v1);
          Ls[lane] = v0;
          if (lane + 32 < 49) Ls[lane + 32] = v1;
        }
        __syncwarp();
      }
    }
    __syncthreads();
  }
  if (tid == 0) state->t[2] = gtimer();
  const bool ok = s_fail == 0;
  // ---- backward substitution: x_j = W^T (y_j - sum_i L_ij^T x_i), levels in reverse ----------------------------
  if (ok) {
    for (int l = P.levels - 1; l >= 0; l--) {
      for (int c = P.level_ptr[l] + warp; c < P.level_ptr[l + 1]; c += SP_WARPS) {
        const int j = P.level_cols[c];
        const double *D = P.Dg + (size_t)j * 49;
        double rhs = lane < 7 ? P.y[(size_t)j * 7 + lane] : 0.0;
        for (int s = P.colptr[j]; s < P.colptr[j + 1]; s++) {
          const double *Ls = P.Lv + (size_t)s * 49, *xi = P.xs + (size_t)P.rowidx[s] * 7;
          if (lane < 7) {
#pragma unroll
            for (int r = 0; r < 7; r++) rhs = fma(-Ls[r * 7 + lane], xi[r], rhs);
          }
        }
        double xv = 0.0;
#pragma unroll
        for (int q = 0; q < 7; q++) {
          const double rq = __shfl_sync(0xffffffffu, rhs, q);
          if (lane < 7 && q >= lane) xv = fma(D[q * 7 + lane], rq, xv);
        }
        if (lane < 7) {
          P.xs[(size_t)j * 7 + lane] = xv;
          P.x[(size_t)P.perm[j] * 7 + lane] = xv;
        }
      }
      __syncthreads();
    }
  }
  if (tid == 0) {
    state->sparse_runs += 1;
    if (!ok) state->pcg_bad = 1;  // not positive definite: the exact dense solver behind reproduces the reference's dx = 0
  }
  if (ok) finish_update(true, P.x, fin_scratch, 7 * M, N, Twc, dx, delta_thresh, state);
  if (tid == 0) state->t[3] = gtimer();
}

}  // namespace b200gn
