// b200slam — stage 3 for LARGE graphs (7(N-1) > 217), default solver: PIPELINED block-Jacobi
// preconditioned conjugate gradients (Ghysels & Vanroose 2014) in fp64 on the block-sparse normal
// equations, ONE cooperative kernel per GN iteration with ONE grid-wide synchronisation per CG
// iteration.
//
// What it replaces: SparseBlock::solve (gn_kernels.cu:131-152, host-side Eigen SimplicialLLT on the
// assembled 7(N-1) system).  The system is the same one (fixed pose dropped, duplicate blocks summed,
// :70-112) but is never assembled: H x restricted to pose r is  D_r x_r - sum_{t in adj(r)} B_t x_o(t)
// with B_t the per-edge 7x7 block the build produced and D_r = sum_t B_t.
//
// Why this shape.  The system is tiny (config 4: 1023 poses, 16 k incident blocks = 6.4 MB) and CG on it
// is pure latency: the first version (gn_pcg.cuh) spent 3 counter barriers + 3 serial 148-term sums per
// iteration (~45 us per iteration, 4.9 ms per solve).  Here:
//   * one warp owns a pose; its 8 CG vectors, D_r, D_r^-1 and its incident blocks live in shared memory
//     for the whole solve, so an iteration touches global memory only for the neighbours' 7-vectors;
//   * the pipelined recurrences need the two dot products (r,u), (w,u) and the neighbours' m = D^-1 w at
//     the SAME point, so one exchange per iteration carries both: every CTA writes its m rows and its
//     two partial sums, then releases a per-CTA sequence tag; every warp acquires all tags (that IS the
//     grid barrier: no counter, no second round trip) and sums the partials in fixed order, so all CTAs
//     compute bit-identical alpha/beta and leave the loop together;
//   * buffers are double-buffered by sequence parity (a CTA can run at most one exchange ahead).
// Pipelined CG loses a few digits of attainable accuracy to its longer recurrences, so the kernel ends
// with the TRUE residual b - H x (two more exchanges); unless that is below 1e-9 relative (and the
// recursive one below 1e-10) within max_cg iterations, or if a curvature is not positive, it leaves the
// poses alone and sets state->pcg_bad: the exact tiled Cholesky (gn_dense.cuh), enqueued right behind,
// then runs instead of being skipped — for this and all later iterations of the call.  Ring-like graphs
// (few loop closures: CG needs ~N iterations) therefore cost one capped PCG attempt per call.
// Deterministic: adjacency in edge order, fixed-order reductions, no atomics on data.
#pragma once

namespace b200gn {

constexpr int P2_THREADS = 512;
constexpr int P2_WARPS = P2_THREADS / 32;
constexpr int P2_MAX_CTAS = 160;
constexpr int P2_MAX_P = 8;                            // poses per warp
constexpr int P2_POSE_DOUBLES = 49 + 49 + 7 + 8 * 7;  // D, Dinv, b, 8 vectors
constexpr int P2_SMEM_BYTES = 190 * 1024;              // dynamic shared memory opt-in (+ ~30 KB static <= 227 KB)

struct Pcg2Bufs {
  const int *adj_ptr;   // (M+1)
  const int *adj_list;  // (e << 1) | side
  double *xch;          // [2][D] exchanged vector (u0, m, x), by sequence parity
  double *part;         // [2][P2_MAX_CTAS][2] partial sums, by sequence parity
  unsigned *tags;       // [2][P2_MAX_CTAS] sequence tags (zeroed before the launch)
  int max_cg;
  double tol2, true_tol2;
};

// vectors of a pose in shared memory
enum { V_X = 0, V_R, V_U, V_W, V_P, V_S, V_Q, V_Z };

struct P2Pose {  // shared-memory view of one pose
  double *D, *Dinv, *b, *v;
};
__device__ __forceinline__ P2Pose p2_pose(double *base, int slot) {
  double *p = base + (size_t)slot * P2_POSE_DOUBLES;
  return P2Pose{p, p + 49, p + 98, p + 105};
}

// Publish: all warps have written their xch rows / wred entries.  One CTA barrier, then warp 0 folds the
// 16 per-warp partials (fixed xor tree) and lane 0 stores them and releases the tag.
__device__ __forceinline__ void p2_publish(const Pcg2Bufs &B, unsigned seq, double (*wred)[2]) {
  __syncthreads();
  if (threadIdx.x < 32) {
    const int lane = threadIdx.x;
    double g = lane < P2_WARPS ? wred[lane][0] : 0.0, d = lane < P2_WARPS ? wred[lane][1] : 0.0;
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) {
      g += __shfl_xor_sync(0xffffffffu, g, o);
      d += __shfl_xor_sync(0xffffffffu, d, o);
    }
    if (lane == 0) {
      const int par = seq & 1;
      double *p = B.part + ((size_t)par * P2_MAX_CTAS + blockIdx.x) * 2;
      __stcg(p, g);
      __stcg(p + 1, d);
      st_release_u32(B.tags + par * P2_MAX_CTAS + blockIdx.x, seq);  // cumulative: covers the CTA's xch rows
    }
  }
  __syncthreads();  // wred may be rewritten
}

// Wait until every CTA has published `seq`, then return the two grid sums (identical bits on every CTA: lane c
// sums slots c, c+32, ... in order, then a fixed xor tree).  ONE warp polls and the rest of the CTA waits at a
// block barrier: tools/micro/xchg_latency.cu measures 2.4 k cycles per exchange at 64 CTAs this way against
// 8.0 k with all 16 warps polling (their loads queue up in front of the tag stores at the L2 slice).
// Polling uses RELAXED gpu-scope loads and NO acquire fence: ld.acquire.gpu / fence.acq_rel.gpu compile to
// CCTL.IVALL (invalidate-all of the SM's L1), which is not needed here — every cross-CTA value read afterwards
// (partial sums, neighbours' vectors, x) is loaded with ld.global.cg, i.e. from L2, the point of coherence, and
// is issued only after the tag load has returned (control dependency); the writer's st.release.gpu made its data
// visible at L2 before the tag.
__device__ __forceinline__ void p2_wait(const Pcg2Bufs &B, unsigned seq, double &g, double &d, double *s_sum) {
  const int lane = threadIdx.x & 31, par = seq & 1, G = gridDim.x;
  if (threadIdx.x < 32) {
    double sg = 0.0, sd = 0.0;
    for (int c = lane; c < G; c += 32) {
      const unsigned *t = B.tags + par * P2_MAX_CTAS + c;
      unsigned v;
      do {
        asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(t) : "memory");
      } while (v < seq);
    }
    __syncwarp();
    for (int c = lane; c < G; c += 32) {
      const double *p = B.part + ((size_t)par * P2_MAX_CTAS + c) * 2;
      sg += __ldcg(p);
      sd += __ldcg(p + 1);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      sg += __shfl_xor_sync(0xffffffffu, sg, o);
      sd += __shfl_xor_sync(0xffffffffu, sd, o);
    }
    if (lane == 0) { s_sum[0] = sg; s_sum[1] = sd; }
  }
  __syncthreads();
  g = s_sum[0];
  d = s_sum[1];
}

// (H v)_k of one pose for lane k < 7 (other lanes: garbage), v_r = own 7-vector, neighbours' vectors read from the
// exchange buffer xv (L2).  Two phases per chunk of up to 32 incident blocks: (1) the warp GATHERS the chunk's
// neighbour entries (7 per block) into its shared-memory buffer nb, lane l taking entries l, l+32, ... — up to 7
// independent L2 loads per lane in flight, ONE round trip (the first version had each lane load the 7 values of
// its own blocks in sequence with the FMAs: 7 k cycles per pass); (2) 28 lanes = 7 rows x 4 slices of the chunk
// multiply from shared memory; slices combined in fixed order.
constexpr int P2_NB_BLOCKS = 32;
__device__ __forceinline__ double p2_matvec(const P2Pose &S, const double *vr, const double *__restrict__ xv,
                                            const double *blocks, const int *others, int cnt, int lane, double *nb) {
  const int k = lane % 7, slice = lane / 7;
  double ac[7];  // one accumulator per column: 7 independent DFMA chains instead of one 7x longer
#pragma unroll
  for (int c = 0; c < 7; c++) ac[c] = 0.0;
  for (int base = 0; base < cnt; base += P2_NB_BLOCKS) {
    const int m = min(P2_NB_BLOCKS, cnt - base), total = 7 * m;
    double val[7];
#pragma unroll
    for (int u = 0; u < 7; u++) {
      const int idx = lane + 32 * u;
      val[u] = 0.0;
      if (idx < total) {
        const int t = idx / 7, c = idx - 7 * t, o = others[base + t];
        if (o >= 0) val[u] = __ldcg(xv + (size_t)o * 7 + c);
      }
    }
#pragma unroll
    for (int u = 0; u < 7; u++)
      if (lane + 32 * u < total) nb[lane + 32 * u] = val[u];
    __syncwarp();
    if (slice < 4) {
      for (int t = slice; t < m; t += 4) {
        const double *blk = blocks + (size_t)(base + t) * 49 + k * 7;
#pragma unroll
        for (int c = 0; c < 7; c++) ac[c] = fma(blk[c], nb[t * 7 + c], ac[c]);
      }
    }
    __syncwarp();
  }
  const double a = ((ac[0] + ac[1]) + (ac[2] + ac[3])) + ((ac[4] + ac[5]) + ac[6]);
  const double a1 = __shfl_sync(0xffffffffu, a, (lane + 7) & 31);
  const double a2 = __shfl_sync(0xffffffffu, a, (lane + 14) & 31);
  const double a3 = __shfl_sync(0xffffffffu, a, (lane + 21) & 31);
  double dv = 0.0;
  if (lane < 7) {
    double dp[7];
#pragma unroll
    for (int c = 0; c < 7; c++) dp[c] = S.D[k * 7 + c] * vr[c];
    dv = ((dp[0] + dp[1]) + (dp[2] + dp[3])) + ((dp[4] + dp[5]) + dp[6]);
  }
  return dv - (((a + a1) + a2) + a3);
}

__global__ void __launch_bounds__(P2_THREADS, 1)
pcg2_solve_retract_kernel(float *Twc, const int64_t *__restrict__ ii_row, const int64_t *__restrict__ jj_row,
                          const double *__restrict__ eb, int N, int E, int P, int blocks_cap, Pcg2Bufs B,
                          float delta_thresh, float *dx, State *state) {
  // written only by earlier launches / copies on the stream: uniform over the grid
  if (state->done || state->pcg_bad || state->solver_sel != 0) return;
  extern __shared__ __align__(16) double p2sm[];
  __shared__ double wred[P2_WARPS][2];
  __shared__ double fin_scratch[64];
  __shared__ int s_off[P2_WARPS * P2_MAX_P + 1];
  __shared__ int s_cnt[P2_WARPS * P2_MAX_P];  // distinct free neighbours per pose after merging
  __shared__ int s_setup_fail;
  __shared__ double s_sum[2];
  __shared__ double s_nb[P2_WARPS][P2_NB_BLOCKS * 7];
  double *nb = s_nb[threadIdx.x >> 5];
  const int M = N - NUM_FIX, D = 7 * M;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (blockIdx.x == 0 && tid == 0) state->t[0] = gtimer();
  const int G = gridDim.x, nW = G * P2_WARPS, wg = blockIdx.x * P2_WARPS + warp;
  double *pose_base = p2sm;                                          // [P2_WARPS * P] poses
  double *blk_base = p2sm + (size_t)P2_WARPS * P * P2_POSE_DOUBLES;  // staged incident blocks
  int *oth_base = reinterpret_cast<int *>(blk_base + (size_t)blocks_cap * 49);

  // ---- setup -----------------------------------------------------------------------------------------
  // offsets of every pose's incident blocks in the CTA's staging area (serial prefix, <= 128 poses)
  if (tid == 0) {
    int s = 0;
    for (int w = 0; w < P2_WARPS; w++)
      for (int pi = 0; pi < P; pi++) {
        const int pose = blockIdx.x * P2_WARPS + w + pi * nW;
        s_off[w * P + pi] = s;
        if (pose < M) s += B.adj_ptr[pose + 1] - B.adj_ptr[pose];
      }
    s_off[P2_WARPS * P] = s;
    s_setup_fail = 0;
  }
  __syncthreads();
  // A CTA whose incident blocks do not fit its shared memory (~480 blocks for 16 poses with 200 KB: graphs
  // with more than ~30 incident directed edges per pose) marks the solve as failed => exact Cholesky.
  const bool staged = s_off[P2_WARPS * P] <= blocks_cap;  // CTA-uniform
  if (!staged && tid == 0) s_setup_fail = 1;

  const int par0 = 1;  // parity of exchange 1
  double acc_g = 0.0, acc_d = 0.0;
  if (staged) {
    for (int pi = 0; pi < P; pi++) {
      const int pose = wg + pi * nW;
      if (pose >= M) continue;  // warp-uniform
      P2Pose S = p2_pose(pose_base, warp * P + pi);
      const int t0 = B.adj_ptr[pose], cnt = B.adj_ptr[pose + 1] - t0;
      double *blocks = blk_base + (size_t)s_off[warp * P + pi] * 49;
      int *others = oth_base + s_off[warp * P + pi];
      double a0 = 0.0, a1 = 0.0;  // entries lane and lane + 32 of [D (49) | b (7)]
      // Incident blocks are MERGED per neighbour while they are staged: the graph lists every edge in both
      // directions (global_opt.py:104-110) and may repeat edges, and all of a pose's blocks towards the same
      // neighbour o multiply the same x_o — summing them (in edge order: deterministic) halves the shared-memory
      // bytes and the DFMAs of every mat-vec.  Blocks towards the fixed pose only enter D_r.
      int m = 0;  // distinct free neighbours so far (warp-uniform)
      for (int t = 0; t < cnt; t++) {
        const int code = B.adj_list[t0 + t], e = code >> 1;
        const double *blk = eb + (size_t)e * 56;
        const double v0 = __ldcg(blk + lane);
        const double v1 = lane + 32 < 56 ? __ldcg(blk + lane + 32) : 0.0;
        a0 += v0;
        if (lane + 32 < 49) a1 += v1;
        else if (lane + 32 < 56) a1 += (code & 1) ? v1 : -v1;  // g_i = -g_j
        const int o = (int)((code & 1) ? ii_row[e] : jj_row[e]) - NUM_FIX;  // < 0: the fixed pose
        if (o < 0) continue;  // warp-uniform
        int j = -1;
        for (int s0 = 0; s0 < m && j < 0; s0 += 32) {
          const unsigned hit = __ballot_sync(0xffffffffu, s0 + lane < m && others[s0 + lane] == o);
          if (hit) j = s0 + __ffs(hit) - 1;
        }
        if (j < 0) {
          j = m++;
          blocks[(size_t)j * 49 + lane] = v0;
          if (lane + 32 < 49) blocks[(size_t)j * 49 + lane + 32] = v1;
          if (lane == 0) others[j] = o;
        } else {
          blocks[(size_t)j * 49 + lane] += v0;
          if (lane + 32 < 49) blocks[(size_t)j * 49 + lane + 32] += v1;
        }
        __syncwarp();
      }
      if (lane == 0) s_cnt[warp * P + pi] = m;
      S.D[lane] = a0;
      if (lane + 32 < 49) S.D[lane + 32] = a1;
      else if (lane + 32 < 56) S.b[lane + 32 - 49] = a1;
      __syncwarp();
      if (lane == 0) {
        double scratch[98];
        if (!spd7_inverse(S.D, scratch, S.Dinv)) s_setup_fail = 1;
      }
      __syncwarp();
      if (lane < 7) {
        double u = 0.0;
#pragma unroll
        for (int c = 0; c < 7; c++) u = fma(S.Dinv[lane * 7 + c], S.b[c], u);
        S.v[V_X * 7 + lane] = 0.0;
        S.v[V_R * 7 + lane] = S.b[lane];
        S.v[V_U * 7 + lane] = u;
        S.v[V_W * 7 + lane] = 0.0;
        S.v[V_P * 7 + lane] = 0.0;
        S.v[V_S * 7 + lane] = 0.0;
        S.v[V_Q * 7 + lane] = 0.0;
        S.v[V_Z * 7 + lane] = 0.0;
        __stcg(B.xch + (size_t)par0 * D + (size_t)pose * 7 + lane, u);  // exchange u0
      }
    }
  }
  __syncwarp();
  // exchange 1: u0 everywhere (+ the setup-failure flag as a "partial": any non-zero => fail)
  if (lane == 0) { wred[warp][0] = 0.0; wred[warp][1] = 0.0; }
  __syncthreads();
  if (tid == 0) wred[0][0] = s_setup_fail ? 1.0 : 0.0;
  unsigned seq = 1;
  p2_publish(B, seq, wred);
  double gsum, dsum;
  p2_wait(B, seq, gsum, dsum, s_sum);
  bool ok = gsum == 0.0;  // uniform over the grid
  if (blockIdx.x == 0 && tid == 0) state->t[1] = gtimer();
  // w0 = H u0
  if (ok) {
    for (int pi = 0; pi < P; pi++) {
      const int pose = wg + pi * nW;
      if (pose >= M) continue;
      P2Pose S = p2_pose(pose_base, warp * P + pi);
      const int cnt = s_cnt[warp * P + pi];
      const double w = p2_matvec(S, S.v + V_U * 7, B.xch + (size_t)par0 * D, blk_base + (size_t)s_off[warp * P + pi] * 49,
                                 oth_base + s_off[warp * P + pi], cnt, lane, nb);
      if (lane < 7) S.v[V_W * 7 + lane] = w;
    }
    __syncwarp();
  }

  // ---- pipelined CG ----------------------------------------------------------------------------------
  int cg_it = 0;
  bool converged = false;
  double gamma0 = 0.0, gamma_old = 1.0, alpha_old = 1.0, gamma_last = 0.0;
  if (ok) {
    long long c0 = clock64(), c_pub = 0, c_wait = 0, c_mat = 0, c_loc = 0, c_sc = 0, c_mv = 0;
    for (int it = 0; it <= B.max_cg; it++) {
      seq++;
      const int par = seq & 1;
      double *xv = B.xch + (size_t)par * D;
      // local: partial (r,u), (w,u); m = Dinv w -> exchange buffer
      acc_g = 0.0;
      acc_d = 0.0;
      double m_reg[P2_MAX_P];
#pragma unroll
      for (int pi = 0; pi < P2_MAX_P; pi++) {
        m_reg[pi] = 0.0;
        if (pi >= P) continue;
        const int pose = wg + pi * nW;
        if (pose >= M) continue;
        P2Pose S = p2_pose(pose_base, warp * P + pi);
        if (lane < 7) {
          const double u = S.v[V_U * 7 + lane];
          acc_g = fma(S.v[V_R * 7 + lane], u, acc_g);
          acc_d = fma(S.v[V_W * 7 + lane], u, acc_d);
          double mp[7];
#pragma unroll
          for (int c = 0; c < 7; c++) mp[c] = S.Dinv[lane * 7 + c] * S.v[V_W * 7 + c];
          const double m = ((mp[0] + mp[1]) + (mp[2] + mp[3])) + ((mp[4] + mp[5]) + mp[6]);
          m_reg[pi] = m;
          __stcg(xv + (size_t)pose * 7 + lane, m);
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        acc_g += __shfl_xor_sync(0xffffffffu, acc_g, o);
        acc_d += __shfl_xor_sync(0xffffffffu, acc_d, o);
      }
      if (lane == 0) { wred[warp][0] = acc_g; wred[warp][1] = acc_d; }
      { const long long c = clock64(); c_loc += c - c0; c0 = c; }
      p2_publish(B, seq, wred);
      { const long long c = clock64(); c_pub += c - c0; c0 = c; }
      double gamma, delta;
      p2_wait(B, seq, gamma, delta, s_sum);
      { const long long c = clock64(); c_wait += c - c0; c0 = c; }
      gamma_last = gamma;
      if (it == 0) gamma0 = gamma;
      if (!(gamma == gamma) || !(delta == delta)) { ok = false; break; }
      if (gamma <= B.tol2 * gamma0) { converged = true; break; }  // covers gamma0 == 0 (b = 0 => x = 0)
      if (it == B.max_cg) break;
      double beta, alpha;
      if (it == 0) {
        beta = 0.0;
        if (!(delta > 0.0)) { ok = false; break; }  // u^T H u <= 0: not positive definite
        alpha = gamma / delta;
      } else {
        beta = gamma / gamma_old;
        const double den = delta - beta * gamma / alpha_old;
        if (!(den > 0.0)) { ok = false; break; }
        alpha = gamma / den;
      }
      gamma_old = gamma;
      alpha_old = alpha;
      cg_it = it + 1;
      { const long long c = clock64(); c_sc += c - c0; c0 = c; }
      // n = H m, then the eight recurrences (all local)
#pragma unroll
      for (int pi = 0; pi < P2_MAX_P; pi++) {
        if (pi >= P) continue;
        const int pose = wg + pi * nW;
        if (pose >= M) continue;
        P2Pose S = p2_pose(pose_base, warp * P + pi);
        const int cnt = s_cnt[warp * P + pi];
        double mr[7];  // own m as a 7-vector
#pragma unroll
        for (int c = 0; c < 7; c++) mr[c] = __shfl_sync(0xffffffffu, m_reg[pi], c);
        const double nk = p2_matvec(S, mr, xv, blk_base + (size_t)s_off[warp * P + pi] * 49,
                                    oth_base + s_off[warp * P + pi], cnt, lane, nb);
        { const long long c = clock64(); c_mv += c - c0; c0 = c; }
        if (lane < 7) {
          double *v = S.v;
          const double z = fma(beta, v[V_Z * 7 + lane], nk);
          const double q = fma(beta, v[V_Q * 7 + lane], m_reg[pi]);
          const double s = fma(beta, v[V_S * 7 + lane], v[V_W * 7 + lane]);
          const double p = fma(beta, v[V_P * 7 + lane], v[V_U * 7 + lane]);
          v[V_Z * 7 + lane] = z;
          v[V_Q * 7 + lane] = q;
          v[V_S * 7 + lane] = s;
          v[V_P * 7 + lane] = p;
          v[V_X * 7 + lane] = fma(alpha, p, v[V_X * 7 + lane]);
          v[V_R * 7 + lane] = fma(-alpha, s, v[V_R * 7 + lane]);
          v[V_U * 7 + lane] = fma(-alpha, q, v[V_U * 7 + lane]);
          v[V_W * 7 + lane] = fma(-alpha, z, v[V_W * 7 + lane]);
        }
      }
      __syncwarp();
      { const long long c = clock64(); c_mat += c - c0; c0 = c; }
    }
    if (tid == 32) {  // extrema over the CTAs of the non-waiting cycles (skew between CTAs shows up as others' wait)
      const unsigned long long busy = (unsigned long long)(c_pub + c_mat + c_loc + c_sc + c_mv);
      atomicMax(&state->prof_busy_max, busy);
      atomicMin(&state->prof_busy_min, busy);
    }
    if (blockIdx.x == 0 && tid == 32) {
      state->prof[0] = c_pub; state->prof[1] = c_wait; state->prof[2] = c_mat;
      state->prof[3] = c_loc; state->prof[4] = c_sc; state->prof[5] = c_mv;
    }
  }

  if (blockIdx.x == 0 && tid == 0) state->t[2] = gtimer();
  // ---- true residual of the returned x, and x for the finishing CTA -------------------------------------
  seq++;
  {
    const unsigned seq_x = seq;
    double *xv = B.xch + (size_t)(seq_x & 1) * D;
    if (staged) {
      for (int pi = 0; pi < P; pi++) {
        const int pose = wg + pi * nW;
        if (pose >= M) continue;
        P2Pose S = p2_pose(pose_base, warp * P + pi);
        if (lane < 7) __stcg(xv + (size_t)pose * 7 + lane, S.v[V_X * 7 + lane]);
      }
    }
    if (lane == 0) { wred[warp][0] = 0.0; wred[warp][1] = 0.0; }
    p2_publish(B, seq, wred);
    double t0, t1;
    p2_wait(B, seq, t0, t1, s_sum);
    acc_g = 0.0;
    if (ok && converged) {
      for (int pi = 0; pi < P; pi++) {
        const int pose = wg + pi * nW;
        if (pose >= M) continue;
        P2Pose S = p2_pose(pose_base, warp * P + pi);
        const int cnt = s_cnt[warp * P + pi];
        const double hx = p2_matvec(S, S.v + V_X * 7, xv, blk_base + (size_t)s_off[warp * P + pi] * 49,
                                    oth_base + s_off[warp * P + pi], cnt, lane, nb);
        const double rt = lane < 7 ? S.b[lane] - hx : 0.0;
        double rv[7];  // (r_true, Dinv r_true)
#pragma unroll
        for (int c = 0; c < 7; c++) rv[c] = __shfl_sync(0xffffffffu, rt, c);
        if (lane < 7) {
          double zt = 0.0;
#pragma unroll
          for (int c = 0; c < 7; c++) zt = fma(S.Dinv[lane * 7 + c], rv[c], zt);
          acc_g = fma(rt, zt, acc_g);
        }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc_g += __shfl_xor_sync(0xffffffffu, acc_g, o);
    if (lane == 0) { wred[warp][0] = acc_g; wred[warp][1] = 0.0; }
    seq++;
    p2_publish(B, seq, wred);
    double gamma_true;
    p2_wait(B, seq, gamma_true, t1, s_sum);
    const bool good = ok && converged && (gamma_true == gamma_true) && gamma_true <= B.true_tol2 * gamma0;
    if (blockIdx.x == 0) {
      if (tid == 0) {
        state->cg_iters = cg_it;
        state->cg_rel = gamma0 > 0.0 ? gamma_last / gamma0 : 0.0;
        state->cg_true_rel = gamma0 > 0.0 ? gamma_true / gamma0 : 0.0;
        state->cg_total += cg_it;
        state->adj_ready = 1;
        if (!good) state->pcg_bad = 1;  // the exact solver behind this launch takes over (this and later iterations)
      }
      if (good) finish_update(true, xv, fin_scratch, D, N, Twc, dx, delta_thresh, state);
      if (tid == 0) state->t[3] = gtimer();
    }
  }
}

}  // namespace b200gn
