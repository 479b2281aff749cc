// b200slam — persistent Gauss-Newton kernel for small graphs (the tracker: N = 2, E = 2).
//
// When all E*n points fit in the shared memory of the 148 SMs (32 B per point; 512x384 pairs:
// E <= 4), the WHOLE max_iter loop is ONE cooperative launch:
//   prologue   every CTA loads its slice of the edge's points ONCE — streams valid/idx/Q/Xj/Cj,
//              gathers the matched Xi/Ci, folds the iteration-independent validity test
//              (match & q > Q_thr & ci > C_thr & cj > C_thr) — into shared memory (SoA, 8 words/point)
//   iteration  relative pose from the current Twc -> 35 accumulators from shared memory -> CTA
//              partial (fp64) -> ticket; the last-arriving CTA runs stages 2-3 (finalize, assemble,
//              fp64 Cholesky, retraction, convergence flag: small_tail) and releases a generation
//              counter the other CTAs spin on (they are co-resident: cooperative launch, 1 CTA/SM).
// Per iteration this replaces a kernel launch + 17.7 MB of L2/HBM reads + dependent gathers by a
// grid barrier + shared-memory reads (measured on B200: see DESIGN.md section 5).
#pragma once

namespace b200gn {

constexpr int PERSIST_WORDS = 8;           // Xi(3) Xj(3) q ind
constexpr int PERSIST_MAX_POINTS = 5632;   // per CTA: 176 KB of dynamic shared memory
constexpr int PERSIST_THREADS = 512;

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned *p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_u32(unsigned *p, unsigned v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

template <int KIND>
__global__ void __launch_bounds__(PERSIST_THREADS, 1)
gn_persistent_kernel(float *Twc, const float *__restrict__ Xs, const float *__restrict__ Cs,
                     const int64_t *__restrict__ ii_row, const int64_t *__restrict__ jj_row,
                     const int64_t *__restrict__ idx, const uint8_t *__restrict__ valid,
                     const float *__restrict__ Q, int n, BuildConsts C, double *partials, State *state,
                     int N, int E, int max_iter, float delta_thresh, float *dx, double *edge_blocks) {
  extern __shared__ __align__(16) float pts_sm[];
  __shared__ float pose_sm[16];
  __shared__ int flag_sm;
  if (KIND == B200SLAM_GN_CALIB) {  // gn_kernels.cu:1273-1278
    C.fx = __ldg(C.Kptr + 0); C.fy = __ldg(C.Kptr + 4); C.cx = __ldg(C.Kptr + 2); C.cy = __ldg(C.Kptr + 5);
  }
  const int e = blockIdx.y, S = gridDim.x, s = blockIdx.x, tid = threadIdx.x;
  const int ix = (int)ii_row[e], jx = (int)jj_row[e];
  const int groups = n >> 2;  // host guarantees n % 4 == 0
  const int gper = (groups + S - 1) / S;
  const int g_begin = s * gper, g_end = min(groups, g_begin + gper);
  const int stride = 4 * gper;  // points per smem array
  const int npts = 4 * max(g_end - g_begin, 0);

  const bool stamper = threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0;  // State::prof[0..2]: start, points staged, loop done
  if (stamper) state->prof[0] = gtimer();
  // ---- prologue: stage this CTA's points ---------------------------------------------------------
  {
    const size_t xi_slot = C.slots ? (size_t)C.slots[ix] : (size_t)ix, xj_slot = C.slots ? (size_t)C.slots[jx] : (size_t)jx;
    const float *Xi_base = Xs + xi_slot * n * 3, *Xj_base = Xs + xj_slot * n * 3;
    const float *Ci_base = Cs + xi_slot * n, *Cj_base = Cs + xj_slot * n;
    const bool second = e >= C.E1;
    const int64_t *idx_e = (second ? C.idx2 : idx) + (size_t)e * n;
    const uint8_t *val_e = (second ? C.valid2 : valid) + (size_t)e * n;
    const float *Q_e = (second ? C.Q2 : Q) + (size_t)e * n;
    // A CTA holds at most PERSIST_MAX_POINTS / 4 = 1408 groups = 2.75 per thread: all (<= 3) groups of a thread
    // are loaded in ONE round of independent streaming loads (24 in flight per thread) followed by one round
    // of gathers, instead of a load -> gather -> store chain per group (6 dependent HBM/L2 round trips).
    constexpr int KMAX = (PERSIST_MAX_POINTS / 4 + PERSIST_THREADS - 1) / PERSIST_THREADS;
    uint32_t vm4[KMAX];
    longlong2 ia[KMAX], ib[KMAX];
    float4 q4[KMAX], c4[KMAX], xa[KMAX], xb[KMAX], xc[KMAX];
#pragma unroll
    for (int k = 0; k < KMAX; k++) {
      const int g = g_begin + tid + k * PERSIST_THREADS;
      if (g < g_end) {
        vm4[k] = __ldg(reinterpret_cast<const uint32_t *>(val_e) + g);
        ia[k] = __ldg(reinterpret_cast<const longlong2 *>(idx_e) + 2 * g);
        ib[k] = __ldg(reinterpret_cast<const longlong2 *>(idx_e) + 2 * g + 1);
        q4[k] = __ldg(reinterpret_cast<const float4 *>(Q_e) + g);
        c4[k] = __ldg(reinterpret_cast<const float4 *>(Cj_base) + g);
        xa[k] = __ldg(reinterpret_cast<const float4 *>(Xj_base) + 3 * g);
        xb[k] = __ldg(reinterpret_cast<const float4 *>(Xj_base) + 3 * g + 1);
        xc[k] = __ldg(reinterpret_cast<const float4 *>(Xj_base) + 3 * g + 2);
      }
    }
#pragma unroll
    for (int k = 0; k < KMAX; k++) {
      const int g = g_begin + tid + k * PERSIST_THREADS;
      if (g >= g_end) continue;
      const long long ind[4] = {ia[k].x, ia[k].y, ib[k].x, ib[k].y};
      const float qq[4] = {q4[k].x, q4[k].y, q4[k].z, q4[k].w};
      const float cj[4] = {c4[k].x, c4[k].y, c4[k].z, c4[k].w};
      float Xi[4][3], ci[4];
#pragma unroll
      for (int t = 0; t < 4; t++) {
        const bool vm = (vm4[k] >> (8 * t)) & 0xffu;
        const long long id = vm ? ind[t] : 0;  // gn_kernels.cu:912-913
        Xi[t][0] = __ldg(Xi_base + 3 * id);
        Xi[t][1] = __ldg(Xi_base + 3 * id + 1);
        Xi[t][2] = __ldg(Xi_base + 3 * id + 2);
        ci[t] = __ldg(Ci_base + id);
      }
      float qs[4];
      int iw[4];
#pragma unroll
      for (int t = 0; t < 4; t++) {
        const bool vm = (vm4[k] >> (8 * t)) & 0xffu;
        const bool ok = vm && (qq[t] > C.Q_thresh) && (ci[t] > C.C_thresh) && (cj[t] > C.C_thresh);
        qs[t] = ok ? qq[t] : -INFINITY;  // fails "q > Q_thresh" in point_contrib => point skipped
        iw[t] = (int)ind[t];
      }
      const int p = 4 * (g - g_begin);
      float4 *a = reinterpret_cast<float4 *>(pts_sm + p);
      const int st4 = stride >> 2;
      a[0 * st4] = make_float4(Xi[0][0], Xi[1][0], Xi[2][0], Xi[3][0]);
      a[1 * st4] = make_float4(Xi[0][1], Xi[1][1], Xi[2][1], Xi[3][1]);
      a[2 * st4] = make_float4(Xi[0][2], Xi[1][2], Xi[2][2], Xi[3][2]);
      a[3 * st4] = make_float4(xa[k].x, xa[k].w, xb[k].z, xc[k].y);
      a[4 * st4] = make_float4(xa[k].y, xb[k].x, xb[k].w, xc[k].z);
      a[5 * st4] = make_float4(xa[k].z, xb[k].y, xc[k].x, xc[k].w);
      a[6 * st4] = make_float4(qs[0], qs[1], qs[2], qs[3]);
      a[7 * st4] = make_float4(__int_as_float(iw[0]), __int_as_float(iw[1]), __int_as_float(iw[2]),
                               __int_as_float(iw[3]));
    }
  }
  __syncthreads();

  // ---- iterations ----------------------------------------------------------------------------------
  const unsigned total = gridDim.x * gridDim.y;
  const int lane = tid & 31, warp = tid >> 5;
  __shared__ float red[PERSIST_THREADS / 32][NACC];
  __shared__ bool is_last;
  if (tid < 16) pose_sm[tid] = __ldcg(Twc + 8 * (tid < 8 ? ix : jx) + (tid & 7));
  __syncthreads();
  if (stamper) state->prof[1] = gtimer();
  for (int it = 0; it < max_iter; it++) {
    const EdgePose P = rel_pose(pose_sm, pose_sm + 8);

    float acc[NACC];
#pragma unroll
    for (int l = 0; l < NACC; l++) acc[l] = 0.0f;
    for (int p = tid; p < npts; p += PERSIST_THREADS) {
      const float q = pts_sm[6 * stride + p];
      if (!(q > C.Q_thresh)) continue;
      const float Xi[3] = {pts_sm[p], pts_sm[stride + p], pts_sm[2 * stride + p]};
      const float Xj[3] = {pts_sm[3 * stride + p], pts_sm[4 * stride + p], pts_sm[5 * stride + p]};
      const long long ind = (long long)__float_as_int(pts_sm[7 * stride + p]);
      point_contrib<KIND>(acc, P, C, Xi, INFINITY, Xj, INFINITY, q, ind);
    }
    // warp tree (fp32) -> cross-warp sum (fp64, fixed order) -> per-CTA partial
#pragma unroll
    for (int l = 0; l < NACC; l++) {
      float v = acc[l];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) red[warp][l] = v;
    }
    __syncthreads();
    if (tid < NACC) {
      double sum = 0.0;
#pragma unroll
      for (int wq = 0; wq < PERSIST_THREADS / 32; wq++) sum += (double)red[wq][tid];
      partials[((size_t)e * S + s) * NACC + tid] = sum;
    }
    __syncthreads();
    if (tid == 0) {
      cuda::atomic_ref<unsigned, cuda::thread_scope_device> ticket(state->ticket);
      is_last = ticket.fetch_add(1u, cuda::memory_order_acq_rel) == total - 1;
    }
    __syncthreads();
    if (is_last) {
      if (tid == 0) { state->ticket = 0; state->t[0] = gtimer(); }
      small_tail<PERSIST_THREADS>(Twc, ii_row, jj_row, partials, S, E, N, delta_thresh, dx, edge_blocks, state);
      __syncthreads();
      if (tid == 0) {
        __threadfence();
        st_release_u32(&state->gen, (unsigned)(it + 1));  // opens the grid barrier
      }
    } else if (tid == 0) {
      while (ld_acquire_u32(&state->gen) < (unsigned)(it + 1)) __nanosleep(20);
    }
    __syncthreads();
    // one L2 round trip for everything the next iteration needs: the two poses of this CTA's edge (the tail
    // on another SM wrote them) and the convergence flag
    if (tid < 16) pose_sm[tid] = __ldcg(Twc + 8 * (tid < 8 ? ix : jx) + (tid & 7));
    if (tid == 16) flag_sm = __ldcg(&state->done);
    __syncthreads();
    if (flag_sm) break;  // ||dx|| < delta: same early exit on every CTA
  }
  if (stamper) state->prof[2] = gtimer();
}

}  // namespace b200gn
