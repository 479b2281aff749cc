// b200slam — frame-to-keyframe pose tracking on the device: replaces
// FrameTracker.opt_pose_ray_dist_sim3 + FrameTracker.solve + check_convergence
// (splatt3r_slam/tracker.py:156-214, geometry.py:17-52, nonlinear_optimizer.py:5-33).
//
// The reference runs, per Gauss-Newton step (<= 50 of them), ~30 PyTorch kernels that materialise a
// (4n x 7) Jacobian, an A^T A GEMM, a Cholesky and an `.item()` host sync.  Here the whole loop is
// ONE cooperative launch (same structure as gn_persistent.cuh):
//   prologue   each CTA stages its slice of the points once into shared memory: Xf (matched frame
//              point), rd_k = [Xk/|Xk|, |Xk|] (tracker.py:185) and sqrt(Qk) for valid points, 8 words
//   step       P = T_CkCf * Xf; residual r = rd_k - [P/d, d]; J = -d(rd)/dP [I, -[P]x, P] in closed
//              form ([-(I - rr^T)/d, [r]x, 0] for the ray rows, [-r^T, 0, -d] for the distance row);
//              Huber on the whitened residual (nonlinear_optimizer.py:28-33); 28 + 7 + 1 sums
//              (H upper triangle, A^T b, b^T b) per thread -> warp tree -> CTA partial (fp64) -> ticket
//   tail       the last CTA sums the partials in a fixed order, factors the 7x7 system in fp64
//              (one thread), retracts T <- Exp(tau) T, applies the reference's stopping rule
//              (|cost decrease| / old cost < rel_error, or |tau| < delta_norm; tested AFTER the
//              retraction like tracker.py:199-210) and opens the grid barrier.
// On exit T_WCf = T_WCk * T_CkCf (tracker.py:212).  Deterministic: no atomics on data.
// Deviation (documented in DESIGN.md): points with valid == 0 are skipped, where the reference
// multiplies their rows by zero — identical unless such a row holds NaN/Inf (then the reference's
// Cholesky throws and the frame is dropped).
#pragma once

namespace b200gn {

struct TrackState {      // head of the tracker workspace (256 B)
  float T[8];            // current T_CkCf
  unsigned ticket, gen;
  int iters, converged, chol_fail, done;
  double cost, old_cost;
};
constexpr int TRACK_NACC = 36;
constexpr int TRACK_GROUPS = PERSIST_THREADS / TRACK_NACC;  // 14 groups of CTAs in the partial reduction

__device__ __forceinline__ void sim3_mul(const float *A, const float *B, float *C) {  // C = A * B
  float t[3];
  rotate(A + 3, B, t);
  const float *a = A + 3, *b = B + 3;
  C[0] = A[7] * t[0] + A[0]; C[1] = A[7] * t[1] + A[1]; C[2] = A[7] * t[2] + A[2];
  C[3] = a[3] * b[0] + a[0] * b[3] + a[1] * b[2] - a[2] * b[1];
  C[4] = a[3] * b[1] - a[0] * b[2] + a[1] * b[3] + a[2] * b[0];
  C[5] = a[3] * b[2] + a[0] * b[1] - a[1] * b[0] + a[2] * b[3];
  C[6] = a[3] * b[3] - a[0] * b[0] - a[1] * b[1] - a[2] * b[2];
  C[7] = A[7] * B[7];
}

// KIND 0: ray + distance residual (opt_pose_ray_dist_sim3, tracker.py:174-214); B = Xk (n,3).
// KIND 1: pixel + log-depth residual (opt_pose_calib_sim3, tracker.py:216-266, project_calib
//         geometry.py:63-105); B = meas_k (n,3) = [u_k, v_k, log z_k], valid2 = valid_meas_k.
struct TrackConsts {
  float inv_sigma_a, inv_sigma_b, huber_k, rel_error, delta_norm;
  float ulo, uhi, vlo, vhi, z_eps;  // calib: pixel window (border, W-1-border) x (border, H-1-border)
  const float *K;                   // calib: device (3,3) intrinsics
};

template <int KIND>
__global__ void __launch_bounds__(PERSIST_THREADS, 1)
track_kernel(const float *__restrict__ Xf, const float *__restrict__ B, const float *__restrict__ Qk,
             const uint8_t *__restrict__ valid, const uint8_t *__restrict__ valid2,
             const float *__restrict__ T_WCf_in, const float *__restrict__ T_WCk, float *T_WCf_out,
             float *T_CkCf_out, int n, int gper, TrackConsts C, int max_iters, double *partials,
             TrackState *state) {
  const float huber_k = C.huber_k, rel_error = C.rel_error, delta_norm = C.delta_norm;
  extern __shared__ __align__(16) float pts_sm[];
  __shared__ float pose_sm[8];
  __shared__ float red[PERSIST_THREADS / 32][TRACK_NACC];
  __shared__ bool is_last;
  __shared__ int flag_sm;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, s = blockIdx.x, S = gridDim.x;
  const int p_begin = min(n, s * gper), p_end = min(n, p_begin + gper);
  const int npts = p_end - p_begin, stride = gper;

  // ---- prologue -------------------------------------------------------------------------------------
  for (int p = tid; p < npts; p += PERSIST_THREADS) {
    const size_t g = (size_t)p_begin + p;
    const float xk = __ldg(B + 3 * g), yk = __ldg(B + 3 * g + 1), zk = __ldg(B + 3 * g + 2);
    bool ok = valid[g] != 0;
    pts_sm[p] = __ldg(Xf + 3 * g);
    pts_sm[stride + p] = __ldg(Xf + 3 * g + 1);
    pts_sm[2 * stride + p] = __ldg(Xf + 3 * g + 2);
    if (KIND == 0) {
      const float dk = sqrtf(xk * xk + yk * yk + zk * zk);  // point_to_ray_dist, geometry.py:17-24
      const float dk_inv = 1.0f / dk;
      pts_sm[3 * stride + p] = dk_inv * xk;
      pts_sm[4 * stride + p] = dk_inv * yk;
      pts_sm[5 * stride + p] = dk_inv * zk;
      pts_sm[6 * stride + p] = dk;
    } else {
      pts_sm[3 * stride + p] = xk;
      pts_sm[4 * stride + p] = yk;
      pts_sm[5 * stride + p] = zk;
      ok = ok && valid2[g] != 0;  // tracker.py:236: valid2 = valid_proj & valid_meas_k (the constant half)
    }
    pts_sm[7 * stride + p] = ok ? sqrtf(__ldg(Qk + g)) : 0.0f;  // tracker.py:177-178: valid * sqrt(Qk)
  }
  if (tid == 0) {  // T_CkCf = T_WCk^-1 * T_WCf (tracker.py:182): every CTA computes the same value
    float Ti[8], Tj[8];
    for (int k = 0; k < 8; k++) { Ti[k] = __ldg(T_WCk + k); Tj[k] = __ldg(T_WCf_in + k); }
    const EdgePose P = rel_pose(Ti, Tj);
    pose_sm[0] = P.t[0]; pose_sm[1] = P.t[1]; pose_sm[2] = P.t[2];
    pose_sm[3] = P.q[0]; pose_sm[4] = P.q[1]; pose_sm[5] = P.q[2]; pose_sm[6] = P.q[3];
    pose_sm[7] = P.s;
  }
  __syncthreads();

  float fx = 0.0f, fy = 0.0f, cx = 0.0f, cy = 0.0f;
  if (KIND == 1) { fx = __ldg(C.K + 0); fy = __ldg(C.K + 4); cx = __ldg(C.K + 2); cy = __ldg(C.K + 5); }
  for (int it = 0; it < max_iters; it++) {
    const float t0 = pose_sm[0], t1 = pose_sm[1], t2 = pose_sm[2], sc = pose_sm[7];
    const float q[4] = {pose_sm[3], pose_sm[4], pose_sm[5], pose_sm[6]};
    float acc[NACC], cost = 0.0f;
#pragma unroll
    for (int l = 0; l < NACC; l++) acc[l] = 0.0f;
    for (int p = tid; p < npts; p += PERSIST_THREADS) {
      const float wq = pts_sm[7 * stride + p];
      if (!(wq > 0.0f)) continue;
      const float X[3] = {pts_sm[p], pts_sm[stride + p], pts_sm[2 * stride + p]};
      float R[3];
      rotate(q, X, R);
      const float P0 = fmaf(sc, R[0], t0), P1 = fmaf(sc, R[1], t1), P2 = fmaf(sc, R[2], t2);
      if (KIND == 0) {
        const float d = sqrtf(P0 * P0 + P1 * P1 + P2 * P2), d_inv = 1.0f / d;
        const float r0 = d_inv * P0, r1 = d_inv * P1, r2 = d_inv * P2;
        const float e[4] = {pts_sm[3 * stride + p] - r0, pts_sm[4 * stride + p] - r1, pts_sm[5 * stride + p] - r2,
                            pts_sm[6 * stride + p] - d};
        const float si_ray = C.inv_sigma_a * wq, si_dist = C.inv_sigma_b * wq;
        const float rr[3] = {r0, r1, r2};
#pragma unroll
        for (int k = 0; k < 3; k++) {
          float J[7];
#pragma unroll
          for (int c = 0; c < 3; c++) J[c] = -d_inv * ((c == k ? 1.0f : 0.0f) - rr[k] * rr[c]);
          // [r]x row k
          J[3] = k == 0 ? 0.0f : k == 1 ? r2 : -r1;
          J[4] = k == 0 ? -r2 : k == 1 ? 0.0f : r0;
          J[5] = k == 0 ? r1 : k == 1 ? -r0 : 0.0f;
          J[6] = 0.0f;
          const float wh = fabsf(si_ray * e[k]);
          const float w = si_ray * si_ray * (wh < huber_k ? 1.0f : huber_k / wh);
          accum_row<0x3fu>(acc, J, w, e[k]);
          cost = fmaf(w * e[k], e[k], cost);
        }
        {
          const float J[7] = {-r0, -r1, -r2, 0.0f, 0.0f, 0.0f, -d};
          const float wh = fabsf(si_dist * e[3]);
          const float w = si_dist * si_dist * (wh < huber_k ? 1.0f : huber_k / wh);
          accum_row<0x47u>(acc, J, w, e[3]);
          cost = fmaf(w * e[3], e[3], cost);
        }
      } else {
        // project_calib (geometry.py:63-105) with K = [fx 0 cx; 0 fy cy; 0 0 1]
        const float z_inv = 1.0f / P2;
        const float u = fmaf(fx * P0, z_inv, cx), v = fmaf(fy * P1, z_inv, cy);
        const bool vp = u > C.ulo && u < C.uhi && v > C.vlo && v < C.vhi && P2 > C.z_eps;
        if (!vp) continue;  // sqrt_info2 = valid2 * sqrt_info (tracker.py:237)
        const float e[3] = {pts_sm[3 * stride + p] - u, pts_sm[4 * stride + p] - v, pts_sm[5 * stride + p] - logf(P2)};
        const float si_px = C.inv_sigma_a * wq, si_dep = C.inv_sigma_b * wq;
        // rows a of dpz_dP; J = -a [I, -[P]x, P] = [-a, a x P, -a.P]
        const float a[3][3] = {{fx * z_inv, 0.0f, -fx * P0 * z_inv * z_inv},
                               {0.0f, fy * z_inv, -fy * P1 * z_inv * z_inv},
                               {0.0f, 0.0f, z_inv}};
#pragma unroll
        for (int k = 0; k < 3; k++) {
          const float J[7] = {-a[k][0], -a[k][1], -a[k][2],
                              a[k][1] * P2 - a[k][2] * P1, a[k][2] * P0 - a[k][0] * P2, a[k][0] * P1 - a[k][1] * P0,
                              -(a[k][0] * P0 + a[k][1] * P1 + a[k][2] * P2)};
          const float si = k < 2 ? si_px : si_dep;
          const float wh = fabsf(si * e[k]);
          const float w = si * si * (wh < huber_k ? 1.0f : huber_k / wh);
          accum_row<0x7fu>(acc, J, w, e[k]);
          cost = fmaf(w * e[k], e[k], cost);
        }
      }
    }
#pragma unroll
    for (int l = 0; l < TRACK_NACC; l++) {
      float v = l < NACC ? acc[l < NACC ? l : 0] : cost;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) red[warp][l] = v;
    }
    __syncthreads();
    if (tid < TRACK_NACC) {
      double sum = 0.0;
#pragma unroll
      for (int wq = 0; wq < PERSIST_THREADS / 32; wq++) sum += (double)red[wq][tid];
      partials[(size_t)s * TRACK_NACC + tid] = sum;
    }
    __syncthreads();
    if (tid == 0) {
      cuda::atomic_ref<unsigned, cuda::thread_scope_device> ticket(state->ticket);
      is_last = ticket.fetch_add(1u, cuda::memory_order_acq_rel) == (unsigned)S - 1;
    }
    __syncthreads();
    if (is_last) {
      // sum of the per-CTA partials in a FIXED order, by the whole CTA: thread (l, grp) adds the CTAs
      // c = grp (mod TRACK_GROUPS) of accumulator l (independent coalesced loads), thread l then adds the
      // group sums in order.  (One thread per accumulator walking all 148 partials cost ~11 us per step.)
      __shared__ double tot[TRACK_NACC], grp_sum[TRACK_GROUPS][TRACK_NACC];
      if (tid < TRACK_GROUPS * TRACK_NACC) {
        const int l = tid % TRACK_NACC, grp = tid / TRACK_NACC;
        double sum = 0.0;
        for (int c = grp; c < S; c += TRACK_GROUPS) sum += __ldcg(partials + (size_t)c * TRACK_NACC + l);
        grp_sum[grp][l] = sum;
      }
      __syncthreads();
      if (tid < TRACK_NACC) {
        double sum = 0.0;
#pragma unroll
        for (int grp = 0; grp < TRACK_GROUPS; grp++) sum += grp_sum[grp][tid];
        tot[tid] = sum;
      }
      __syncthreads();
      if (tid == 0) {
        state->ticket = 0;
        double g[7], tau[7];
        for (int a = 0; a < 7; a++) g[a] = -tot[28 + a];  // g = -A^T b with A = robust * J, b = robust * r
        const double new_cost = 0.5 * tot[35];
        const bool ok = solve7(tot, g, tau);
        int done = 0;
        if (!ok) {
          state->chol_fail = 1;  // tracker.py:88-90: torch.linalg.cholesky raises, the frame is skipped
          for (int k = 0; k < 8; k++) state->T[k] = pose_sm[k];
          done = 1;
        } else {
          float xi[7], T[8];
          double nrm2 = 0.0;
          for (int a = 0; a < 7; a++) { xi[a] = (float)tau[a]; nrm2 += (double)xi[a] * xi[a]; }
          for (int k = 0; k < 8; k++) T[k] = pose_sm[k];
          retract_pose(T, xi);
          for (int k = 0; k < 8; k++) state->T[k] = T[k];
          const double old_cost = it == 0 ? INFINITY : state->old_cost;
          const double rel_dec = fabs((old_cost - new_cost) / old_cost);  // NaN on the first step: not converged
          done = (rel_dec < (double)rel_error) || ((float)sqrt(nrm2) < delta_norm);
          state->old_cost = new_cost;
          state->cost = new_cost;
          state->iters = it + 1;
          state->converged = done;
        }
        flag_sm = done;
        state->done = done;
        __threadfence();
        st_release_u32(&state->gen, (unsigned)(it + 1));
      }
    } else if (tid == 0) {
      while (ld_acquire_u32(&state->gen) < (unsigned)(it + 1)) __nanosleep(20);
      flag_sm = __ldcg(&state->done);
    }
    __syncthreads();
    if (tid < 8) pose_sm[tid] = __ldcg(&state->T[tid]);
    __syncthreads();
    if (flag_sm) break;
  }
  if (s == 0 && tid == 0) {  // tracker.py:212
    float Tk[8], Tr[8], Tw[8];
    for (int k = 0; k < 8; k++) { Tk[k] = __ldg(T_WCk + k); Tr[k] = pose_sm[k]; }
    sim3_mul(Tk, Tr, Tw);
    for (int k = 0; k < 8; k++) { T_WCf_out[k] = Tw[k]; T_CkCf_out[k] = Tr[k]; }
  }
}

}  // namespace b200gn
