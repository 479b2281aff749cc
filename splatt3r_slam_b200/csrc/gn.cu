// b200slam — Gauss-Newton driver for sm_100a: stage-1 build kernel, stage-2 per-edge
// finalisation (M^T H0 M in fp64), assembly, fp64 Cholesky, Sim3 retraction, convergence flag.
// Replaces gauss_newton_{rays,calib,points}_cuda + SparseBlock + pose_retr_kernel of the
// reference (gn_kernels.cu:56-158, :414-452, :724-810, :1139-1227, :1545-1637).
// The whole max_iter loop is enqueued on one stream with NO host round trip per iteration
// (the reference does D2H -> Eigen SimplicialLLT on the CPU -> H2D -> .item() each iteration).
#include <cuda/atomic>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <vector>

#include "../../include/b200slam.h"
#include "dev_cache.cuh"
#include "gn_build.cuh"

namespace b200gn {

// ---- workspace layout -------------------------------------------------------------------------
#ifndef B200SLAM_SINGLE_CTA_MAX_D
#define B200SLAM_SINGLE_CTA_MAX_D 217  // largest system of the single-CTA packed solver (A/B: lower it to route mid-size graphs to the tiled solver)
#endif
struct State {  // head of the workspace
  int iters;      // iterations executed so far
  int done;       // set when ||dx|| < delta_thresh: later launches are no-ops
  int chol_fail;  // number of failed factorisations (dx = 0)
  unsigned ticket;  // arrival counter of the fused small-problem path (reset by the last CTA)
  unsigned long long t[4];  // %globaltimer stamps of the last fused tail: start, finalized, solved, end
  unsigned gen;  // grid-barrier generation of the persistent kernel (= iterations completed)
  int cg_iters;        // PCG iterations of the last large-graph solve
  double cg_rel;       // its final r^T M^-1 r relative to the initial value
  double cg_true_rel;  // pipelined PCG: the same quantity for the TRUE residual b - H x of the returned x
  int pcg_bad;         // pipelined PCG gave up (not converged / not SPD): the exact Cholesky runs instead, for the rest of the call
  int adj_ready;       // pose adjacency lists of the PCG solvers are built (once per call, the topology is fixed)
  int cg_total;        // PCG iterations summed over the call
  int dense_runs;      // number of exact tiled-Cholesky solves of the call
  unsigned long long prof[6];
  unsigned long long prof_busy_max, prof_busy_min;
  int solver_sel;      // large graphs: 0 = pipelined PCG, 1 = block-sparse Cholesky (set by b200slam_gn_analyze for this call)
  int sparse_runs;     // number of block-sparse Cholesky solves of the call  // per-CTA non-waiting cycles of the CG loop: extrema over the grid  // pipelined PCG, CTA 0 / warp 1, cycles summed over the CG loop: local+publish, wait, matvec+update
};
__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
constexpr size_t STATE_BYTES = 256;
static_assert(sizeof(State) <= STATE_BYTES, "State must fit the workspace head");

__host__ __device__ inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

struct Layout {
  size_t partials, edge_blocks, A, dense_ctrl, pcg, pcg2, sparse_ints, sparse_vals, rhs, rows, first, total;
  int sparse_nl_cap, sparse_work_cap;
};

// Number of CTAs (S) that split the n points of one edge when `Ecount` edges are built by one
// launch.  The build kernel is capped at 128 registers => 2 resident CTAs per SM => 296 slots on
// 148 SMs.  S is chosen to fill whole waves of 296 (measured on B200, N=32/E=128: 384 CTAs =
// 1.3 waves ran 27% slower than 256 CTAs), with at least 512 points per CTA.
// B200SLAM_GN_SPLITS overrides S (tuning aid).
constexpr int CTA_SLOTS = 444;  // 148 SMs x BUILD_MIN_CTAS (3) resident CTAs
constexpr int MAX_SPLITS = 160;
inline int choose_splits(int Ecount, int n) {
  static const int forced = getenv("B200SLAM_GN_SPLITS") ? atoi(getenv("B200SLAM_GN_SPLITS")) : 0;
  if (Ecount < 1) Ecount = 1;
  int maxS = (n + 511) / 512;
  if (maxS > MAX_SPLITS) maxS = MAX_SPLITS;
  if (maxS < 1) maxS = 1;
  if (forced > 0) return forced < maxS ? forced : maxS;
  int best = 1;
  double best_eff = 0.0;
  for (int S = 1; S <= maxS; S++) {
    const long long C = (long long)Ecount * S;
    const long long waves = (C + CTA_SLOTS - 1) / CTA_SLOTS;
    const double eff = (double)C / (double)(waves * CTA_SLOTS);
    if (eff > best_eff + 0.02) { best_eff = eff; best = S; }  // smallest S within 2% of the best fill
    if (waves >= 8 && eff > 0.9) break;
  }
  return best;
}

inline size_t partial_slots(int E) { return (size_t)E * 2 + 4096; }
inline int splits_for(int E, int Ecount, int n) {
  int S = choose_splits(Ecount, n);
  const size_t cap = partial_slots(E) / (size_t)(Ecount > 0 ? Ecount : 1);
  if ((size_t)S > cap) S = (int)cap;
  return S < 1 ? 1 : S;
}

inline Layout make_layout(int N, int E, int n) {
  Layout L;
  const size_t D = (size_t)7 * (N > 1 ? N - 1 : 1);
  size_t off = STATE_BYTES;
  // capacity in (edge, split) slots; splits_for() clamps S so that any shard of the E edges fits
  L.partials = off; off = align256(off + partial_slots(E) * NACC * sizeof(double));
  L.edge_blocks = off; off = align256(off + (size_t)E * 56 * sizeof(double));
  // stage-3 scratch: the dense matrix (single-CTA solver for D <= 217; above that the multi-CTA tiled
  // solver, which also needs a 256-byte control block and the inverted diagonal tiles), and the
  // buffers of the optional PCG solver
  // [ctrl 256 B | cnt M | ptr M+1 | list 2E (ints) | Minv 49M | b,x,r,p 4*7M | partials 4*148 (doubles)]
  L.A = off;
  off = align256(off + D * D * sizeof(double));
  L.dense_ctrl = off;
  if (D > B200SLAM_SINGLE_CTA_MAX_D) off = align256(off + 256 + ((D + 63) / 64) * 64 * 64 * sizeof(double));
  L.pcg = off;
  if (D > B200SLAM_SINGLE_CTA_MAX_D)
    off = align256(off + 256 + align256((2 * (D / 7) + 2 + 2 * (size_t)E) * sizeof(int)) +
                   (49 * (D / 7) + 4 * D + 4 * 148) * sizeof(double));
  // pipelined PCG (gn_pcg2.cuh): [xch 2*D | partial sums 2*160*2 (doubles) | tags 2*160 (u32)]; shares the adjacency of L.pcg
  L.pcg2 = off;
  if (D > B200SLAM_SINGLE_CTA_MAX_D)
    off = align256(off + align256(2 * D * sizeof(double)) + align256((size_t)2 * 160 * 2 * sizeof(double)) +
                   align256((size_t)2 * 160 * sizeof(unsigned)));
  // block-sparse Cholesky (gn_sparse.cuh): the plan's integer arrays and the value arrays, sized for the caps
  // under which the sparse solver is chosen at all (b200slam_gn_analyze)
  L.sparse_ints = L.sparse_vals = off;
  L.sparse_nl_cap = L.sparse_work_cap = 0;
  if (D > B200SLAM_SINGLE_CTA_MAX_D) {
    const size_t Mp = D / 7;
    L.sparse_nl_cap = (int)(48 * Mp < 65536 ? 48 * Mp : 65536);
    L.sparse_work_cap = (int)(256 * Mp < 200000 ? 256 * Mp : 200000);
    off = align256(off + (8 * Mp + 16 + 4 * (size_t)L.sparse_nl_cap + 3 * (size_t)L.sparse_work_cap + 5 * (size_t)E) * sizeof(int));
    L.sparse_vals = off;
    off = align256(off + (49 * Mp + 49 * (size_t)L.sparse_nl_cap + 28 * Mp) * sizeof(double));
  }
  L.rhs = off; off = align256(off + D * sizeof(double));
  L.rows = off; off = align256(off + (size_t)2 * E * sizeof(int64_t));
  L.first = off; off = align256(off + (size_t)2 * E);
  L.total = off;
  return L;
}

}  // namespace b200gn

#include "gn_solve.cuh"

namespace b200gn {

// ---- stage 1 ----------------------------------------------------------------------------------
template <int KIND, bool FUSED>
#ifndef BUILD_MIN_CTAS
#define BUILD_MIN_CTAS 3  // resident CTAs per SM the build kernel is compiled for: 80 registers, 24 warps/SM (A/B r02q at N=256: 2 CTAs = 117 registers 3.99 ms, 3 CTAs 3.46 ms, 4 CTAs = 64 registers + spills 3.81 ms)
#endif
__global__ void __launch_bounds__(BUILD_THREADS, BUILD_MIN_CTAS)
build_kernel(float *Twc, const float *__restrict__ Xs, const float *__restrict__ Cs,
             const int64_t *__restrict__ ii_row, const int64_t *__restrict__ jj_row,
             const int64_t *__restrict__ idx, const uint8_t *__restrict__ valid,
             const float *__restrict__ Q, int n, int e0, BuildConsts C, double *partials, State *state,
             int N, int E, float delta_thresh, float *dx, double *edge_blocks) {
  if (state->done) return;
  if (KIND == B200SLAM_GN_CALIB) {  // gn_kernels.cu:1273-1278
    C.fx = __ldg(C.Kptr + 0); C.fy = __ldg(C.Kptr + 4); C.cx = __ldg(C.Kptr + 2); C.cy = __ldg(C.Kptr + 5);
  }
  const int e = e0 + blockIdx.y;
  const int S = gridDim.x, s = blockIdx.x;
  const int ix = (int)ii_row[e], jx = (int)jj_row[e];
  const EdgePose P = rel_pose(Twc + 8 * ix, Twc + 8 * jx);

  const size_t xi_slot = C.slots ? (size_t)C.slots[ix] : (size_t)ix, xj_slot = C.slots ? (size_t)C.slots[jx] : (size_t)jx;
  const float *Xi_base = Xs + xi_slot * n * 3, *Xj_base = Xs + xj_slot * n * 3;
  const float *Ci_base = Cs + xi_slot * n, *Cj_base = Cs + xj_slot * n;
  const bool second = e >= C.E1;
  const int64_t *idx_e = (second ? C.idx2 : idx) + (size_t)e * n;
  const uint8_t *val_e = (second ? C.valid2 : valid) + (size_t)e * n;
  const float *Q_e = (second ? C.Q2 : Q) + (size_t)e * n;

  float acc[NACC];
#pragma unroll
  for (int l = 0; l < NACC; l++) acc[l] = 0.0f;

  const bool vec_ok = C.vec_ok != 0;         // 16-byte loads: n % 4 == 0 and aligned bases (set_alignment)
  const int groups = vec_ok ? (n >> 2) : 0;  // groups of 4 points
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
  // scalar path (n % 4 != 0): the split takes a contiguous range of points, one per thread
  if (!vec_ok) {
    const int per = (n + S - 1) / S;
    const int k_end = min(n, (s + 1) * per);
    for (int k = s * per + threadIdx.x; k < k_end; k += BUILD_THREADS) {
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
    partials[((size_t)blockIdx.y * S + s) * NACC + threadIdx.x] = sum;  // slot of the edge WITHIN this launch
  }
  if (FUSED) {
    // last-arriving CTA finishes the iteration: finalize + assemble + solve + retract
    // (release: this CTA's partials; acquire: everyone else's) through one acq_rel RMW
    __shared__ bool is_last;
    __syncthreads();
    if (threadIdx.x == 0) {
      cuda::atomic_ref<unsigned, cuda::thread_scope_device> ticket(state->ticket);
      is_last = ticket.fetch_add(1u, cuda::memory_order_acq_rel) == gridDim.x * gridDim.y - 1;
    }
    __syncthreads();
    if (!is_last) return;
    if (threadIdx.x == 0) { state->ticket = 0; state->t[0] = gtimer(); }
    small_tail<BUILD_THREADS>(Twc, ii_row, jj_row, partials, S, E, N, delta_thresh, dx, edge_blocks, state);
  }
}

// ---- keyframe ids -> rows (replaces torch::_unique + searchsorted, gn_kernels.cu:160-169) --------
// row(x) = number of distinct ids smaller than x among cat(ii, jj).
__device__ __forceinline__ int64_t id_at(const int64_t *ii, const int64_t *jj, int E, int p) {
  return p < E ? ii[p] : jj[p - E];
}
__global__ void first_occurrence_kernel(const int64_t *__restrict__ ii, const int64_t *__restrict__ jj, int E,
                                        uint8_t *__restrict__ first, const int64_t *gate) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= 2 * E || (gate && gate[p] >= 0)) return;  // gate: rows already computed by rows_bitmap_kernel
  const int64_t x = id_at(ii, jj, E, p);
  bool f = true;
  for (int q = 0; q < p && f; q++) f = id_at(ii, jj, E, q) != x;
  first[p] = f ? 1 : 0;
}
__global__ void rank_kernel(const int64_t *__restrict__ ii, const int64_t *__restrict__ jj, int E,
                            const uint8_t *__restrict__ first, int64_t *rows, int gated) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= 2 * E || (gated && rows[p] >= 0)) return;  // (a thread reads only its own slot before writing it)
  const int64_t x = id_at(ii, jj, E, p);
  int r = 0;
  for (int q = 0; q < 2 * E; q++) r += (first[q] && id_at(ii, jj, E, q) < x) ? 1 : 0;
  rows[p] = r;  // rows[0..E) = ii_row, rows[E..2E) = jj_row
}

// Short edge lists (2E <= 1024, every tracker / local-window call): both quadratic phases in ONE single-CTA launch
// (shared-memory copy of the ids, block barrier between the phases) instead of two launches.
__global__ void __launch_bounds__(1024)
rows_small_kernel(const int64_t *__restrict__ ii, const int64_t *__restrict__ jj, int E, int64_t *__restrict__ rows) {
  __shared__ long long ids[1024];
  __shared__ unsigned char first[1024];
  const int p = threadIdx.x, n2 = 2 * E;
  if (p < n2) ids[p] = id_at(ii, jj, E, p);
  __syncthreads();
  if (p < n2) {
    const long long x = ids[p];
    bool f = true;
    for (int q = 0; q < p && f; q++) f = ids[q] != x;
    first[p] = f ? 1 : 0;
  }
  __syncthreads();
  if (p < n2) {
    const long long x = ids[p];
    int r = 0;
    for (int q = 0; q < n2; q++) r += (first[q] && ids[q] < x) ? 1 : 0;
    rows[p] = r;
  }
}

// Large edge lists (2E > 1024): O(E) instead of O(E^2).  ONE CTA: id range by a block reduction; when the range fits
// the scratch as a presence bitmap (range <= 16 E ids and <= ROWS_BM_WORDS words: keyframe ids are small integers),
// set the bits, exclusive prefix-popcount over the words in shared memory, row = prefix[word] + popcount(lower bits).
// Otherwise every rows[p] is set to -1 and the two quadratic kernels above (gated per thread on rows[p] < 0) do the
// work, so the result never depends on the id values.
constexpr int ROWS_BM_WORDS = 8192;  // 32 KB of shared memory: id range <= 262 144
__global__ void __launch_bounds__(1024)
rows_bitmap_kernel(const int64_t *__restrict__ ii, const int64_t *__restrict__ jj, int E, unsigned *__restrict__ bm,
                   size_t bm_bytes, int64_t *__restrict__ rows) {
  __shared__ long long red_lo[32], red_hi[32];
  __shared__ int pref[ROWS_BM_WORDS];
  __shared__ int wsum[32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, n2 = 2 * E;
  long long lo = 0x7fffffffffffffffll, hi = -0x7fffffffffffffffll - 1;
  for (int p = tid; p < n2; p += 1024) {
    const long long x = id_at(ii, jj, E, p);
    lo = x < lo ? x : lo;
    hi = x > hi ? x : hi;
  }
  for (int o = 16; o > 0; o >>= 1) {
    const long long a = __shfl_xor_sync(0xffffffffu, lo, o), b = __shfl_xor_sync(0xffffffffu, hi, o);
    lo = a < lo ? a : lo;
    hi = b > hi ? b : hi;
  }
  if (lane == 0) { red_lo[warp] = lo; red_hi[warp] = hi; }
  __syncthreads();
  lo = red_lo[0]; hi = red_hi[0];
  for (int k = 1; k < 32; k++) { lo = red_lo[k] < lo ? red_lo[k] : lo; hi = red_hi[k] > hi ? red_hi[k] : hi; }
  const unsigned long long range = (unsigned long long)hi - (unsigned long long)lo + 1ull;  // 0 on full-range wrap
  const unsigned long long words = (range + 31ull) >> 5;
  if (range == 0 || words > (unsigned long long)ROWS_BM_WORDS || words * 4ull > bm_bytes) {
    for (int p = tid; p < n2; p += 1024) rows[p] = -1;  // the quadratic kernels take over
    return;
  }
  const int W = (int)words;
  for (int k = tid; k < W; k += 1024) bm[k] = 0u;
  __syncthreads();
  for (int p = tid; p < n2; p += 1024) {
    const unsigned long long d = (unsigned long long)id_at(ii, jj, E, p) - (unsigned long long)lo;
    atomicOr(bm + (d >> 5), 1u << (d & 31ull));
  }
  __syncthreads();
  // exclusive prefix of the word popcounts: thread t owns words [t*C, (t+1)*C)
  const int C = (W + 1023) / 1024;
  int local = 0;
  for (int k = tid * C; k < min(W, (tid + 1) * C); k++) local += __popc(__ldcg(bm + k));
  int incl = local;
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) wsum[warp] = incl;
  __syncthreads();
  int base = incl - local;
  for (int k = 0; k < warp; k++) base += wsum[k];
  for (int k = tid * C; k < min(W, (tid + 1) * C); k++) {
    pref[k] = base;
    base += __popc(__ldcg(bm + k));
  }
  __syncthreads();
  for (int p = tid; p < n2; p += 1024) {
    const unsigned long long d = (unsigned long long)id_at(ii, jj, E, p) - (unsigned long long)lo;
    const unsigned w = __ldcg(bm + (d >> 5)), bit = (unsigned)(d & 31ull);
    rows[p] = pref[d >> 5] + __popc(w & ((1u << bit) - 1u));
  }
}

// ---- stage 2 (large problems): one warp per edge -----------------------------------------------
__global__ void __launch_bounds__(128)
finalize_kernel(const float *__restrict__ Twc, const int64_t *__restrict__ ii_row,
                const double *__restrict__ partials, int S, int e0, int Ecount,
                double *__restrict__ edge_blocks, const State *__restrict__ state) {
  if (state->done) return;
  const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (wid >= Ecount) return;
  __shared__ double sc[4][FIN_SCRATCH];
  const int e = e0 + wid;
  finalize_edge(Twc, ii_row, partials, S, e, wid, edge_blocks + (size_t)e * 56, sc[threadIdx.x >> 5], lane);
}

// ---- stage 2 fused with the exchange of the edge-sharded multi-GPU solve ------------------------
// One warp per edge finalizes its block and stores the 56 doubles into EVERY rank's edge-block array through
// NVLink peer pointers (symmetric memory), so the "all-gather" is the kernel's own epilogue: no separate
// collective, no staging copy.  The last CTA of the launch (ticket) then publishes flags[rank] = seq in every
// rank's flag array with a system-scope release; consumers wait with peers_wait_kernel.  The buffers are
// double-buffered by the caller (a rank can run at most one iteration ahead of its peers).
constexpr int MAX_PEERS = 16;
struct PeerPtrs {
  double *blocks[MAX_PEERS];    // each rank's (E,56) edge-block array
  unsigned *flags[MAX_PEERS];   // each rank's flag array (>= world entries)
};
__global__ void __launch_bounds__(128)
finalize_push_kernel(const float *__restrict__ Twc, const int64_t *__restrict__ ii_row,
                     const double *__restrict__ partials, int S, int e0, int Ecount, PeerPtrs P, int world, int rank,
                     unsigned seq, State *state) {
  const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  __shared__ double sc[4][FIN_SCRATCH];
  __shared__ bool is_last;
  if (!state->done && wid < Ecount) {
    const int e = e0 + wid;
    double *own = P.blocks[rank] + (size_t)e * 56;
    finalize_edge(Twc, ii_row, partials, S, e, wid, own, sc[threadIdx.x >> 5], lane);
    const double v0 = own[lane], v1 = lane + 32 < 56 ? own[lane + 32] : 0.0;  // written by this warp (syncwarp inside)
    for (int p = 0; p < world; p++) {
      if (p == rank) continue;
      double *dst = P.blocks[p] + (size_t)e * 56;
      dst[lane] = v0;
      if (lane + 32 < 56) dst[lane + 32] = v1;
    }
  }
  // completion: every CTA's peer stores happen-before its ticket increment; the last CTA's system-scope release
  // fence is cumulative over them
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    cuda::atomic_ref<unsigned, cuda::thread_scope_device> ticket(state->ticket);
    is_last = ticket.fetch_add(1u, cuda::memory_order_acq_rel) == gridDim.x - 1;
  }
  __syncthreads();
  if (is_last) {
    if (threadIdx.x == 0) state->ticket = 0;
    if ((int)threadIdx.x < world) {
      __threadfence_system();
      asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(P.flags[threadIdx.x] + rank), "r"(seq) : "memory");
    }
  }
}

// Wait until every rank has published `seq` (its rows of the edge-block array are in THIS rank's memory).
__global__ void peers_wait_kernel(const unsigned *flags, int world, unsigned seq) {
  const int p = threadIdx.x;
  if (p < world) {
    unsigned v;
    do {
      asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + p) : "memory");
    } while ((int)(v - seq) < 0);
  }
  __syncwarp();
  asm volatile("fence.acq_rel.sys;" ::: "memory");
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

#include "gn_persistent.cuh"
#include "gn_track.cuh"
#include "gn_pcg.cuh"
#include "gn_pcg2.cuh"
#include "gn_sparse_plan.h"
#include "gn_sparse.cuh"
#include "gn_dense.cuh"
#include "gn_api.cuh"
#include "gn_track_api.cuh"
