// b200slam — refine_matches, tiled kernel with bulk-copy (TMA engine) staging of the descriptor
// windows in shared memory.  Same arithmetic and scan order as refine_f16_kernel (bit-exact with
// matching_kernels.cu:24-80); different data path:
//
//   A CTA owns a TW x TH (32x16) tile of query pixels.  At every dilation level the CTA reduces the bounding
//   box of its queries' current centres, grows it by the level's reach (radius*d), clips it to the
//   image and lets ONE thread issue one `cp.async.bulk` (UBLKCP) per box row: (box_w * fdim * 2)
//   contiguous bytes of D11 -> shared memory, completion signalled on an mbarrier.  Every query
//   whose window lies inside the box then scores its 49 candidates from shared memory with
//   conflict-free 16-byte LDS (48-byte pixel stride) instead of 12-sector scattered LDGs — the
//   reference layout costs ~25 L1 cycles per warp-load, which bounded the first version of this
//   kernel at 95% L1 throughput.  Queries whose window leaves the box (outliers, or a tile whose
//   matches are not spatially coherent) take the global-memory path for that level; the result is
//   identical either way.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace rtile {

// Query tile and box size (build macros for A/B runs).  Measured on B200, 512x384 pairs:
//   16x16 queries, 112 KB box, 2 CTAs/SM   265 us  (the dilation-5 box of a 16x16 tile is 48x48 px = 110.6 KB: any
//                                                   spread of the match centres overflows it, and the queries
//                                                   left outside take the global-memory path)
//   16x16 queries, 224 KB box, 1 CTA/SM    257 us  (no overflow, half the warps)
//   32x16 queries, 224 KB box, 1 CTA/SM    242 us  (no overflow, same 16 warps per SM)   <- default
#ifndef RT_TW
#define RT_TW 32
#endif
#ifndef RT_TH
#define RT_TH 16
#endif
#ifndef RT_BOX_KB
#define RT_BOX_KB 224
#endif
#ifndef RT_SLACK
#define RT_SLACK 2  // spread of the match centres (px) a level's box must absorb before the level is split in two
#endif
#ifndef RT_ISSUE_LANES
#define RT_ISSUE_LANES 32
#endif
#ifndef RT_PAIRS
#define RT_PAIRS 1  // two candidates per HMUL2/HADD2 (register transpose), see mac_chunk_pairs
#endif
#ifndef RT_DEAL
#define RT_DEAL 1  // how the re-deal places the queries that do not fit the one-per-class pattern (see the kernel)
#endif
#ifndef RT_MERGE
#define RT_MERGE 1  // stage one box for all remaining levels as soon as it fits
#endif
#ifndef RT_PERMUTE_FIRST
#define RT_PERMUTE_FIRST 1  // re-deal at the first (coarsest) level too (191.9 vs 193.4 us with the class-balanced deal)
#endif
#ifndef RT_LINEAR
#define RT_LINEAR 0  // (measured: slower, see experiments/README.md) all candidates of a level as one scan-order list in groups of 8 (no odd seventh candidate per column; centre skipped after the first level)
#endif
#ifndef RT_PREFETCH
#define RT_PREFETCH 0  // (measured: slower) software-pipelined candidate loads in scan_level
#endif
#ifndef RT_OVERFLOW_FIT
#define RT_OVERFLOW_FIT 1  // a tile whose bounding box overflows stages the largest box that fits, not the nominal one
#endif
#ifndef RT_PERMUTE
#define RT_PERMUTE 1  // re-deal the queries of a warp to its lanes per level so that quarter-warps hit distinct banks
#endif
#ifndef RT_STATS
#define RT_STATS 0  // A/B builds only: per-level counters (live queries, queries outside the staged box, ...) in g_rt_stats
#endif
#if RT_STATS
__device__ unsigned long long g_rt_stats[8][4];  // [dilation][live, out of box, tiles whose bounding box overflowed, tiles merged]
#endif
constexpr int TW = RT_TW, TH = RT_TH, THREADS = TW * TH, SLACK = RT_SLACK;
constexpr int BOX_BYTES = RT_BOX_KB * 1024;  // shared-memory box per CTA

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE;\n"
      "bra WAIT_LOOP;\n"
      "DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}

// Sequential fp16-rounded dot product over one 16-byte chunk (8 channels) for NJ candidates at
// once (NJ independent dependency chains -> ILP); q4 = the query's 8 channels as 4 half2.
template <int NJ>
__device__ __forceinline__ void mac_chunk(__half (&s)[NJ], const __half2 (&q4)[4], const uint4 (&c)[NJ]) {
#pragma unroll
  for (int t = 0; t < 4; t++) {
    __half2 pr[NJ];
#pragma unroll
    for (int j = 0; j < NJ; j++) {
      const uint32_t cw = t == 0 ? c[j].x : t == 1 ? c[j].y : t == 2 ? c[j].z : c[j].w;
      pr[j] = __hmul2(q4[t], *reinterpret_cast<const __half2 *>(&cw));
    }
#pragma unroll
    for (int j = 0; j < NJ; j++) s[j] = __hadd(s[j], __low2half(pr[j]));
#pragma unroll
    for (int j = 0; j < NJ; j++) s[j] = __hadd(s[j], __high2half(pr[j]));
  }
}

// The same sums for candidate PAIRS: candidates 2p and 2p+1 of the column run in the two lanes of one HMUL2 +
// HADD2 (their channel words are transposed in registers, 2 PRMT per word pair; the query channel is
// broadcast by the operand selector).  Per 4 MACs: 2 PRMT + 2 HMUL2 + 2 HADD2 instead of 2 HMUL2 + 4 HADD +
// the half extractions of the scalar form.  `_rn` intrinsics: a plain __hadd2(s, __hmul2(a, b)) would be
// contracted into a single-rounding HFMA2.  The last odd candidate is paired with itself.
template <int NJ>
__device__ __forceinline__ void mac_chunk_pairs(__half2 (&s2)[(NJ + 1) / 2], const __half2 (&q4)[4], const uint4 (&c)[NJ]) {
#pragma unroll
  for (int t = 0; t < 4; t++) {
    const __half2 qlo = __low2half2(q4[t]), qhi = __high2half2(q4[t]);
#pragma unroll
    for (int p = 0; p < (NJ + 1) / 2; p++) {
      const int ja = 2 * p, jb = 2 * p + 1 < NJ ? 2 * p + 1 : 2 * p;
      const uint32_t wa = t == 0 ? c[ja].x : t == 1 ? c[ja].y : t == 2 ? c[ja].z : c[ja].w;
      const uint32_t wb = t == 0 ? c[jb].x : t == 1 ? c[jb].y : t == 2 ? c[jb].z : c[jb].w;
      const __half2 a = *reinterpret_cast<const __half2 *>(&wa), bb = *reinterpret_cast<const __half2 *>(&wb);
      s2[p] = __hadd2_rn(s2[p], __hmul2_rn(qlo, __lows2half2(a, bb)));
      s2[p] = __hadd2_rn(s2[p], __hmul2_rn(qhi, __highs2half2(a, bb)));
    }
  }
}

struct Box {
  int x0, y0, bw, bh;  // origin and extent in pixels (bw*bh == 0: nothing staged)
};

// One column of candidates (fixed u, NJ = 2*RADIUS+1 values of v at stride d) for one query.
// LOADER(v_clamped, chunk) returns the 16-byte chunk of pixel (u, v).
template <int FDIM, int RADIUS, class Loader>
__device__ __forceinline__ void scan_column(const __half2 (&q)[FDIM / 2], int u, int v0, int d, int h,
                                            const Loader &load, float &max_score, int &u_new, int &v_new) {
  constexpr int NJ = 2 * RADIUS + 1;
  __half s[NJ];
  int vv[NJ];
#pragma unroll
  for (int j = 0; j < NJ; j++) {
    s[j] = __ushort_as_half(0);
    vv[j] = v0 + (j - RADIUS) * d;
  }
#if RT_PAIRS
  __half2 s2[(NJ + 1) / 2];
#pragma unroll
  for (int p = 0; p < (NJ + 1) / 2; p++) s2[p] = __half2half2(__ushort_as_half(0));
#endif
#pragma unroll
  for (int ch = 0; ch < FDIM / 8; ch++) {
    uint4 c[NJ];
#pragma unroll
    for (int j = 0; j < NJ; j++) c[j] = load(min(max(vv[j], 0), h - 1), ch);
    const __half2 q4[4] = {q[ch * 4 + 0], q[ch * 4 + 1], q[ch * 4 + 2], q[ch * 4 + 3]};
#if RT_PAIRS
    mac_chunk_pairs<NJ>(s2, q4, c);
#else
    mac_chunk<NJ>(s, q4, c);
#endif
  }
#if RT_PAIRS
#pragma unroll
  for (int j = 0; j < NJ; j++) s[j] = (j & 1) ? __high2half(s2[j / 2]) : __low2half(s2[j / 2]);
#endif
#pragma unroll
  for (int j = 0; j < NJ; j++) {  // reference scan order: j (v) inner, strict >
    const float sf = __half2float(s[j]);
    if (vv[j] >= 0 && vv[j] < h && sf > max_score) {
      max_score = sf;
      u_new = u;
      v_new = vv[j];
    }
  }
}

// All (2*RADIUS+1)^2 candidates of one level for one query as ONE list in the reference's scan order (u outer, v
// inner, strict >), eight at a time = four HMUL2/HADD2 pairs with no idle half (a column of 7 leaves one).  The
// centre candidate is scored at the first level only: at every later level the centre is the pixel whose score
// IS max_score (the previous level's winner: same operands, same sum) or, when nothing has won yet, the
// unchanged input pixel whose score already failed `> max_score` at the first level - it can never win again.  It
// sits in the middle of the scan order (index 24 of 49), so the 48 others are exactly 3 + 3 groups around it.
// roff / coff: element offsets (uint4 units) of the 7 clamped rows / columns; rmask / cmask: which lie in the image.
// LOADER(offset, chunk) returns one 16-byte chunk.
template <int FDIM, int RADIUS, class Loader>
__device__ __forceinline__ void scan_level(const __half2 (&q)[FDIM / 2], const int (&roff)[2 * RADIUS + 1],
                                           const int (&coff)[2 * RADIUS + 1], unsigned rmask, unsigned cmask,
                                           bool with_centre, const Loader &load, float &max_score, int &best) {
  constexpr int NJ = 2 * RADIUS + 1, NC = NJ * NJ, CT = RADIUS * NJ + RADIUS, G = 8, NG = (NC - 1) / G, NCH = FDIM / 8;
  static_assert((NC - 1) % (2 * G) == 0, "the candidates before and after the centre must form whole groups");
  // candidate number (scan order) of slot k of group g: the centre is not part of any group
  auto cand = [](int g, int k) { return g * G + k < CT ? g * G + k : g * G + k + 1; };
#if RT_PREFETCH
  // software pipeline over the NG * NCH (group, chunk) steps: the eight 16-byte chunks of step s + 1 are requested
  // before the 96 HMUL2/HADD2/PRMT of step s (ncu: short-scoreboard, i.e. waiting for LDS, was the top stall)
  uint4 cb[2][G];
#pragma unroll
  for (int k = 0; k < G; k++) cb[0][k] = load(roff[cand(0, k) % NJ] + coff[cand(0, k) / NJ], 0);
#endif
  __half2 s2[G / 2];
#pragma unroll
  for (int g = 0; g < NG; g++) {
    if (g == NG / 2 && with_centre) {
      const int off = roff[RADIUS] + coff[RADIUS];
      __half2 s1[1] = {__half2half2(__ushort_as_half(0))};
#pragma unroll
      for (int ch = 0; ch < NCH; ch++) {
        const uint4 c[1] = {load(off, ch)};
        const __half2 q4[4] = {q[ch * 4 + 0], q[ch * 4 + 1], q[ch * 4 + 2], q[ch * 4 + 3]};
        mac_chunk_pairs<1>(s1, q4, c);
      }
      const float sf = __half2float(__low2half(s1[0]));
      if (((rmask >> RADIUS) & (cmask >> RADIUS) & 1u) && sf > max_score) {
        max_score = sf;
        best = CT;
      }
    }
#pragma unroll
    for (int p = 0; p < G / 2; p++) s2[p] = __half2half2(__ushort_as_half(0));
#pragma unroll
    for (int ch = 0; ch < NCH; ch++) {
      const __half2 q4[4] = {q[ch * 4 + 0], q[ch * 4 + 1], q[ch * 4 + 2], q[ch * 4 + 3]};
#if RT_PREFETCH
      const int s = g * NCH + ch, gn = (s + 1) / NCH, chn = (s + 1) % NCH;
      if (s + 1 < NG * NCH) {
#pragma unroll
        for (int k = 0; k < G; k++) cb[(s + 1) & 1][k] = load(roff[cand(gn, k) % NJ] + coff[cand(gn, k) / NJ], chn);
      }
      mac_chunk_pairs<G>(s2, q4, cb[s & 1]);
#else
      uint4 c[G];
#pragma unroll
      for (int k = 0; k < G; k++) c[k] = load(roff[cand(g, k) % NJ] + coff[cand(g, k) / NJ], ch);
      mac_chunk_pairs<G>(s2, q4, c);
#endif
    }
#pragma unroll
    for (int k = 0; k < G; k++) {
      const int t = cand(g, k);
      const float sf = __half2float((k & 1) ? __high2half(s2[k / 2]) : __low2half(s2[k / 2]));
      if (((rmask >> (t % NJ)) & (cmask >> (t / NJ)) & 1u) && sf > max_score) {
        max_score = sf;
        best = t;
      }
    }
  }
}

template <int FDIM, int RADIUS>
__global__ void __launch_bounds__(THREADS, RT_BOX_KB > 113 ? 1 : 2)
refine_tile_kernel(const __half *__restrict__ D11, const __half *__restrict__ D21,
                   const int64_t *__restrict__ p1, int64_t *__restrict__ p1_new, int h, int w, int n,
                   int qw, int bands, int dilation_max, int lin_out) {
  extern __shared__ __align__(128) unsigned char box_mem[];
  __shared__ uint64_t bar;
  __shared__ int red[2][THREADS / 32][8];
  constexpr int PIX_BYTES = FDIM * 2;
  constexpr int MAX_PX = BOX_BYTES / PIX_BYTES;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tiles_x = (qw + TW - 1) / TW;
  const int qx = (blockIdx.x % tiles_x) * TW + (tid % TW);
  // query rows are cut into `bands` bands of <= TH rows (band k = rows [k*qh/bands, (k+1)*qh/bands)): the host
  // picks the band count whose tile count fills whole waves of SMs (512x384: 27 bands of 14-15 rows = 432
  // tiles in 3 waves of 148, instead of 24 bands of 16 rows = 384 tiles in the same 3 waves)
  const int band = blockIdx.x / tiles_x, qh = n / qw;
  const int qy0 = (int)((long long)band * qh / bands), qy1 = (int)((long long)(band + 1) * qh / bands);
  const int qy = qy0 + (tid / TW);
  const long long qlin = (long long)qy * qw + qx;
  const bool active = qx < qw && qy < qy1;
  const size_t bq = (size_t)blockIdx.y * n + (active ? qlin : 0);
  const __half *img = D11 + (size_t)blockIdx.y * h * w * FDIM;

  if (tid == 0) mbar_init(&bar, 1);

  __half2 q[FDIM / 2];
  {
    const uint4 *src = reinterpret_cast<const uint4 *>(D21 + bq * FDIM);
#pragma unroll
    for (int v = 0; v < FDIM / 8; v++) {
      const uint4 c = __ldg(src + v);
      q[v * 4 + 0] = *reinterpret_cast<const __half2 *>(&c.x);
      q[v * 4 + 1] = *reinterpret_cast<const __half2 *>(&c.y);
      q[v * 4 + 2] = *reinterpret_cast<const __half2 *>(&c.z);
      q[v * 4 + 3] = *reinterpret_cast<const __half2 *>(&c.w);
    }
  }
  const longlong2 pin = __ldg(reinterpret_cast<const longlong2 *>(p1) + bq);
  // centres far outside the image can never produce an in-image candidate: clamp them into int
  // range (reach <= radius*dilation_max) so the arithmetic below stays in 32 bits
  const long long reach = (long long)RADIUS * dilation_max;
  const bool can_hit = pin.x >= -reach && pin.x < w + reach && pin.y >= -reach && pin.y < h + reach;
  int u0 = can_hit ? (int)pin.x : -(1 << 20), v0 = can_hit ? (int)pin.y : -(1 << 20);
  int u_new = u0, v_new = v0;
  float max_score = 0.0f;
  uint32_t parity = 0;
  int buf = 0;
  // queries that could never hit the image keep their input pixel verbatim (64-bit): written here, so
  // that only (centre, score, descriptor, output slot) travel with a query when lanes are re-dealt
  if (active && !can_hit) {
    if (lin_out) p1_new[bq] = pin.x + (long long)w * pin.y;
    else reinterpret_cast<longlong2 *>(p1_new)[bq] = pin;
  }
  bool live = active && can_hit;
  int qout = (int)qlin;  // output slot of the query this lane currently holds (n < 2^31 on this path)
#if RT_PERMUTE
  __shared__ unsigned char lane_of_slot[THREADS / 32][32];
#endif
  __syncthreads();  // mbarrier initialised

  Box b{0, 0, 0, 0};
  bool merged = false;  // b already holds every pixel the remaining levels can touch: no more barriers or copies
  for (int pass = 0, d = dilation_max; d > 0; pass++) {
    const int rd = RADIUS * d;
    // A level whose nominal box (tile + reach + SLACK px) exceeds the shared-memory box is
    // scanned in two passes of candidate columns, i <= 0 then i > 0 (same i-outer scan order), each
    // with a box of half the reach in x.  With the default 224 KB box no level splits.
#if RT_LINEAR
    constexpr bool split = false;  // the host launches this build only when every level's nominal box fits
#else
    const bool split = (TW + 2 * rd + SLACK) * (TH + 2 * rd + SLACK) > MAX_PX;
#endif
    const int ilo = split && pass ? d : -rd, ihi = split && !pass ? 0 : rd;
    bool fresh = false;  // a copy was issued in this pass
    if (!merged) {
    // ---- bounding box of the centres of this tile (queries that can hit the image only) --------
    // One block barrier per pass: every warp leaves its partial box in shared memory, and after the
    // barrier EVERY warp reduces the partials itself (shuffles), so nobody waits for an elected thread;
    // warp 0 then issues the bulk copies, one box row per lane.  `red` is double-buffered by pass: a
    // fast warp writes the next pass's partials while a slow one may still be reading this pass's.
    const bool part = live;
    int vals[7] = {part ? u0 : (1 << 29), part ? -u0 : (1 << 29), part ? v0 : (1 << 29),
                   part ? -v0 : (1 << 29), part ? u0 : 0, part ? v0 : 0, part ? 1 : 0};
#pragma unroll
    for (int k = 0; k < 4; k++) vals[k] = __reduce_min_sync(0xffffffffu, vals[k]);
#pragma unroll
    for (int k = 4; k < 7; k++) vals[k] = __reduce_add_sync(0xffffffffu, vals[k]);
    int(*redp)[8] = red[buf];
    buf ^= 1;
    if (lane == 0) {
#pragma unroll
      for (int k = 0; k < 7; k++) redp[warp][k] = vals[k];
    }
    __syncthreads();  // also: every thread has finished reading the previous pass's box
    {
      const bool has = lane < THREADS / 32;
#pragma unroll
      for (int k = 0; k < 4; k++) vals[k] = __reduce_min_sync(0xffffffffu, has ? redp[lane][k] : (1 << 29));
#pragma unroll
      for (int k = 4; k < 7; k++) vals[k] = __reduce_add_sync(0xffffffffu, has ? redp[lane][k] : 0);
    }
    b = Box{0, 0, 0, 0};
    {
      const int umin = vals[0], numax = vals[1], vmin = vals[2], nvmax = vals[3], su = vals[4], sv = vals[5], c = vals[6];
      if (c > 0 && RT_MERGE && !split) {
        // One box for ALL remaining levels (total reach RADIUS * (d + ... + 1)) if it fits: the finer levels
        // then run without block barriers, box reductions or copies (warps drift apart freely).
        const int rall = RADIUS * d * (d + 1) / 2;
        const int x0 = max(umin - rall, 0), x1 = min(-numax + rall, w - 1);
        const int y0 = max(vmin - rall, 0), y1 = min(-nvmax + rall, h - 1);
        if (x1 >= x0 && y1 >= y0 && (x1 - x0 + 1) * (y1 - y0 + 1) <= MAX_PX) {
          b = Box{x0, y0, x1 - x0 + 1, y1 - y0 + 1};
          merged = true;
#if RT_STATS
          if (tid == 0) atomicAdd(&g_rt_stats[d & 7][3], 1ull);
#endif
        }
      }
      if (c > 0 && !merged) {
        int x0 = max(umin + ilo, 0), x1 = min(-numax + ihi, w - 1);
        int y0 = max(vmin - rd, 0), y1 = min(-nvmax + rd, h - 1);
        if (x1 >= x0 && y1 >= y0) {
          if ((x1 - x0 + 1) * (y1 - y0 + 1) > MAX_PX) {
#if RT_STATS
            if (tid == 0) atomicAdd(&g_rt_stats[d & 7][2], 1ull);
#endif
#if RT_OVERFLOW_FIT
            // incoherent tile: the largest box of the bounding box's shape that fits, around the mean centre
            const float shrink = sqrtf((float)MAX_PX / (float)((x1 - x0 + 1) * (y1 - y0 + 1)));
            int side_x = max(min((int)((x1 - x0 + 1) * shrink), w), min(TW + ihi - ilo + 2, w));
            int side_y = min(min(y1 - y0 + 1, h), MAX_PX / side_x);
            side_x = min(min(x1 - x0 + 1, w), MAX_PX / side_y);
#else
            // incoherent tile: keep a box of the nominal size around the mean centre
            const int side_x = min(TW + ihi - ilo + 2, w), side_y = min(min(TH + 2 * rd + 2, h), MAX_PX / side_x);
#endif
            const int cx = su / c + (ilo + ihi) / 2, cy = sv / c;
            x0 = min(max(cx - side_x / 2, 0), w - side_x); x1 = x0 + side_x - 1;
            y0 = min(max(cy - side_y / 2, 0), h - side_y); y1 = y0 + side_y - 1;
          }
          b = Box{x0, y0, x1 - x0 + 1, y1 - y0 + 1};
        }
      }
    }
    if (warp == 0 && b.bw * b.bh > 0) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic reads -> async writes
      const uint32_t row_bytes = (uint32_t)b.bw * PIX_BYTES;
      if (lane == 0) mbar_expect_tx(&bar, row_bytes * (uint32_t)b.bh);
      __syncwarp();
      if (lane < RT_ISSUE_LANES)
        for (int r = lane; r < b.bh; r += RT_ISSUE_LANES)
          bulk_g2s(box_mem + (size_t)r * row_bytes, img + ((size_t)(b.y0 + r) * w + b.x0) * FDIM, row_bytes, &bar);
    }
    fresh = b.bw * b.bh > 0;
    }
    const bool staged = b.bw * b.bh > 0;
#if RT_PERMUTE
    // Candidate (i, j) of a query sits at box pixel pix0 + j*bw + i, and a 16-byte chunk of pixel p lies in
    // bank group (3p + chunk) mod 8: the 8 lanes of a quarter-warp (one LDS.128 phase) are conflict-free
    // for ALL candidates of the level iff their pix0 differ mod 8.  The first level's centres are
    // nearly adjacent pixels (truncation skips / repeats some); afterwards they scatter (2.3 wavefronts per
    // phase measured), so the warp re-deals its 32 queries: the k-th query of residue class r goes to lane
    // 8k + r, the rest (5th+ of a class, dead queries) fill the empty slots from the top quarter down
    // (RT_DEAL=0) or are spread by class over the quarters that are not one-per-class (RT_DEAL=1, default).
    if ((RT_PERMUTE_FIRST || d < dilation_max) && !(split && pass)) {
      const unsigned lt = (1u << lane) - 1u;
      const int cls = live ? (((v0 - b.y0) * b.bw + (u0 - b.x0)) & 7) : 8;
#if RT_DEAL == 0
      int n_slot = 0, rank = 0;
#pragma unroll
      for (int r = 0; r < 8; r++) {
        const unsigned m = __ballot_sync(0xffffffffu, cls == r);
        if ((lane & 7) == r) n_slot = __popc(m);
        if (cls == r) rank = __popc(m & lt);
      }
      const unsigned empty = __ballot_sync(0xffffffffu, n_slot <= (lane >> 3));  // slot 8k + r is empty iff n_r <= k
      const bool spare = cls == 8 || rank >= 4;
      const unsigned spares = __ballot_sync(0xffffffffu, spare);
      const int dest = spare ? (int)__fns(empty, 31, -(__popc(spares & lt) + 1)) : (rank << 3 | cls);
#else
      // m = min(3, smallest class) quarters take exactly one query of every class (conflict-free); the other
      // queries, ordered by class (dead ones last), are dealt round-robin to the remaining 4 - m quarters, so
      // that a class with c members left puts at most ceil(c / (4 - m)) of them into one quarter
      // (simulation, uniform residues: 1.82 wavefronts per phase against 2.00 for RT_DEAL=0 and 2.60 undealt).
      int rank = 0, before = 0, nmin = 3;
#pragma unroll
      for (int r = 0; r < 9; r++) {
        const unsigned m = __ballot_sync(0xffffffffu, cls == r);
        const int nr = __popc(m);
        if (cls == r) rank = __popc(m & lt);
        if (cls > r) before += nr;  // queries of lower classes
        if (r < 8) nmin = min(nmin, nr);
      }
      int dest;
      if (cls < 8 && rank < nmin) dest = rank << 3 | cls;
      else {
        const int pos = before - nmin * min(cls, 8) + rank - (cls < 8 ? nmin : 0), nq = 4 - nmin;
        dest = ((nmin + pos % nq) << 3) + pos / nq;
      }
#endif
      lane_of_slot[warp][dest] = (unsigned char)lane;
      __syncwarp();
      const int src = lane_of_slot[warp][lane];
      __syncwarp();
#pragma unroll
      for (int k = 0; k < FDIM / 2; k++) {
        const unsigned x = __shfl_sync(0xffffffffu, *reinterpret_cast<unsigned *>(&q[k]), src);
        q[k] = *reinterpret_cast<const __half2 *>(&x);
      }
      u0 = __shfl_sync(0xffffffffu, u0, src);
      v0 = __shfl_sync(0xffffffffu, v0, src);
      max_score = __shfl_sync(0xffffffffu, max_score, src);
      qout = __shfl_sync(0xffffffffu, qout, src);
      live = __shfl_sync(0xffffffffu, (int)live, src) != 0;
      u_new = u0;
      v_new = v0;
    }
#endif
    if (fresh) {  // (the re-deal above needs only the box geometry: it overlaps the copy)
      mbar_wait(&bar, parity);
      parity ^= 1u;
    }
    if (live) {
      // window clipped to the image must lie inside the staged box to use shared memory
      const int wx0 = max(u0 + ilo, 0), wx1 = min(u0 + ihi, w - 1), wy0 = max(v0 - rd, 0), wy1 = min(v0 + rd, h - 1);
      const bool in_box = staged && wx0 >= b.x0 && wx1 < b.x0 + b.bw && wy0 >= b.y0 && wy1 < b.y0 + b.bh;
#if RT_STATS
      atomicAdd(&g_rt_stats[d & 7][0], 1ull);
      if (!in_box) atomicAdd(&g_rt_stats[d & 7][1], 1ull);
#endif
#if RT_LINEAR
      if (wx0 <= wx1 && wy0 <= wy1) {
        constexpr int NJ = 2 * RADIUS + 1;
        int uc[NJ], vc[NJ], roff[NJ], coff[NJ], best = -1;
        unsigned rmask = 0, cmask = 0;
#pragma unroll
        for (int k = 0; k < NJ; k++) {
          const int u = u0 + (k - RADIUS) * d, v = v0 + (k - RADIUS) * d;
          cmask |= (u >= 0 && u < w) ? 1u << k : 0u;
          rmask |= (v >= 0 && v < h) ? 1u << k : 0u;
          uc[k] = min(max(u, 0), w - 1);  // clamped coordinates lie in the clipped window, hence in the box
          vc[k] = min(max(v, 0), h - 1);
        }
        if (in_box) {
#pragma unroll
          for (int k = 0; k < NJ; k++) {
            roff[k] = (vc[k] - b.y0) * b.bw * (FDIM / 8);
            coff[k] = (uc[k] - b.x0) * (FDIM / 8);
          }
          const uint4 *base = reinterpret_cast<const uint4 *>(box_mem);
          auto load = [&](int off, int ch) { return base[off + ch]; };
          scan_level<FDIM, RADIUS>(q, roff, coff, rmask, cmask, d == dilation_max, load, max_score, best);
        } else {
#pragma unroll
          for (int k = 0; k < NJ; k++) {
            roff[k] = vc[k] * w * (FDIM / 8);  // h*w*FDIM/8 < 2^31: checked by the host
            coff[k] = uc[k] * (FDIM / 8);
          }
          const uint4 *base = reinterpret_cast<const uint4 *>(img);
          auto load = [&](int off, int ch) { return __ldg(base + off + ch); };
          scan_level<FDIM, RADIUS>(q, roff, coff, rmask, cmask, d == dilation_max, load, max_score, best);
        }
        if (best >= 0) {
          u_new = u0 + (best / NJ - RADIUS) * d;
          v_new = v0 + (best % NJ - RADIUS) * d;
        }
      }
#else
      if (wx0 <= wx1 && wy0 <= wy1) {
        if (in_box) {
          const uint4 *base = reinterpret_cast<const uint4 *>(box_mem);
          for (int i = ilo; i <= ihi; i += d) {
            const int u = u0 + i;
            if (u < 0 || u >= w) continue;
            const int col = (u - b.x0) * (FDIM / 8);
            auto load = [&](int v, int ch) { return base[(v - b.y0) * b.bw * (FDIM / 8) + col + ch]; };
            scan_column<FDIM, RADIUS>(q, u, v0, d, h, load, max_score, u_new, v_new);
          }
        } else {
          const uint4 *base = reinterpret_cast<const uint4 *>(img);
          for (int i = ilo; i <= ihi; i += d) {
            const int u = u0 + i;
            if (u < 0 || u >= w) continue;
            auto load = [&](int v, int ch) { return __ldg(base + ((size_t)v * w + u) * (FDIM / 8) + ch); };
            scan_column<FDIM, RADIUS>(q, u, v0, d, h, load, max_score, u_new, v_new);
          }
        }
      }
#endif
    }
    if (split && !pass) continue;  // second half of this level next, same centres
    u0 = u_new;
    v0 = v_new;
    d--;
    pass = -1;
  }
  if (live) {
    const size_t bo = (size_t)blockIdx.y * n + qout;
    if (lin_out) p1_new[bo] = u_new + (long long)w * v_new;  // pixel_to_lin (matching.py:13-15) folded in
    else reinterpret_cast<longlong2 *>(p1_new)[bo] = make_longlong2(u_new, v_new);
  }
}

}  // namespace rtile
