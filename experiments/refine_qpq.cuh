// MEASURED DEAD END (experiments/README.md): bit-exact, conflict-free, and slower than refine_tile_kernel.  Built only by
// `tools/build_ab.sh qpq -DRT_QPQ=1`.
//
// b200slam — refine_matches, "quarter-warp per query pair" kernel: the tiled kernel of refine_tile.cuh with a data
// path that is free of shared-memory bank conflicts BY CONSTRUCTION, whatever the match centres do.
//
// refine_tile_kernel gives every lane one query; after the first dilation level the centres scatter and the eight
// lanes of a quarter-warp (one LDS.128 wavefront) hit random 16-byte bank groups: 1.8 wavefronts per ideal one even
// after re-dealing the queries of a warp to its lanes by bank residue (ncu r02e: 33.6 M wavefronts, 14.7 M of them
// conflicts; waiting for LDS is the top stall).  Here a QUARTER-WARP owns a PAIR of queries (A, B) at a time:
//   * lane j (0..6) of the quarter scores ROW j of both 7x7 windows, one column at a time: the seven lanes of one
//     LDS.128 read pixels (u, v + (j-3) d) of the same column in seven different box rows.  Every box row is staged
//     by its own bulk copy, so its base is free: row y starts at a 128-byte multiple plus 16 * (y mod 7) bytes.
//     Rows v + (j-3) d, j = 0..6, are distinct mod 7 for every d that is not a multiple of 7 (7 is prime), so the
//     seven lanes always hit seven different bank groups: exactly one wavefront per quarter, for every dilation
//     level, every centre and also for a box that serves several levels.  The eighth lane idles.
//   * the two candidates a lane scores at a time are A's and B's pixel of the same (column, row): the packed HMUL2 /
//     HADD2 carry (A, B) in their two halves with the query operand (qA_k, qB_k) pre-transposed once per tile in
//     shared memory, so no half of a packed instruction idles on the odd seventh candidate of a column.
//   * the argmax in the reference's scan order (u outer, v inner, strict >) is a quarter-wide integer maximum of
//     (score bits << 6 | 63 - scan index): scores that can win are positive fp16 values, whose bit patterns order
//     like the values; equal scores resolve to the smaller scan index, as the sequential strict > does.
// Same arithmetic per candidate (sequential fp16-rounded MACs) and same results as refine_f16_kernel; queries stay in
// their home lanes (state travels by shuffle), so there is no re-deal and no block barrier beyond the box staging.
#pragma once
#include "../splatt3r_slam_b200/csrc/refine_tile.cuh"

namespace rtile {

#ifndef RT_QPQ_BOX_KB
#define RT_QPQ_BOX_KB 200  // box bytes; the rest of the 227 KB holds the transposed query pairs (24 KB)
#endif
constexpr int QPQ_BOX_BYTES = RT_QPQ_BOX_KB * 1024;

// row pitch in 16-byte units: pixels of a row + up to 6 units of (y mod 7) offset, rounded to a multiple of 8 units
__device__ __host__ __forceinline__ int qpq_pitch16(int bw, int fdim8) { return (bw * fdim8 + 6 + 7) & ~7; }

template <int FDIM, int RADIUS>
__global__ void __launch_bounds__(THREADS, 1)
refine_qpq_kernel(const __half *__restrict__ D11, const __half *__restrict__ D21, const int64_t *__restrict__ p1,
                  int64_t *__restrict__ p1_new, int h, int w, int n, int qw, int bands, int dilation_max, int lin_out) {
  extern __shared__ __align__(128) unsigned char box_mem[];  // [box QPQ_BOX_BYTES | qT: THREADS/2 pairs x FDIM half2]
  __shared__ uint64_t bar;
  __shared__ int red[2][THREADS / 32][8];
  constexpr int C8 = FDIM / 8;  // 16-byte chunks per pixel
  constexpr int NJ = 2 * RADIUS + 1;
  static_assert(NJ == 7, "one row of the window per lane of a quarter-warp, rows distinct mod 7");
  uint4 *qT = reinterpret_cast<uint4 *>(box_mem + QPQ_BOX_BYTES);  // pair p: FDIM half2 = FDIM/4 uint4

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tiles_x = (qw + TW - 1) / TW;
  const int qx = (blockIdx.x % tiles_x) * TW + (tid % TW);
  const int band = blockIdx.x / tiles_x, qh = n / qw;
  const int qy0 = (int)((long long)band * qh / bands), qy1 = (int)((long long)(band + 1) * qh / bands);
  const int qy = qy0 + (tid / TW);
  const long long qlin = (long long)qy * qw + qx;
  const bool active = qx < qw && qy < qy1;
  const size_t bq = (size_t)blockIdx.y * n + (active ? qlin : 0);
  const __half *img = D11 + (size_t)blockIdx.y * h * w * FDIM;

  if (tid == 0) mbar_init(&bar, 1);

  // ---- query descriptors: transposed per pair (even lane A, odd lane B) into shared memory, once per tile --------
  {
    uint32_t mine[FDIM / 2], other[FDIM / 2];
    const uint4 *src = reinterpret_cast<const uint4 *>(D21 + bq * FDIM);
#pragma unroll
    for (int v = 0; v < C8; v++) {
      const uint4 c = __ldg(src + v);
      mine[v * 4 + 0] = c.x; mine[v * 4 + 1] = c.y; mine[v * 4 + 2] = c.z; mine[v * 4 + 3] = c.w;
    }
#pragma unroll
    for (int m = 0; m < FDIM / 2; m++) other[m] = __shfl_xor_sync(0xffffffffu, mine[m], 1);
    // word m of a query = channels (2m, 2m+1); qT[2m] = (A_2m, B_2m), qT[2m+1] = (A_2m+1, B_2m+1).
    // The even lane writes the first half of the pair's FDIM words, the odd lane the second half.
    const bool odd = lane & 1;
    uint32_t out[FDIM / 2];
#pragma unroll
    for (int m = 0; m < FDIM / 4; m++) {
      const uint32_t a_lo = odd ? other[m + FDIM / 4] : mine[m];            // A's word
      const uint32_t b_lo = odd ? mine[m + FDIM / 4] : other[m];            // B's word
      out[2 * m] = __byte_perm(a_lo, b_lo, 0x5410);      // (A.lo, B.lo)
      out[2 * m + 1] = __byte_perm(a_lo, b_lo, 0x7632);  // (A.hi, B.hi)
    }
    uint4 *dst = qT + (size_t)(tid >> 1) * (FDIM / 4) + (odd ? FDIM / 8 : 0);
#pragma unroll
    for (int v = 0; v < FDIM / 8; v++) dst[v] = make_uint4(out[4 * v], out[4 * v + 1], out[4 * v + 2], out[4 * v + 3]);
  }
  const longlong2 pin = __ldg(reinterpret_cast<const longlong2 *>(p1) + bq);
  const long long reach = (long long)RADIUS * dilation_max;
  const bool can_hit = pin.x >= -reach && pin.x < w + reach && pin.y >= -reach && pin.y < h + reach;
  int u0 = can_hit ? (int)pin.x : -(1 << 14), v0 = can_hit ? (int)pin.y : -(1 << 14);
  unsigned ms_bits = 0;  // max_score as fp16 bits (0.0 seed: numeric_limits<c10::Half>::min() == 0 in the reference build)
  uint32_t parity = 0;
  int buf = 0;
  if (active && !can_hit) {
    if (lin_out) p1_new[bq] = pin.x + (long long)w * pin.y;
    else reinterpret_cast<longlong2 *>(p1_new)[bq] = pin;
  }
  const bool live = active && can_hit;
  __syncthreads();  // mbarrier initialised, qT written

  const int quarter = lane >> 3, j = lane & 7;  // j: window row of this lane (7 = idle)
  const unsigned qmask = 0xffu << (quarter * 8);
  Box b{0, 0, 0, 0};
  int pitch = 0;  // row pitch of the staged box, 16-byte units
  bool merged = false;
  for (int d = dilation_max; d > 0; d--) {
    const int rd = RADIUS * d;
    bool fresh = false;
    if (!merged) {
      int vals[7] = {live ? u0 : (1 << 29), live ? -u0 : (1 << 29), live ? v0 : (1 << 29),
                     live ? -v0 : (1 << 29), live ? u0 : 0, live ? v0 : 0, live ? 1 : 0};
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
      __syncthreads();  // also: every thread has finished reading the previous level's box
      {
        const bool has = lane < THREADS / 32;
#pragma unroll
        for (int k = 0; k < 4; k++) vals[k] = __reduce_min_sync(0xffffffffu, has ? redp[lane][k] : (1 << 29));
#pragma unroll
        for (int k = 4; k < 7; k++) vals[k] = __reduce_add_sync(0xffffffffu, has ? redp[lane][k] : 0);
      }
      b = Box{0, 0, 0, 0};
      auto fits = [&](int bw, int bh) { return (long long)qpq_pitch16(bw, C8) * 16 * bh <= QPQ_BOX_BYTES; };
      {
        const int umin = vals[0], numax = vals[1], vmin = vals[2], nvmax = vals[3], su = vals[4], sv = vals[5], c = vals[6];
        if (c > 0) {
          const int rall = RADIUS * d * (d + 1) / 2;
          const int x0 = max(umin - rall, 0), x1 = min(-numax + rall, w - 1);
          const int y0 = max(vmin - rall, 0), y1 = min(-nvmax + rall, h - 1);
          if (x1 >= x0 && y1 >= y0 && fits(x1 - x0 + 1, y1 - y0 + 1)) {
            b = Box{x0, y0, x1 - x0 + 1, y1 - y0 + 1};
            merged = true;
          }
        }
        if (c > 0 && !merged) {
          int x0 = max(umin - rd, 0), x1 = min(-numax + rd, w - 1);
          int y0 = max(vmin - rd, 0), y1 = min(-nvmax + rd, h - 1);
          if (x1 >= x0 && y1 >= y0) {
            if (!fits(x1 - x0 + 1, y1 - y0 + 1)) {
              // incoherent tile: the largest box of the bounding box's shape that fits, around the mean centre
              const float shrink = sqrtf((float)(QPQ_BOX_BYTES / (FDIM * 2)) / (float)((x1 - x0 + 1) * (y1 - y0 + 1)));
              const int side_x = min(max(min((int)((x1 - x0 + 1) * shrink) - 2, w), min(TW + 2 * rd + 2, w)), x1 - x0 + 1);
              const int side_y = max(min(min(y1 - y0 + 1, h), QPQ_BOX_BYTES / (qpq_pitch16(side_x, C8) * 16)), 1);
              const int cx = su / c, cy = sv / c;
              x0 = min(max(cx - side_x / 2, 0), w - side_x); x1 = x0 + side_x - 1;
              y0 = min(max(cy - side_y / 2, 0), h - side_y); y1 = y0 + side_y - 1;
            }
            b = Box{x0, y0, x1 - x0 + 1, y1 - y0 + 1};
          }
        }
      }
      pitch = qpq_pitch16(b.bw, C8);
      if (warp == 0 && b.bw * b.bh > 0) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic reads -> async writes
        const uint32_t row_bytes = (uint32_t)b.bw * FDIM * 2;
        if (lane == 0) mbar_expect_tx(&bar, row_bytes * (uint32_t)b.bh);
        __syncwarp();
        for (int r = lane; r < b.bh; r += 32)
          bulk_g2s(box_mem + ((size_t)r * pitch + (unsigned)(b.y0 + r) % 7u) * 16,
                   img + ((size_t)(b.y0 + r) * w + b.x0) * FDIM, row_bytes, &bar);
      }
      fresh = b.bw * b.bh > 0;
    }
    const bool staged = b.bw * b.bh > 0;
    // home lane: is this query's window (clipped to the image) non-empty, inside the staged box, inside the image?
    const int wx0 = max(u0 - rd, 0), wx1 = min(u0 + rd, w - 1), wy0 = max(v0 - rd, 0), wy1 = min(v0 + rd, h - 1);
    const bool work = live && wx0 <= wx1 && wy0 <= wy1;
    const bool in_box = staged && wx0 >= b.x0 && wx1 < b.x0 + b.bw && wy0 >= b.y0 && wy1 < b.y0 + b.bh;
    const bool interior = u0 - rd >= 0 && u0 + rd < w && v0 - rd >= 0 && v0 + rd < h;
    // state that travels to the quarter: centre (16 + 16 bits, biased) and score bits | flags
    const unsigned st_uv = ((unsigned)(u0 + 16384) << 16) | (unsigned)(v0 + 16384);
    const unsigned st_ms = ms_bits | (work ? 1u << 16 : 0u) | (in_box ? 1u << 17 : 0u) | (work && in_box && interior ? 1u << 18 : 0u);
    if (fresh) {
      mbar_wait(&bar, parity);
      parity ^= 1u;
    }
    const int jj = min(j, NJ - 1);  // the idle eighth lane shadows lane 6 (same addresses: broadcast)
    const int dj = (jj - RADIUS) * d, step = d * C8;
    unsigned my_key = 0;
#pragma unroll 1
    for (int r = 0; r < 4; r++) {
      const int srcA = r * 8 + quarter * 2;  // home lanes of the pair this quarter scores in round r
      const unsigned uvA = __shfl_sync(0xffffffffu, st_uv, srcA), msA = __shfl_sync(0xffffffffu, st_ms, srcA);
      const unsigned uvB = __shfl_sync(0xffffffffu, st_uv, srcA + 1), msB = __shfl_sync(0xffffffffu, st_ms, srcA + 1);
      unsigned keyA = 0, keyB = 0;
      const bool workA = (msA >> 16) & 1u, workB = (msB >> 16) & 1u;
      if (workA || workB) {
        const uint4 *qp = qT + (size_t)(warp * 16 + r * 4 + quarter) * (FDIM / 4);
        int offA[NJ], offB[NJ];
        __half2 s2[NJ];
#pragma unroll
        for (int i = 0; i < NJ; i++) s2[i] = __half2half2(__ushort_as_half(0));
        // sequential fp16-rounded MACs of (A, B) x 7 columns over the FDIM channels; load(off) = one 16-byte chunk
        auto run = [&](auto load) {
#pragma unroll
          for (int ch = 0; ch < C8; ch++) {
            uint4 cA[NJ], cB[NJ];
#pragma unroll
            for (int i = 0; i < NJ; i++) {
              cA[i] = load(offA[i] + ch);
              cB[i] = load(offB[i] + ch);
            }
            const uint4 qa = qp[2 * ch], qb = qp[2 * ch + 1];
            const uint32_t qw8[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};  // (A_k, B_k), k = 8 ch + 0..7
#pragma unroll
            for (int t = 0; t < 4; t++) {
              const __half2 q_lo = *reinterpret_cast<const __half2 *>(&qw8[2 * t]);
              const __half2 q_hi = *reinterpret_cast<const __half2 *>(&qw8[2 * t + 1]);
#pragma unroll
              for (int i = 0; i < NJ; i++) {
                const uint32_t wa = t == 0 ? cA[i].x : t == 1 ? cA[i].y : t == 2 ? cA[i].z : cA[i].w;
                const uint32_t wb = t == 0 ? cB[i].x : t == 1 ? cB[i].y : t == 2 ? cB[i].z : cB[i].w;
                const __half2 a = *reinterpret_cast<const __half2 *>(&wa), bb = *reinterpret_cast<const __half2 *>(&wb);
                s2[i] = __hadd2_rn(s2[i], __hmul2_rn(q_lo, __lows2half2(a, bb)));
                s2[i] = __hadd2_rn(s2[i], __hmul2_rn(q_hi, __highs2half2(a, bb)));
              }
            }
          }
        };
        unsigned killA = 0, killB = 0;  // columns (bit i) whose candidates must not win; 0x7f = the whole row
        if ((msA >> 18) & (msB >> 18) & 1u) {
          // fast path: both windows inside the image and inside the staged box - affine addresses, nothing to mask
          const int uA = (int)(uvA >> 16) - 16384, yA = (int)(uvA & 0xffffu) - 16384 + dj;
          const int uB = (int)(uvB >> 16) - 16384, yB = (int)(uvB & 0xffffu) - 16384 + dj;
          const int baseA = (yA - b.y0) * pitch + (int)((unsigned)yA % 7u) + (uA - rd - b.x0) * C8;
          const int baseB = (yB - b.y0) * pitch + (int)((unsigned)yB % 7u) + (uB - rd - b.x0) * C8;
#pragma unroll
          for (int i = 0; i < NJ; i++) {
            offA[i] = baseA + i * step;
            offB[i] = baseB + i * step;
          }
          const uint4 *base = reinterpret_cast<const uint4 *>(box_mem);
          run([&](int off) { return base[off]; });
        } else {
          // general path: image borders (rows / columns clamped and masked), windows outside the box (global
          // loads), a dead partner (shadows the live query; its result is dropped)
          const unsigned uvA2 = workA ? uvA : uvB, uvB2 = workB ? uvB : uvA;
          const int uA = (int)(uvA2 >> 16) - 16384, yA = (int)(uvA2 & 0xffffu) - 16384 + dj;
          const int uB = (int)(uvB2 >> 16) - 16384, yB = (int)(uvB2 & 0xffffu) - 16384 + dj;
          const bool boxed = ((workA ? msA : msB) >> 17) & ((workB ? msB : msA) >> 17) & 1u;
          const int yAc = min(max(yA, 0), h - 1), yBc = min(max(yB, 0), h - 1);
          const int rowA = boxed ? (yAc - b.y0) * pitch + (int)((unsigned)yAc % 7u) - b.x0 * C8 : yAc * w * C8;
          const int rowB = boxed ? (yBc - b.y0) * pitch + (int)((unsigned)yBc % 7u) - b.x0 * C8 : yBc * w * C8;
#pragma unroll
          for (int i = 0; i < NJ; i++) {
            const int xA = uA + (i - RADIUS) * d, xB = uB + (i - RADIUS) * d;
            killA |= (xA >= 0 && xA < w) ? 0u : 1u << i;
            killB |= (xB >= 0 && xB < w) ? 0u : 1u << i;
            offA[i] = rowA + min(max(xA, 0), w - 1) * C8;  // h*w*C8 < 2^31: checked by the host
            offB[i] = rowB + min(max(xB, 0), w - 1) * C8;
          }
          if (!workA || yA < 0 || yA >= h) killA = 0x7fu;
          if (!workB || yB < 0 || yB >= h) killB = 0x7fu;
          const uint4 *sbase = reinterpret_cast<const uint4 *>(box_mem), *gbase = reinterpret_cast<const uint4 *>(img);
          run([&](int off) { return boxed ? sbase[off] : __ldg(gbase + off); });
        }
        if (j >= NJ) killA = killB = 0x7fu;
        // Lane-local argmax over the 7 columns of this row, A and B at once in the two halves: ordered fp16
        // compares against the score carried from the coarser levels (NaN never wins, strict > keeps the first
        // column among equals), then the quarter's maximum of (score bits << 6 | 63 - scan index).
        const unsigned carried = (msA & 0xffffu) | (msB << 16);
        unsigned best = carried, bidx = 0x00070007u;  // column of the best so far, 7 = none
#pragma unroll
        for (int i = 0; i < NJ; i++) {
          unsigned x = *reinterpret_cast<const unsigned *>(&s2[i]);
          if (killA | killB)  // (uniformly zero on the fast path)
            x = (((killA >> i) & 1u) ? 0xfc00u : (x & 0xffffu)) | (((killB >> i) & 1u) ? 0xfc000000u : (x & 0xffff0000u));
          const unsigned g = __hgt2_mask(*reinterpret_cast<const __half2 *>(&x), *reinterpret_cast<const __half2 *>(&best));
          best = (best & ~g) | (x & g);
          bidx = (bidx & ~g) | ((unsigned)i * 0x00010001u & g);
        }
        const unsigned iA = bidx & 0xffffu, iB = bidx >> 16;
        keyA = iA < (unsigned)NJ ? ((best & 0xffffu) << 6 | (63u - (NJ * iA + (unsigned)jj))) : 0u;
        keyB = iB < (unsigned)NJ ? ((best >> 16) << 6 | (63u - (NJ * iB + (unsigned)jj))) : 0u;
      }
      keyA = __reduce_max_sync(qmask, keyA);
      keyB = __reduce_max_sync(qmask, keyB);
      // back to the home lanes: lane l was scored in round l / 8 by quarter (l % 8) / 2, as A if even
      const unsigned gotA = __shfl_sync(0xffffffffu, keyA, ((lane & 7) >> 1) * 8);
      const unsigned gotB = __shfl_sync(0xffffffffu, keyB, ((lane & 7) >> 1) * 8);
      if ((lane >> 3) == r) my_key = (lane & 1) ? gotB : gotA;
    }
    if (my_key) {  // some candidate beat the score carried from the coarser levels (strict >)
      ms_bits = my_key >> 6;
      const int t = 63 - (int)(my_key & 63u);
      u0 += (t / NJ - RADIUS) * d;
      v0 += (t % NJ - RADIUS) * d;
    }
  }
  if (live) {
    const size_t bo = (size_t)blockIdx.y * n + (size_t)qlin;
    if (lin_out) p1_new[bo] = u0 + (long long)w * v0;
    else reinterpret_cast<longlong2 *>(p1_new)[bo] = make_longlong2(u0, v0);
  }
}

}  // namespace rtile
