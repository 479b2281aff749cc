// b200slam — refine_matches, pair-interleaved tiled kernel (second generation of refine_tile.cuh).
//
// Same arithmetic, scan order and tie-breaking as matching_kernels.cu:24-80 (bit-exact), same
// staging (one cp.async.bulk per box row, mbarrier completion).  What changes is the layout of the
// staged box, so that the sequential fp16 dot products of TWO candidates run in the two lanes of
// one HMUL2 + one HADD2 with no per-use shuffles:
//
//   ncu on the first tiled kernel showed 130 M warp instructions for 196 608 queries — 86 per
//   candidate, of which ~24 are PRMT byte-permutes: HMUL2 naturally yields (q_k c_k, q_k+1 c_k+1)
//   of ONE candidate, but the rounding order forces a strictly sequential sum over k, so the two
//   lanes of an HADD2 can only be two different CANDIDATES, and building (cA_k, cB_k) operands
//   costs a permute per use.  Here the permute is done ONCE per staged pixel instead of once per
//   use (each pixel is used by ~49 candidates): after the bulk copy lands, the CTA interleaves, in
//   place, rows r and r+d of the box (d = dilation of this level; rows grouped in blocks of d,
//   even block paired with the next odd block): slot r holds half2 (A_k, B_k) for k = 0..11, slot
//   r+d for k = 12..23.  A query's 7 candidates of one column (rows r0, r0+d, ..., r0+6d) are then
//   covered by exactly 4 pair slots whatever the parity of r0's block, and one candidate pair
//   costs 6 LDS.128 + 24 HMUL2 (query channel broadcast by operand swizzle) + 24 HADD2.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "refine_tile.cuh"

namespace rpair {

using rtile::TW;
using rtile::TH;
using rtile::THREADS;
constexpr int BOX_BYTES = 115200;  // 2400 staged pixels of 48 B; two CTAs per SM (2 x ~113 KB)

struct PBox {
  int x0, Y0, bw, bh;  // x0 clipped to the image; Y0 VIRTUAL (may be < 0); bh a multiple of 2d
};

__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
  return r;
}

// In-place interleave of box rows (rA, rA + d) at column x: see the header comment.
__device__ __forceinline__ void interleave_pair(uint4 *base, int bw, int rA, int d, int x) {
  uint4 *pa = base + ((size_t)rA * bw + x) * 3, *pb = base + ((size_t)(rA + d) * bw + x) * 3;
  const uint4 a0 = pa[0], a1 = pa[1], a2 = pa[2], b0 = pb[0], b1 = pb[1], b2 = pb[2];
  const uint32_t A[12] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w, a2.x, a2.y, a2.z, a2.w};
  const uint32_t B[12] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w, b2.x, b2.y, b2.z, b2.w};
  uint32_t o[24];  // o[k] = half2 (A_k, B_k)
#pragma unroll
  for (int t = 0; t < 12; t++) {
    o[2 * t] = prmt(A[t], B[t], 0x5410);
    o[2 * t + 1] = prmt(A[t], B[t], 0x7632);
  }
  pa[0] = make_uint4(o[0], o[1], o[2], o[3]);
  pa[1] = make_uint4(o[4], o[5], o[6], o[7]);
  pa[2] = make_uint4(o[8], o[9], o[10], o[11]);
  pb[0] = make_uint4(o[12], o[13], o[14], o[15]);
  pb[1] = make_uint4(o[16], o[17], o[18], o[19]);
  pb[2] = make_uint4(o[20], o[21], o[22], o[23]);
}

// One column (fixed u) of 7 candidates for one query, from the interleaved box.
// r0 = box row of candidate j = 0 (>= 0), col3 = 3 * (u - x0).  q2[t] = (q_2t, q_2t+1).
template <int RADIUS>
__device__ __forceinline__ void scan_column_pairs(const __half2 (&q2)[12], const uint4 *__restrict__ base, int bw,
                                                  int r0, int col3, int d, int u, int v0, int h,
                                                  float &max_score, int &u_new, int &v_new) {
  static_assert(RADIUS == 3, "7 candidates per column = 4 pair slots");
  const int m0 = r0 / d, o = r0 - m0 * d, par = m0 & 1;
  const int rowA0 = (m0 - par) * d + o;  // slot A of pair 0; pair p: + 2*p*d; slot B: + d
  __half2 S[4];
#pragma unroll
  for (int p = 0; p < 4; p++) S[p] = __half2(__ushort_as_half(0), __ushort_as_half(0));
#pragma unroll
  for (int c = 0; c < 6; c++) {
    uint4 W[4];
#pragma unroll
    for (int p = 0; p < 4; p++) {
      const int row = rowA0 + 2 * p * d + (c >= 3 ? d : 0);
      W[p] = base[(size_t)row * bw * 3 + col3 + (c % 3)];
    }
#pragma unroll
    for (int t = 0; t < 4; t++) {
      const int k = 4 * c + t;
      const __half2 qk = (k & 1) ? __high2half2(q2[k >> 1]) : __low2half2(q2[k >> 1]);
#pragma unroll
      for (int p = 0; p < 4; p++) {
        const uint32_t wv = t == 0 ? W[p].x : t == 1 ? W[p].y : t == 2 ? W[p].z : W[p].w;
        S[p] = __hadd2_rn(S[p], __hmul2_rn(qk, *reinterpret_cast<const __half2 *>(&wv)));  // _rn: never contracted into HFMA2
      }
    }
  }
  // reference scan order within the column: j = 0..6 (v ascending), strict >
#pragma unroll
  for (int p = 0; p < 4; p++) {
#pragma unroll
    for (int l = 0; l < 2; l++) {
      const int j = 2 * p + l - par;
      const int vv = v0 + (j - RADIUS) * d;
      const float sf = l == 0 ? __low2float(S[p]) : __high2float(S[p]);
      if (j >= 0 && j <= 2 * RADIUS && vv >= 0 && vv < h && sf > max_score) {
        max_score = sf;
        u_new = u;
        v_new = vv;
      }
    }
  }
}


template <int FDIM, int RADIUS>
__global__ void __launch_bounds__(THREADS, 2)
refine_pair_kernel(const __half *__restrict__ D11, const __half *__restrict__ D21,
                   const int64_t *__restrict__ p1, int64_t *__restrict__ p1_new, int h, int w, int n,
                   int qw, int dilation_max) {
  static_assert(FDIM == 24, "interleaved layout is written for 24 channels (3 x 16 B per pixel)");
  extern __shared__ __align__(128) unsigned char box_mem[];
  __shared__ uint64_t bar;
  __shared__ int red[THREADS / 32][7];
  __shared__ PBox sbox;
  constexpr int PIX_BYTES = FDIM * 2;
  constexpr int MAX_PX = BOX_BYTES / PIX_BYTES;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tiles_x = (qw + TW - 1) / TW;
  const int qx = (blockIdx.x % tiles_x) * TW + (tid % TW);
  const int qy = (blockIdx.x / tiles_x) * TH + (tid / TW);
  const long long qlin = (long long)qy * qw + qx;
  const bool active = qx < qw && qlin < n;
  const size_t bq = (size_t)blockIdx.y * n + (active ? qlin : 0);
  const __half *img = D11 + (size_t)blockIdx.y * h * w * FDIM;

  if (tid == 0) rtile::mbar_init(&bar, 1);

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
  const long long reach = (long long)RADIUS * dilation_max;
  const bool can_hit = pin.x >= -reach && pin.x < w + reach && pin.y >= -reach && pin.y < h + reach;
  int u0 = can_hit ? (int)pin.x : -(1 << 20), v0 = can_hit ? (int)pin.y : -(1 << 20);
  int u_new = u0, v_new = v0;
  float max_score = 0.0f;  // numeric_limits<c10::Half>::min() == 0 in the reference build
  uint32_t parity = 0;
  __syncthreads();  // mbarrier initialised

  for (int d = dilation_max; d > 0; d--) {
    const int rd = RADIUS * d;
    // ---- bounding box of the centres of this tile ------------------------------------------------
    const bool part = active && can_hit;
    const int BIG = 1 << 29;
    int vals[4] = {part ? u0 : BIG, part ? -u0 : BIG, part ? v0 : BIG, part ? -v0 : BIG};
#pragma unroll
    for (int k = 0; k < 4; k++) vals[k] = __reduce_min_sync(0xffffffffu, vals[k]);
    const int su_w = __reduce_add_sync(0xffffffffu, part ? u0 : 0);
    const int sv_w = __reduce_add_sync(0xffffffffu, part ? v0 : 0);
    const int cnt_w = __popc(__ballot_sync(0xffffffffu, part));
    if (lane == 0) {
#pragma unroll
      for (int k = 0; k < 4; k++) red[warp][k] = vals[k];
      red[warp][4] = su_w; red[warp][5] = sv_w; red[warp][6] = cnt_w;
    }
    __syncthreads();  // also: every thread has finished reading the previous level's box
    if (tid == 0) {
      int umin = BIG, numax = BIG, vmin = BIG, nvmax = BIG, su = 0, sv = 0, c = 0;
      for (int k = 0; k < THREADS / 32; k++) {
        umin = min(umin, red[k][0]); numax = min(numax, red[k][1]);
        vmin = min(vmin, red[k][2]); nvmax = min(nvmax, red[k][3]);
        su += red[k][4]; sv += red[k][5]; c += red[k][6];
      }
      PBox b{0, 0, 0, 0};
      if (c > 0) {
        int x0 = max(umin - rd, 0), x1 = min(-numax + rd, w - 1);
        int Y0 = vmin - rd;                                   // virtual: rows outside the image stay unfilled
        int bh = ((-nvmax + rd - Y0 + 1) + 2 * d - 1) / (2 * d) * (2 * d);
        if (x1 >= x0) {
          if ((long long)(x1 - x0 + 1) * bh > MAX_PX) {
            // incoherent tile: a box of the nominal size around the mean centre; queries whose
            // window leaves it take the global-memory path for this level
            const int side_x = min(TW + 2 * rd + 2, w);
            bh = min((TH + 2 * rd + 2 * d - 1) / (2 * d) * (2 * d), MAX_PX / side_x / (2 * d) * (2 * d));
            const int cx = su / c, cy = sv / c;
            x0 = min(max(cx - side_x / 2, 0), w - side_x); x1 = x0 + side_x - 1;
            Y0 = cy - bh / 2;
          }
          b = PBox{x0, Y0, x1 - x0 + 1, bh};
        }
      }
      // rows of the box that exist in the image
      const int r_lo = max(0, -b.Y0), r_hi = min(b.bh, h - b.Y0);  // [r_lo, r_hi)
      if (b.bw <= 0 || r_hi <= r_lo) b = PBox{0, 0, 0, 0};
      sbox = b;
      if (b.bw > 0) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic accesses -> async writes
        const uint32_t row_bytes = (uint32_t)b.bw * PIX_BYTES;
        rtile::mbar_expect_tx(&bar, row_bytes * (uint32_t)(r_hi - r_lo));
        for (int r = r_lo; r < r_hi; r++)
          rtile::bulk_g2s(box_mem + (size_t)r * row_bytes, img + ((size_t)(b.Y0 + r) * w + b.x0) * FDIM, row_bytes, &bar);
      }
    }
    __syncthreads();
    const PBox b = sbox;
    const bool staged = b.bw > 0;
    uint4 *base = reinterpret_cast<uint4 *>(box_mem);
    if (staged) {
      rtile::mbar_wait(&bar, parity);
      parity ^= 1u;
      // in-place pair interleave (rows outside the image hold stale bytes: their lanes are never read out)
      const int pairs = (b.bh >> 1) * b.bw;
      for (int t = tid; t < pairs; t += THREADS) {
        const int pr = t / b.bw, x = t - pr * b.bw;
        const int m2 = pr / d, o = pr - m2 * d;
        interleave_pair(base, b.bw, 2 * m2 * d + o, d, x);
      }
      __syncthreads();
    }
    if (active && can_hit) {
      const int wx0 = max(u0 - rd, 0), wx1 = min(u0 + rd, w - 1);
      const bool in_box = staged && wx0 >= b.x0 && wx1 < b.x0 + b.bw && v0 - rd >= b.Y0 && v0 + rd < b.Y0 + b.bh;
      if (wx0 <= wx1 && v0 + rd >= 0 && v0 - rd < h) {
        if (in_box) {
          const int r0 = v0 - rd - b.Y0;
          for (int i = -rd; i <= rd; i += d) {
            const int u = u0 + i;
            if (u < 0 || u >= w) continue;
            scan_column_pairs<RADIUS>(q, base, b.bw, r0, 3 * (u - b.x0), d, u, v0, h, max_score, u_new, v_new);
          }
        } else {
          const uint4 *gbase = reinterpret_cast<const uint4 *>(img);
          for (int i = -rd; i <= rd; i += d) {
            const int u = u0 + i;
            if (u < 0 || u >= w) continue;
            auto load = [&](int v, int ch) { return __ldg(gbase + ((size_t)v * w + u) * (FDIM / 8) + ch); };
            rtile::scan_column<FDIM, RADIUS>(q, u, v0, d, h, load, max_score, u_new, v_new);
          }
        }
      }
    }
    u0 = u_new;
    v0 = v_new;
  }
  if (active) {
    const longlong2 out = can_hit ? make_longlong2(u_new, v_new) : pin;
    reinterpret_cast<longlong2 *>(p1_new)[bq] = out;
  }
}

}  // namespace rpair
