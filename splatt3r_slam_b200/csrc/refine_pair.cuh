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
        S[p] = __hadd2(S[p], __hmul2(qk, *reinterpret_cast<const __half2 *>(&wv)));
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

}  // namespace rpair
