// b200slam — forward 3D-Gaussian-splat tile rasteriser for sm_100a (last item: the
// CUDA compute part of the reference's rendering path; no graphics API involved).
//
// Replaces the FORWARD pass of the reference's vendored rasteriser, thirdparty/diff-gaussian-rasterization-modified:
// preprocessCUDA (cuda_rasterizer/forward.cu:157-260), duplicateWithKeys / identifyTileRanges and the sort
// (rasterizer_impl.cu:78-150, :195-330), renderCUDA (forward.cu:262-370), computeCov3D (:118-153), computeColorFromSH
// in the e3nn axis convention of that fork (:20-72).  The SLAM system only renders (visualization.py:563-596,
// pixelsplat_src/cuda_splatting.py:96-124); the backward pass (training of the network) is out of scope.
//
// Same image: the same 16 x 16 pixel tiles decide which splats a pixel sees (a splat's 3-sigma rectangle in tiles),
// the same near-plane cull (z_view <= 0.2), low-pass (+0.3 px^2), alpha clamp (0.99), 1/255 and T < 1e-4 cut-offs,
// front-to-back order by view depth with ties in splat-index order.  Different pipeline:
//   * the reference sorts every (tile, splat) instance by a 64-bit (tile | depth) key — 32 + log2(tiles) key bits
//     over P' instances (P' = 4-10 x the number of splats).  Here the P splats are sorted ONCE by depth (32-bit
//     keys), instances are emitted in that order, and one STABLE sort on the log2(tiles) tile bits brings them into
//     (tile, depth) order: ~5x fewer bytes through the radix passes.  (cub::DeviceRadixSort / DeviceScan from the
//     CUDA toolkit do the passes — library code, as in the reference.)
//   * per-splat screen data is packed once into 3 x float4 with the conic pre-scaled (power is three FMAs) and two
//     conservative culling aids; the render kernel stages it per batch with 16-byte loads, lets every warp pick the
//     splats that can reach its 8 x 4 pixel block with one ballot per 32 splats, and skips expf below the cut-off.
#include <cub/cub.cuh>
#include <thrust/iterator/counting_iterator.h>
#include <thrust/iterator/transform_iterator.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200slam.h"

namespace b200gs {

constexpr int TILE = 16;
constexpr int RENDER_THREADS = TILE * TILE;

struct Camera {
  float view[16], proj[16], campos[3], bg[3];
};

__device__ __forceinline__ float ndc2pix(float v, int S) { return ((v + 1.0f) * S - 1.0f) * 0.5f; }

// ---- per-splat projection (forward.cu:157-260) -------------------------------------------------------------------
// rect (tile coordinates, x0 y0 x1 y1 as u16), depth bits, packed screen data; radius 0 / no tiles = culled.
__global__ void __launch_bounds__(256)
gs_preprocess_kernel(int P, const float *__restrict__ means, const float *__restrict__ cov3D_precomp,
                     const float *__restrict__ scales, const float *__restrict__ rotations, float scale_modifier,
                     const float *__restrict__ opacities, const float *__restrict__ colors_precomp,
                     const float *__restrict__ shs, int sh_degree, int sh_coeffs, const float *__restrict__ viewmatrix,
                     const float *__restrict__ projmatrix, const float *__restrict__ campos, int W, int H, float tan_fovx,
                     float tan_fovy, int tiles_x, int tiles_y, int *__restrict__ radii, uint32_t *__restrict__ depth_key,
                     uint32_t *__restrict__ ids, uint32_t *__restrict__ tiles_touched, ushort4 *__restrict__ rects,
                     float4 *__restrict__ s0, float4 *__restrict__ s1, float4 *__restrict__ s2) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  radii[idx] = 0;
  tiles_touched[idx] = 0;
  depth_key[idx] = 0u;
  ids[idx] = (uint32_t)idx;
  rects[idx] = make_ushort4(0, 0, 0, 0);
  const float px = means[3 * idx], py = means[3 * idx + 1], pz = means[3 * idx + 2];
  const float *V = viewmatrix, *Pm = projmatrix;  // column-major 4x4 (row vectors multiply from the left)
  const float vx = V[0] * px + V[4] * py + V[8] * pz + V[12];
  const float vy = V[1] * px + V[5] * py + V[9] * pz + V[13];
  const float vz = V[2] * px + V[6] * py + V[10] * pz + V[14];
  if (vz <= 0.2f) return;  // near cull (auxiliary.h:176)
  const float hx = Pm[0] * px + Pm[4] * py + Pm[8] * pz + Pm[12];
  const float hy = Pm[1] * px + Pm[5] * py + Pm[9] * pz + Pm[13];
  const float hw = Pm[3] * px + Pm[7] * py + Pm[11] * pz + Pm[15];
  const float pw = 1.0f / (hw + 0.0000001f);
  const float projx = hx * pw, projy = hy * pw;

  float c3[6];
  if (cov3D_precomp) {
#pragma unroll
    for (int k = 0; k < 6; k++) c3[k] = cov3D_precomp[6 * idx + k];
  } else {
    // Sigma = (S R)^T (S R), quaternion (r, x, y, z) NOT normalised (forward.cu:118-153)
    const float sx = scale_modifier * scales[3 * idx], sy = scale_modifier * scales[3 * idx + 1], sz = scale_modifier * scales[3 * idx + 2];
    const float r = rotations[4 * idx], x = rotations[4 * idx + 1], y = rotations[4 * idx + 2], z = rotations[4 * idx + 3];
    // rows of the rotation matrix as glm stores its columns
    const float R[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                           {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                           {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
    const float s[3] = {sx, sy, sz};
    // M = S * R in glm (column-major) terms: M[c][r] = s[r] * Rg[c][r] with Rg[c] = R[c] above; Sigma = M^T M
    float Mm[3][3];
#pragma unroll
    for (int c = 0; c < 3; c++)
#pragma unroll
      for (int rr = 0; rr < 3; rr++) Mm[c][rr] = s[rr] * R[c][rr];
    auto sig = [&](int a, int b) { return Mm[a][0] * Mm[b][0] + Mm[a][1] * Mm[b][1] + Mm[a][2] * Mm[b][2]; };
    c3[0] = sig(0, 0); c3[1] = sig(0, 1); c3[2] = sig(0, 2); c3[3] = sig(1, 1); c3[4] = sig(1, 2); c3[5] = sig(2, 2);
  }
  // EWA projection of the covariance (forward.cu:75-113)
  const float focal_y = H / (2.0f * tan_fovy), focal_x = W / (2.0f * tan_fovx);
  const float limx = 1.3f * tan_fovx, limy = 1.3f * tan_fovy;
  const float tz = vz;
  const float tx = fminf(limx, fmaxf(-limx, vx / tz)) * tz, ty = fminf(limy, fmaxf(-limy, vy / tz)) * tz;
  const float j00 = focal_x / tz, j02 = -(focal_x * tx) / (tz * tz), j11 = focal_y / tz, j12 = -(focal_y * ty) / (tz * tz);
  // T = W * J with W = rotation part of the view matrix (rows V[0],V[4],V[8] / ...); cov = T^T Vrk T, 2x2 part
  const float t00 = V[0] * j00 + V[2] * j02, t01 = V[4] * j00 + V[6] * j02, t02 = V[8] * j00 + V[10] * j02;
  const float t10 = V[1] * j11 + V[2] * j12, t11 = V[5] * j11 + V[6] * j12, t12 = V[9] * j11 + V[10] * j12;
  const float a0 = c3[0] * t00 + c3[1] * t01 + c3[2] * t02, a1 = c3[1] * t00 + c3[3] * t01 + c3[4] * t02,
              a2 = c3[2] * t00 + c3[4] * t01 + c3[5] * t02;
  const float b0 = c3[0] * t10 + c3[1] * t11 + c3[2] * t12, b1 = c3[1] * t10 + c3[3] * t11 + c3[4] * t12,
              b2 = c3[2] * t10 + c3[4] * t11 + c3[5] * t12;
  const float cxx = t00 * a0 + t01 * a1 + t02 * a2 + 0.3f, cxy = t00 * b0 + t01 * b1 + t02 * b2,
              cyy = t10 * b0 + t11 * b1 + t12 * b2 + 0.3f;
  const float det = cxx * cyy - cxy * cxy;
  if (det == 0.0f) return;
  const float det_inv = 1.f / det;
  const float conA = cyy * det_inv, conB = -cxy * det_inv, conC = cxx * det_inv;
  const float mid = 0.5f * (cxx + cyy);
  const float root = sqrtf(fmaxf(0.1f, mid * mid - det));
  const float my_radius = ceilf(3.f * sqrtf(fmaxf(mid + root, mid - root)));
  const float ix = ndc2pix(projx, W), iy = ndc2pix(projy, H);
  const int rad = (int)my_radius;  // getRect takes the radius as int (auxiliary.h:68)
  const int x0 = min(tiles_x, max(0, (int)((ix - rad) / TILE))), y0 = min(tiles_y, max(0, (int)((iy - rad) / TILE)));
  const int x1 = min(tiles_x, max(0, (int)((ix + rad + TILE - 1) / TILE))), y1 = min(tiles_y, max(0, (int)((iy + rad + TILE - 1) / TILE)));
  if ((x1 - x0) * (y1 - y0) == 0) return;

  float cr, cg, cb;
  if (colors_precomp) {
    cr = colors_precomp[3 * idx]; cg = colors_precomp[3 * idx + 1]; cb = colors_precomp[3 * idx + 2];
  } else {
    // spherical harmonics, degree <= 3, e3nn axis convention of the reference fork (forward.cu:20-72)
    constexpr float C0 = 0.28209479177387814f, C1 = 0.4886025119029199f;
    constexpr float C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f, -1.0925484305920792f, 0.5462742152960396f};
    constexpr float C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                             -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};
    float dx = px - campos[0], dy = py - campos[1], dz = pz - campos[2];
    const float len = sqrtf(dx * dx + dy * dy + dz * dz);
    dx /= len; dy /= len; dz /= len;
    const float *sh = shs + (size_t)idx * sh_coeffs * 3;
    float res[3];
#pragma unroll
    for (int ch = 0; ch < 3; ch++) {
      auto S = [&](int k) { return sh[3 * k + ch]; };
      float v = C0 * S(0);
      if (sh_degree > 0) {
        const float x = dx, y = dy, z = dz;
        v = v - C1 * x * S(1) + C1 * y * S(2) - C1 * z * S(3);
        if (sh_degree > 1) {
          const float zz = z * z, xx = x * x, yy = y * y, zx = z * x, xy = x * y, zy = z * y;
          v = v + C2[0] * zx * S(4) + C2[1] * xy * S(5) + C2[2] * (2.0f * yy - zz - xx) * S(6) + C2[3] * zy * S(7) +
              C2[4] * (zz - xx) * S(8);
          if (sh_degree > 2) {
            v = v + C3[0] * x * (3.0f * zz - xx) * S(9) + C3[1] * zx * y * S(10) + C3[2] * x * (4.0f * yy - zz - xx) * S(11) +
                C3[3] * y * (2.0f * yy - 3.0f * zz - 3.0f * xx) * S(12) + C3[4] * z * (4.0f * yy - zz - xx) * S(13) +
                C3[5] * z * (zz - xx) * S(14) + C3[6] * z * (zz - 3.0f * xx) * S(15);
          }
        }
      }
      res[ch] = fmaxf(v + 0.5f, 0.0f);
    }
    cr = res[0]; cg = res[1]; cb = res[2];
  }
  radii[idx] = rad;
  depth_key[idx] = __float_as_uint(vz);  // vz > 0.2: the bit pattern orders like the value
  tiles_touched[idx] = (uint32_t)((x1 - x0) * (y1 - y0));
  rects[idx] = make_ushort4((unsigned short)x0, (unsigned short)y0, (unsigned short)x1, (unsigned short)y1);
  // power = -0.5 (A dx^2 + C dy^2) - B dx dy = a dx^2 + c dy^2 + b dx dy with the conic pre-scaled (exact: powers of two)
  const float op = opacities[idx];
  s0[idx] = make_float4(ix, iy, -0.5f * conA, -conB);
  // s1 = {conic c, cut-off exponent, opacity, red}, s2 = {green, blue, extent x, extent y}: a pixel that rejects the
  // splat reads s0 and s1 only
  // Culling aids for the render kernel, CONSERVATIVE by construction (they only skip work whose result is known):
  // a pixel takes a term iff alpha = min(0.99, op * exp(power)) >= 1/255, i.e. power >= -log(255 op).  thr sits 1e-3
  // below that (far more than the rounding of expf and of the product), and (ex, ey) is the half-extent of the
  // ellipse power >= thr — sqrt(2 |thr| cov_xx), the conic's inverse being the 2-D covariance — plus half a pixel.
  const float thr = -logf(255.0f * op) - 1e-3f;
  const bool never = !(thr < 0.0f);  // op <= 1/255 (or NaN): no pixel passes... NaN is left to the exact test
  // (an indefinite conic — det <= 0, a covariance that is not positive definite — has no bounded extent: no culling)
  const bool sane = op == op && det > 0.0f && cxx > 0.0f && cyy > 0.0f;
  const float ex = !sane ? 3.0e38f : never ? -1.0f : sqrtf(-2.0f * thr * cxx) + 0.5f;
  const float ey = !sane ? 3.0e38f : never ? -1.0f : sqrtf(-2.0f * thr * cyy) + 0.5f;
  s1[idx] = make_float4(-0.5f * conC, sane ? thr : -3.0e38f, op, cr);
  s2[idx] = make_float4(cg, cb, ex, ey);
}

// tiles touched by the splat at depth rank s: the scan reads it through this functor (no gathered copy)
struct CountInDepthOrder {
  const uint32_t *ids_sorted, *tiles_touched;
  __host__ __device__ __forceinline__ uint32_t operator()(int s) const { return tiles_touched[ids_sorted[s]]; }
};
using CountIter = thrust::transform_iterator<CountInDepthOrder, thrust::counting_iterator<int>>;

// instances in depth order: key = tile, value = splat (rasterizer_impl.cu:78-120 emits (tile | depth) keys instead)
__global__ void gs_emit_kernel(int P, const uint32_t *__restrict__ ids_sorted, const uint32_t *__restrict__ incl,
                               const ushort4 *__restrict__ rects, int tiles_x, long long capacity,
                               uint32_t *__restrict__ keys, uint32_t *__restrict__ vals) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= P) return;
  const uint32_t g = ids_sorted[s];
  const ushort4 r = rects[g];
  long long off = (long long)incl[s] - (long long)((r.z - r.x) * (r.w - r.y));  // inclusive sum minus this splat's own tiles
  for (int y = r.y; y < r.w; y++)
    for (int x = r.x; x < r.z; x++) {
      if (off < capacity) {
        keys[off] = (uint32_t)(y * tiles_x + x);
        vals[off] = g;
      }
      off++;
    }
}

__global__ void gs_ranges_kernel(long long L, const uint32_t *__restrict__ keys, uint2 *__restrict__ ranges) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= L) return;
  const uint32_t t = keys[i];
  if (i == 0) ranges[t].x = 0;
  else {
    const uint32_t p = keys[i - 1];
    if (p != t) { ranges[p].y = (uint32_t)i; ranges[t].x = (uint32_t)i; }
  }
  if (i == L - 1) ranges[t].y = (uint32_t)L;
}

// ---- blending (forward.cu:262-370) --------------------------------------------------------------------------------
// One CTA per 16 x 16 tile as in the reference (the tile decides which splats a pixel sees); inside it a warp owns an
// 8 x 4 pixel block.  Per batch of 256 staged splats a warp first tests 32 splats at a time, one per lane, against
// its block (the splat's conservative extent from the preprocess) and then walks only the set bits of the ballot, in
// order; a pixel skips expf when the exponent is surely below the 1/255 cut-off.  Both tests only skip terms the
// reference's own tests reject, so the blended sequence per pixel is the reference's.  (ncu of the first version —
// every pixel evaluating every splat of its tile — 77 % issue-active, 277 M warp instructions at 786 k splats.)
__global__ void __launch_bounds__(RENDER_THREADS)
gs_render_kernel(const uint2 *__restrict__ ranges, const uint32_t *__restrict__ point_list, int W, int H,
                 const float4 *__restrict__ s0, const float4 *__restrict__ s1, const float4 *__restrict__ s2,
                 const float *__restrict__ bg, float *__restrict__ out_color) {
  __shared__ float4 c0[RENDER_THREADS], c1[RENDER_THREADS], c2[RENDER_THREADS];
  const int tiles_x = (W + TILE - 1) / TILE;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int bx0 = blockIdx.x * TILE + (warp & 1) * 8, by0 = blockIdx.y * TILE + (warp >> 1) * 4;  // the warp's 8 x 4 block
  const int pxi = bx0 + (lane & 7), pyi = by0 + (lane >> 3);
  const bool inside = pxi < W && pyi < H;
  const float pixx = (float)pxi, pixy = (float)pyi;
  const float fx0 = (float)bx0, fx1 = (float)(bx0 + 7), fy0 = (float)by0, fy1 = (float)(by0 + 3);
  const uint2 range = ranges[blockIdx.y * tiles_x + blockIdx.x];
  int todo = (int)(range.y - range.x);
  bool done = !inside;
  float T = 1.0f, Cr = 0.0f, Cg = 0.0f, Cb = 0.0f;
  for (uint32_t base = range.x; todo > 0; base += RENDER_THREADS, todo -= RENDER_THREADS) {
    if (__syncthreads_count(done) == RENDER_THREADS) break;
    if ((int)threadIdx.x < todo) {
      const uint32_t g = point_list[base + threadIdx.x];
      c0[threadIdx.x] = __ldg(s0 + g);
      c1[threadIdx.x] = __ldg(s1 + g);
      c2[threadIdx.x] = __ldg(s2 + g);
    }
    __syncthreads();
    const int nb = min(RENDER_THREADS, todo);
    if (__all_sync(0xffffffffu, done)) continue;  // (still takes part in the staging above)
    for (int j0 = 0; j0 < nb; j0 += 32) {
      unsigned mask;
      {
        const int j = j0 + lane;
        const float4 a = c0[j < nb ? j : 0], e = c2[j < nb ? j : 0];
        const float ddx = a.x - fminf(fmaxf(a.x, fx0), fx1), ddy = a.y - fminf(fmaxf(a.y, fy0), fy1);
        mask = __ballot_sync(0xffffffffu, j < nb && fabsf(ddx) <= e.z && fabsf(ddy) <= e.w);
      }
      while (mask) {
        const int j = j0 + __ffs(mask) - 1;
        mask &= mask - 1;
        if (done) continue;
        const float4 a = c0[j];
        const float4 b = c1[j];
        const float dx = a.x - pixx, dy = a.y - pixy;
        const float power = a.z * dx * dx + b.x * dy * dy + a.w * dx * dy;
        if (power > 0.0f || power < b.y) continue;  // b.y: surely below the 1/255 cut-off (the exact test follows)
        const float alpha = fminf(0.99f, b.z * expf(power));
        if (alpha < 1.0f / 255.0f) continue;
        const float test_T = T * (1.0f - alpha);
        if (test_T < 0.0001f) { done = true; continue; }
        const float4 e = c2[j];
        const float wgt = alpha * T;
        Cr += b.w * wgt; Cg += e.x * wgt; Cb += e.y * wgt;
        T = test_T;
      }
      if (__all_sync(0xffffffffu, done)) break;
    }
  }
  if (inside) {
    const size_t pix = (size_t)pyi * W + pxi, hw = (size_t)H * W;
    out_color[pix] = Cr + T * bg[0];
    out_color[hw + pix] = Cg + T * bg[1];
    out_color[2 * hw + pix] = Cb + T * bg[2];
  }
}

inline size_t al(size_t x) { return (x + 255) & ~(size_t)255; }
struct GsLayout {
  size_t key_a, key_b, id_a, id_b, touched, counts, offsets, rects, s0, s1, s2, ikey_a, ikey_b, ival_a, ival_b, ranges, temp,
      temp_bytes, total;
};
inline int tile_bits(int tiles) {
  int b = 1;
  while ((1 << b) < tiles) b++;
  return b;
}
inline GsLayout gs_layout(int P, int H, int W, long long cap) {
  GsLayout L;
  const size_t p = (size_t)(P > 0 ? P : 1), c = (size_t)(cap > 0 ? cap : 1);
  const int tiles = ((W + TILE - 1) / TILE) * ((H + TILE - 1) / TILE);
  size_t off = 0;
  auto take = [&](size_t bytes) { const size_t o = off; off = al(off + bytes); return o; };
  L.key_a = take(p * 4); L.key_b = take(p * 4); L.id_a = take(p * 4); L.id_b = take(p * 4);
  L.touched = take(p * 4); L.counts = take(p * 4); L.offsets = take(p * 4); L.rects = take(p * 8);
  L.s0 = take(p * 16); L.s1 = take(p * 16); L.s2 = take(p * 16);
  L.ikey_a = take(c * 4); L.ikey_b = take(c * 4); L.ival_a = take(c * 4); L.ival_b = take(c * 4);
  L.ranges = take((size_t)(tiles > 0 ? tiles : 1) * 8);
  size_t t1 = 0, t2 = 0, t3 = 0;
  cub::DoubleBuffer<uint32_t> dk(nullptr, nullptr), dv(nullptr, nullptr);
  cub::DeviceRadixSort::SortPairs(nullptr, t1, dk, dv, (int)p, 0, 32);
  cub::DeviceScan::InclusiveSum(nullptr, t2, CountIter(thrust::counting_iterator<int>(0), CountInDepthOrder{nullptr, nullptr}), (uint32_t *)nullptr, (int)p);
  cub::DeviceRadixSort::SortPairs(nullptr, t3, dk, dv, (int)(c < 0x7fffffffull ? c : 0x7fffffffull), 0, tile_bits(tiles));
  L.temp_bytes = t1 > t2 ? (t1 > t3 ? t1 : t3) : (t2 > t3 ? t2 : t3);
  L.temp = take(L.temp_bytes);
  L.total = off;
  return L;
}

}  // namespace b200gs

using namespace b200gs;

extern "C" size_t b200slam_gs_workspace_bytes(int P, int H, int W, long long max_instances) {
  if (P < 0 || H < 1 || W < 1 || max_instances < 0) return 0;
  return gs_layout(P, H, W, max_instances).total;
}

extern "C" int b200slam_gs_render(const float *means3D, const float *cov3D_precomp, const float *scales,
                                  const float *rotations, float scale_modifier, const float *opacities,
                                  const float *colors_precomp, const float *shs, int sh_degree, int sh_coeffs,
                                  const float *campos, const float *viewmatrix, const float *projmatrix, float tan_fovx,
                                  float tan_fovy, const float *bg, int P, int H, int W, float *out_color, int *radii,
                                  long long max_instances, long long *num_instances, void *workspace,
                                  size_t workspace_bytes, void *stream) {
  if (P < 0 || H < 1 || W < 1 || max_instances < 0 || H > 0xffff * TILE || W > 0xffff * TILE) return B200SLAM_EINVAL;
  if (!out_color || !bg || !viewmatrix || !projmatrix || !workspace) return B200SLAM_EINVAL;
  if (P > 0 && (!means3D || !opacities || !radii || (!cov3D_precomp && (!scales || !rotations)) ||
                (!colors_precomp && (!shs || !campos || sh_degree < 0 || sh_degree > 3 || sh_coeffs < (sh_degree + 1) * (sh_degree + 1)))))
    return B200SLAM_EINVAL;
  const GsLayout L = gs_layout(P, H, W, max_instances);
  if (workspace_bytes < L.total) return B200SLAM_EWORKSPACE;
  cudaStream_t st = (cudaStream_t)stream;
  char *ws = (char *)workspace;
  const int tiles_x = (W + TILE - 1) / TILE, tiles_y = (H + TILE - 1) / TILE, tiles = tiles_x * tiles_y;
  uint2 *ranges = reinterpret_cast<uint2 *>(ws + L.ranges);
  cudaError_t err = cudaMemsetAsync(ranges, 0, (size_t)tiles * 8, st);
  if (err != cudaSuccess) return (int)err;
  long long total = 0;
  const uint32_t *point_list = nullptr;
  float4 *s0 = reinterpret_cast<float4 *>(ws + L.s0), *s1 = reinterpret_cast<float4 *>(ws + L.s1);
  float4 *s2 = reinterpret_cast<float4 *>(ws + L.s2);
  if (P > 0) {
    uint32_t *touched = reinterpret_cast<uint32_t *>(ws + L.touched), *counts = reinterpret_cast<uint32_t *>(ws + L.counts);
    uint32_t *offsets = reinterpret_cast<uint32_t *>(ws + L.offsets);
    ushort4 *rects = reinterpret_cast<ushort4 *>(ws + L.rects);
    cub::DoubleBuffer<uint32_t> dk(reinterpret_cast<uint32_t *>(ws + L.key_a), reinterpret_cast<uint32_t *>(ws + L.key_b));
    cub::DoubleBuffer<uint32_t> dv(reinterpret_cast<uint32_t *>(ws + L.id_a), reinterpret_cast<uint32_t *>(ws + L.id_b));
    const int pb = (P + 255) / 256;
    gs_preprocess_kernel<<<pb, 256, 0, st>>>(P, means3D, cov3D_precomp, scales, rotations, scale_modifier, opacities,
                                             colors_precomp, shs, sh_degree, sh_coeffs, viewmatrix, projmatrix, campos, W, H,
                                             tan_fovx, tan_fovy, tiles_x, tiles_y, radii, dk.Current(), dv.Current(), touched,
                                             rects, s0, s1, s2);
    size_t tb = L.temp_bytes;
    err = cub::DeviceRadixSort::SortPairs(ws + L.temp, tb, dk, dv, P, 0, 32, st);  // splats by view depth (stable)
    if (err != cudaSuccess) return (int)err;
    tb = L.temp_bytes;
    err = cub::DeviceScan::InclusiveSum(ws + L.temp, tb, CountIter(thrust::counting_iterator<int>(0), CountInDepthOrder{dv.Current(), touched}),
                                        offsets, P, st);
    if (err != cudaSuccess) return (int)err;
    // the instance count decides the launch sizes below: one host read per call, like the reference
    // (rasterizer_impl.cu:289)
    uint32_t last = 0;
    err = cudaMemcpyAsync(&last, offsets + (P - 1), 4, cudaMemcpyDeviceToHost, st);
    if (err == cudaSuccess) err = cudaStreamSynchronize(st);
    if (err != cudaSuccess) return (int)err;
    total = (long long)last;
    (void)counts;
    if (num_instances) *num_instances = total;
    if (total > max_instances || total > 0x7fffffffll) return B200SLAM_EWORKSPACE;  // retry with a larger workspace
    if (total > 0) {
      cub::DoubleBuffer<uint32_t> ik(reinterpret_cast<uint32_t *>(ws + L.ikey_a), reinterpret_cast<uint32_t *>(ws + L.ikey_b));
      cub::DoubleBuffer<uint32_t> iv(reinterpret_cast<uint32_t *>(ws + L.ival_a), reinterpret_cast<uint32_t *>(ws + L.ival_b));
      gs_emit_kernel<<<pb, 256, 0, st>>>(P, dv.Current(), offsets, rects, tiles_x, max_instances, ik.Current(), iv.Current());
      tb = L.temp_bytes;
      err = cub::DeviceRadixSort::SortPairs(ws + L.temp, tb, ik, iv, (int)total, 0, tile_bits(tiles), st);  // stable: keeps depth order
      if (err != cudaSuccess) return (int)err;
      gs_ranges_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(total, ik.Current(), ranges);
      point_list = iv.Current();
    }
  } else if (num_instances) {
    *num_instances = 0;
  }
  gs_render_kernel<<<dim3(tiles_x, tiles_y), RENDER_THREADS, 0, st>>>(ranges, point_list, W, H, s0, s1, s2, bg, out_color);
  return (int)cudaGetLastError();
}
