// Experiment (north_star: "tensor cores only for the batched descriptor-correlation contraction"): what does the
// DENSE fp16 correlation of a refine_matches tile cost on the 5th-generation tensor cores?
//
// refine_matches scores, per query, 245 candidates x 24 channels with a sequentially fp16-rounded sum that no MMA
// reproduces, so tensor cores could only PRE-RANK: score every query of a tile against every pixel of the tile's
// staged box (fp32 accumulate), keep the candidates within an error bound of the maximum, and re-score those in the
// reference's arithmetic.  This program measures the first step alone, at refine's real shapes: per 512-query tile the
// boxes of the five dilation levels hold 9252 pixels (3072 + 2436 + 1872 + 1872 for the merged levels 2+1), i.e. 37
// chunks of 256 pixels for each of 4 slabs of 128 queries; 432 tiles per 512x384 pair.
//   A = 128 queries x K, B = 256 pixels x K, K = 32 (24 channels + 8 zero: UMMA K is a multiple of 16),
//   D = A * B^T in TMEM (128 lanes x 256 fp32 columns), tcgen05.mma.cta_group::1.kind::f16, M = 128, N = 256.
// Operands are laid out in shared memory in the canonical K-major no-swizzle form (8 x 16-byte core matrices,
// LBO = 128 B between K chunks, SBO = 512 B between 8-row groups); the accumulator is drained with tcgen05.ld
// 32x32b.x32 and reduced to a per-query maximum (the cheapest possible consumer: a real pre-rank has to pick the
// query's 49 candidates of the level out of the 256 columns).
// Modes: 0 = MMA + drain (one accumulator buffer, serial), 1 = MMA only (no drain), 2 = two accumulator buffers
// (the drain of chunk c overlaps the MMA of chunk c+1).
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -o tc_corr tc_corr.cu && ./tc_corr
#include <cstdio>
#include <cstdlib>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <vector>

constexpr int KP = 32;                // padded channels
constexpr int MQ = 128, NB = 256;     // UMMA M, N
constexpr int SLAB_BYTES = MQ * KP * 2, CHUNK_BYTES = NB * KP * 2;  // 8 KB, 16 KB

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// K-major, SWIZZLE_NONE shared-memory matrix descriptor (cute/arch/mma_sm100_desc.hpp: SmemDescriptor)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3fff);         // start address
  d |= (uint64_t)((128u >> 4) & 0x3fff) << 16;    // leading byte offset: next 16-byte K chunk
  d |= (uint64_t)((512u >> 4) & 0x3fff) << 32;    // stride byte offset: next 8-row group
  d |= (uint64_t)1 << 46;                         // descriptor version (sm_100)
  return d;                                       // layout_type = 0: no swizzle
}
// kind::f16 instruction descriptor: D = f32, A = B = f16, both K-major, M = 128, N = 256
__host__ __device__ constexpr uint32_t make_idesc() {
  return (1u << 4) | (0u << 7) | (0u << 10) | (0u << 15) | (0u << 16) | ((uint32_t)(NB >> 3) << 17) | ((uint32_t)(MQ >> 4) << 24);
}

// Stores row r (24 halves at src) of an operand tile in the canonical layout: [r/8][k chunk][r%8][8 halves].
__device__ __forceinline__ void store_row(unsigned char *tile, int r, const __half *src) {
  const uint4 *s = reinterpret_cast<const uint4 *>(src);
  unsigned char *base = tile + (r >> 3) * 512 + (r & 7) * 16;
  *reinterpret_cast<uint4 *>(base) = __ldg(s);
  *reinterpret_cast<uint4 *>(base + 128) = __ldg(s + 1);
  *reinterpret_cast<uint4 *>(base + 256) = __ldg(s + 2);
  *reinterpret_cast<uint4 *>(base + 384) = make_uint4(0, 0, 0, 0);
}

template <int MODE>
__global__ void __launch_bounds__(256, 1)
corr_kernel(const __half *__restrict__ Q, const __half *__restrict__ Bx, int chunks, int slabs, float *__restrict__ out) {
  extern __shared__ __align__(1024) unsigned char sm[];
  unsigned char *sA = sm, *sB0 = sm + SLAB_BYTES, *sB1 = sB0 + CHUNK_BYTES;
  __shared__ uint64_t bar[2];
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_base_s)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  if (tid == 0) {
    for (int k = 0; k < 2; k++) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar[k])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_base_s;
  const size_t tile = blockIdx.x;
  const __half *Qt = Q + tile * (size_t)slabs * MQ * 24, *Bt = Bx + tile * (size_t)chunks * NB * 24;
  uint32_t phase[2] = {0, 0};
  float best = -1e30f;
  for (int s = 0; s < slabs; s++) {
    for (int r = tid; r < MQ; r += 256) store_row(sA, r, Qt + ((size_t)s * MQ + r) * 24);
    for (int c = 0; c <= chunks; c++) {
      unsigned char *sB = (c & 1) ? sB1 : sB0;
      const int buf = (MODE == 2) ? (c & 1) : 0;
      if (c < chunks) {
        for (int r = tid; r < NB; r += 256) store_row(sB, r, Bt + ((size_t)c * NB + r) * 24);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic writes -> tensor-core (async proxy) reads
      }
      __syncthreads();
      if (c < chunks && tid == 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t d_tmem = tmem + buf * NB;
#pragma unroll
        for (int k = 0; k < KP / 16; k++) {
          const uint64_t da = make_desc(smem_u32(sA) + k * 256), db = make_desc(smem_u32(sB) + k * 256);
          const uint32_t acc = k > 0;
          asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
                       "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}" ::"r"(d_tmem), "l"(da), "l"(db),
                       "r"(make_idesc()), "r"(acc)
                       : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar[buf]))
                     : "memory");
      }
      // drain: MODE 0 the chunk just issued; MODE 2 the previous chunk (its MMA overlapped this chunk's loads)
      const int dc = (MODE == 2) ? c - 1 : c;
      if (dc >= 0 && dc < chunks) {
        const int dbuf = (MODE == 2) ? (dc & 1) : 0;
        asm volatile("{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}" ::"r"(
                         smem_u32(&bar[dbuf])),
                     "r"(phase[dbuf])
                     : "memory");
        phase[dbuf] ^= 1u;
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (MODE != 1 && warp < 4) {
          const uint32_t taddr = tmem + dbuf * NB + ((uint32_t)(warp * 32) << 16);
#pragma unroll 1
          for (int col = 0; col < NB; col += 32) {
            uint32_t v[32];
            asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,"
                "%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
                  "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
                  "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
                  "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                : "r"(taddr + col));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int j = 0; j < 32; j++) best = fmaxf(best, __uint_as_float(v[j]));
          }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      }
      __syncthreads();  // sB / accumulator buffer may be overwritten
    }
    if (warp < 4 && MODE != 1) out[(tile * slabs + s) * MQ + warp * 32 + lane] = best;
    best = -1e30f;
  }
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem));
}

// plain reference for validation: per-query maximum over the tile's pixels of the fp32 dot product
__global__ void ref_kernel(const __half *Q, const __half *Bx, int chunks, int slabs, float *out) {
  const size_t tile = blockIdx.x;
  for (int q = threadIdx.x; q < slabs * MQ; q += blockDim.x) {
    const __half *qq = Q + (tile * slabs * MQ + q) * 24;
    float best = -1e30f;
    for (int p = 0; p < chunks * NB; p++) {
      const __half *pp = Bx + (tile * chunks * NB + p) * 24;
      float s = 0.f;
      for (int k = 0; k < 24; k++) s += __half2float(qq[k]) * __half2float(pp[k]);
      best = fmaxf(best, s);
    }
    out[tile * slabs * MQ + q] = best;
  }
}

template <int MODE>
float run(const char *name, const __half *Q, const __half *B, int tiles, int chunks, int slabs, float *out) {
  const int smem = SLAB_BYTES + 2 * CHUNK_BYTES;
  cudaFuncSetAttribute(corr_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  for (int w = 0; w < 2; w++) corr_kernel<MODE><<<tiles, 256, smem>>>(Q, B, chunks, slabs, out);
  cudaEventRecord(a);
  const int reps = 10;
  for (int w = 0; w < reps; w++) corr_kernel<MODE><<<tiles, 256, smem>>>(Q, B, chunks, slabs, out);
  cudaEventRecord(b);
  cudaError_t e = cudaDeviceSynchronize();
  float ms = 0;
  cudaEventElapsedTime(&ms, a, b);
  const double macs = (double)tiles * slabs * chunks * MQ * NB * KP;
  printf("%-44s %8.1f us per pair   %6.1f TFLOP/s (K padded to 32)   %s\n", name, ms / reps * 1e3,
         2 * macs / (ms / reps * 1e-3) / 1e12, e == cudaSuccess ? "" : cudaGetErrorString(e));
  return ms / reps;
}

int main(int argc, char **argv) {
  const int tiles = argc > 1 ? atoi(argv[1]) : 432, chunks = 37, slabs = 4;
  const size_t nq = (size_t)tiles * slabs * MQ * 24, nb = (size_t)tiles * chunks * NB * 24;
  std::vector<__half> hq(nq), hb(nb);
  srand(1);
  for (auto &x : hq) x = __float2half((rand() % 2001 - 1000) * 2e-4f);
  for (auto &x : hb) x = __float2half((rand() % 2001 - 1000) * 2e-4f);
  __half *Q, *B;
  float *out, *ref;
  cudaMalloc(&Q, nq * 2);
  cudaMalloc(&B, nb * 2);
  cudaMalloc(&out, (size_t)tiles * slabs * MQ * 4);
  cudaMalloc(&ref, (size_t)tiles * slabs * MQ * 4);
  cudaMemcpy(Q, hq.data(), nq * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(B, hb.data(), nb * 2, cudaMemcpyHostToDevice);
  printf("dense pre-rank correlation, %d tiles x %d slabs of 128 queries x %d chunks of 256 box pixels\n", tiles, slabs, chunks);
  run<0>("MMA + drain (serial, 1 accumulator)", Q, B, tiles, chunks, slabs, out);
  // validate against the plain kernel on the first 8 tiles
  const int vt = tiles < 8 ? tiles : 8;
  ref_kernel<<<vt, 256>>>(Q, B, chunks, slabs, ref);
  std::vector<float> ho((size_t)vt * slabs * MQ), hr((size_t)vt * slabs * MQ);
  cudaMemcpy(ho.data(), out, ho.size() * 4, cudaMemcpyDeviceToHost);
  cudaMemcpy(hr.data(), ref, hr.size() * 4, cudaMemcpyDeviceToHost);
  double maxerr = 0;
  for (size_t i = 0; i < ho.size(); i++) maxerr = fmax(maxerr, fabs((double)ho[i] - hr[i]));
  printf("validation: max |tcgen05 - plain| over %zu per-query maxima = %.3e (values ~ %.3f)\n", ho.size(), maxerr, hr[0]);
  run<1>("MMA only (accumulator never read)", Q, B, tiles, chunks, slabs, out);
  run<2>("MMA + drain, 2 accumulators (overlapped)", Q, B, tiles, chunks, slabs, out);
  return 0;
}
