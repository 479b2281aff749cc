// Micro-benchmark: cost of one grid-wide "publish tag / wait for all tags" exchange between G co-resident CTAs of
// 512 threads (the synchronisation of the pipelined PCG, csrc/gn_pcg2.cuh), for several implementations.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o xchg_latency xchg_latency.cu && ./xchg_latency
// Prints cycles per exchange (CTA 0, warp 1) for G in {4, 16, 64, 148}.
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned ld_relaxed(const unsigned *p) {
  unsigned v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned ld_acquire(const unsigned *p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release(unsigned *p, unsigned v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed(unsigned *p, unsigned v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// MODE bit 0: pollers — 0 every warp polls, 1 warp 0 polls then bar.sync
// MODE bit 1: release — 0 st.release.gpu, 1 __threadfence() + st.relaxed
// MODE bit 2: poll load — 0 ld.relaxed.gpu, 1 ld.acquire.gpu
// MODE bit 3: tag stride — 0 packed (4 B), 1 one tag per 128-byte line
// MODE bit 4: payload — each warp also stores 7 doubles before the publish (like the m rows)
template <int MODE>
__global__ void __launch_bounds__(512, 1) xchg(unsigned *tags, double *payload, int iters, long long *out) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, G = gridDim.x;
  const int stride = (MODE & 8) ? 32 : 1;
  long long c0 = clock64();
  for (int it = 1; it <= iters; it++) {
    unsigned *t = tags + (size_t)(it & 1) * 160 * 32;
    if (MODE & 16) {
      if (lane < 7) __stcg(payload + ((size_t)(it & 1) * G * 16 + blockIdx.x * 16 + warp) * 8 + lane, (double)it);
    }
    __syncthreads();
    if (tid == 0) {
      if (MODE & 2) { __threadfence(); st_relaxed(t + blockIdx.x * stride, (unsigned)it); }
      else st_release(t + blockIdx.x * stride, (unsigned)it);
    }
    __syncthreads();
    if (!(MODE & 1) || warp == 0) {
      for (int c = lane; c < G; c += 32) {
        if (MODE & 4) { while (ld_acquire(t + c * stride) < (unsigned)it) {} }
        else { while (ld_relaxed(t + c * stride) < (unsigned)it) {} }
      }
      __syncwarp();
    }
    if (MODE & 1) __syncthreads();
  }
  if (blockIdx.x == 0 && tid == 32) out[0] = (clock64() - c0) / iters;
}

template <int MODE>
void run(const char *name, unsigned *tags, double *payload, long long *out) {
  printf("%-58s", name);
  for (int G : {4, 16, 64, 148}) {
    cudaMemset(tags, 0, 2 * 160 * 32 * sizeof(unsigned));
    int iters = 2000;
    void *args[] = {&tags, &payload, &iters, &out};
    cudaFuncSetAttribute(xchg<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaError_t e = cudaLaunchCooperativeKernel((const void *)xchg<MODE>, dim3(G), dim3(512), args, 200 * 1024, 0);
    cudaDeviceSynchronize();
    long long h = -1;
    cudaMemcpy(&h, out, sizeof(h), cudaMemcpyDeviceToHost);
    if (e != cudaSuccess || cudaGetLastError() != cudaSuccess) h = -1;
    printf("  G=%-3d %7lld", G, h);
  }
  printf("   cycles/exchange\n");
}

int main() {
  unsigned *tags;
  double *payload;
  long long *out;
  cudaMalloc(&tags, 2 * 160 * 32 * sizeof(unsigned));
  cudaMalloc(&payload, 2 * 160 * 16 * 8 * sizeof(double));
  cudaMalloc(&out, 64);
  run<0>("all warps poll, st.release, ld.relaxed, packed tags", tags, payload, out);
  run<16>("  + payload stores", tags, payload, out);
  run<1>("warp 0 polls + bar.sync, st.release, ld.relaxed", tags, payload, out);
  run<17>("  + payload stores", tags, payload, out);
  run<2>("all warps poll, threadfence + st.relaxed", tags, payload, out);
  run<4>("all warps poll, st.release, ld.ACQUIRE", tags, payload, out);
  run<5>("warp 0 polls + bar.sync, st.release, ld.ACQUIRE", tags, payload, out);
  run<8>("all warps poll, st.release, ld.relaxed, tag per line", tags, payload, out);
  run<9>("warp 0 polls + bar.sync, st.release, ld.relaxed, tag per line", tags, payload, out);
  run<25>("  + payload stores", tags, payload, out);
  return 0;
}
