// fp64 issue-rate microbenchmark for sm_100a: DFMA vs DMMA (mma.sync.m8n8k4.f64), per SM per clock.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/micro/fp64_rate tools/micro/fp64_rate.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int ILP>
__global__ void dfma_kernel(double *out, int iters, long long *cyc) {
  double a[ILP];
  for (int i = 0; i < ILP; i++) a[i] = threadIdx.x * 1e-3 + i;
  const double b = 1.0000001, c = 1e-9;
  const long long t0 = clock64();
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < ILP; i++) a[i] = fma(a[i], b, c);
  }
  const long long t1 = clock64();
  double s = 0;
  for (int i = 0; i < ILP; i++) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int ILP>
__global__ void dmma_kernel(double *out, int iters, long long *cyc) {
  double c0[ILP], c1[ILP];
  for (int i = 0; i < ILP; i++) { c0[i] = threadIdx.x * 1e-3 + i; c1[i] = 0.5 * i; }
  const double a = 1.0000001, b = 1e-3 * (threadIdx.x & 3);
  const long long t0 = clock64();
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < ILP; i++)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                   : "+d"(c0[i]), "+d"(c1[i]) : "d"(a), "d"(b));
  }
  const long long t1 = clock64();
  double s = 0;
  for (int i = 0; i < ILP; i++) s += c0[i] + c1[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

int main() {
  double *out; long long *cyc, h[4];
  cudaMalloc(&out, 148 * 1024 * sizeof(double));
  cudaMalloc(&cyc, 148 * sizeof(long long));
  const int iters = 4096;
  for (int threads = 128; threads <= 1024; threads *= 2) {
    dfma_kernel<8><<<148, threads>>>(out, iters, cyc);
    cudaMemcpy(h, cyc, sizeof(long long), cudaMemcpyDeviceToHost);
    const double dfma = (double)iters * 8 * threads / h[0];
    dmma_kernel<8><<<148, threads>>>(out, iters, cyc);
    cudaMemcpy(h, cyc, sizeof(long long), cudaMemcpyDeviceToHost);
    const double dmma = (double)iters * 8 * (threads / 32) * 256 / h[0];
    printf("{\"threads_per_sm\": %d, \"dfma_fma_per_clk_per_sm\": %.1f, \"dmma_fma_per_clk_per_sm\": %.1f}\n", threads, dfma, dmma);
  }
  printf("err=%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
