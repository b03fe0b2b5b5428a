// DMMA / DFMA dependent-issue latency on B200: one warp per SMSP, C independent accumulator chains.
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void dmma884(double &d0, double &d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
template <int C>
__global__ void k_dmma(double *out, long long *cyc, double a, double b, int iters) {
  double c[2 * C];
  for (int i = 0; i < 2 * C; ++i) c[i] = threadIdx.x * 1e-3 + i;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < C; ++i) dmma884(c[2 * i], c[2 * i + 1], a, b);
  }
  long long t1 = clock64();
  double s = 0;
  for (int i = 0; i < 2 * C; ++i) s += c[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}
template <int C>
__global__ void k_dfma(double *out, long long *cyc, double a, double b, int iters) {
  double c[C];
  for (int i = 0; i < C; ++i) c[i] = threadIdx.x * 1e-3 + i;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < C; ++i) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(c[i]) : "d"(a), "d"(b));
  }
  long long t1 = clock64();
  double s = 0;
  for (int i = 0; i < C; ++i) s += c[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}
template <int C>
void run(double *out, long long *cyc, int warps) {
  const int iters = 2000;
  long long h;
  k_dmma<C><<<1, 32 * warps>>>(out, cyc, 1.0000001, 1e-9, iters);
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("DMMA warps/SM %d chains %d: %.1f cycles per DMMA issue (per chain step %.1f)\n", warps, C, (double)h / iters / C, (double)h / iters);
  k_dfma<C><<<1, 32 * warps>>>(out, cyc, 1.0000001, 1e-9, iters);
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("DFMA warps/SM %d chains %d: %.1f cycles per DFMA issue (per chain step %.1f)\n", warps, C, (double)h / iters / C, (double)h / iters);
}
int main() {
  double *out; long long *cyc;
  cudaMalloc(&out, 8 * 1024 * 64); cudaMalloc(&cyc, 64);
  for (int warps : {4, 8, 12}) {
    run<1>(out, cyc, warps); run<2>(out, cyc, warps); run<4>(out, cyc, warps); run<6>(out, cyc, warps); run<8>(out, cyc, warps); run<12>(out, cyc, warps);
  }
  return 0;
}
