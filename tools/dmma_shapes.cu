// fp64 tensor-core instruction shapes on B200 (sm_100a): issue cost and whole-GPU throughput of
// mma.sync m8n8k4 / m16n8k4 / m16n8k8 / m16n8k16 (f64).  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dmma_shapes tools/dmma_shapes.cu
#include <cstdio>
#include <cuda_runtime.h>
struct M884 {
  static constexpr int NC = 2, FL = 8 * 8 * 4 * 2;
  static __device__ __forceinline__ void mma(double *c, const double *a, const double *b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[0]), "+d"(c[1]) : "d"(a[0]), "d"(b[0]));
  }
};
struct M1684 {
  static constexpr int NC = 4, FL = 16 * 8 * 4 * 2;
  static __device__ __forceinline__ void mma(double *c, const double *a, const double *b) {
    asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                 : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3]) : "d"(a[0]), "d"(a[1]), "d"(b[0]));
  }
};
struct M1688 {
  static constexpr int NC = 4, FL = 16 * 8 * 8 * 2;
  static __device__ __forceinline__ void mma(double *c, const double *a, const double *b) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3]) : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
  }
};
struct M16816 {
  static constexpr int NC = 4, FL = 16 * 8 * 16 * 2;
  static __device__ __forceinline__ void mma(double *c, const double *a, const double *b) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
                 : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
                 : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]), "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
  }
};
template <class S, int C>
__global__ void k(double *out, long long *cyc, double av, double bv, int iters) {
  double c[S::NC * C], a[8], b[4];
  for (int i = 0; i < S::NC * C; ++i) c[i] = threadIdx.x * 1e-3 + i;
  for (int i = 0; i < 8; ++i) a[i] = av + i * 1e-9;
  for (int i = 0; i < 4; ++i) b[i] = bv + i * 1e-12;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < C; ++i) S::mma(c + S::NC * i, a, b);
  }
  long long t1 = clock64();
  double s = 0;
  for (int i = 0; i < S::NC * C; ++i) s += c[i];
  out[(blockIdx.x * blockDim.x + threadIdx.x) % 4096] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}
template <class S, int C>
void run(const char *name, double *out, long long *cyc) {
  const int iters = 4000;
  long long h;
  k<S, C><<<1, 128>>>(out, cyc, 1.0000001, 1e-9, iters);  // one warp per SMSP
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  const double cpi = (double)h / iters / C;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<S, C><<<148 * 4, 256>>>(out, cyc, 1.0000001, 1e-9, iters);
  cudaEventRecord(e0);
  k<S, C><<<148 * 4, 256>>>(out, cyc, 1.0000001, 1e-9, iters);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  const double tf = 148.0 * 4 * 8 * iters * C * S::FL / (ms * 1e-3) / 1e12;
  printf("%-9s chains %d: %6.1f cycles per issue per SMSP (1 warp/SMSP) = %5.1f flop/cycle/SMSP; whole GPU %.1f TFLOP/s (%s)\n", name, C, cpi, S::FL / cpi,
         tf, cudaGetErrorString(cudaGetLastError()));
}
int main() {
  double *out; long long *cyc;
  cudaMalloc(&out, 8 * 4096); cudaMalloc(&cyc, 64);
  run<M884, 1>("m8n8k4", out, cyc); run<M884, 4>("m8n8k4", out, cyc); run<M884, 8>("m8n8k4", out, cyc);
  run<M1684, 1>("m16n8k4", out, cyc); run<M1684, 4>("m16n8k4", out, cyc);
  run<M1688, 1>("m16n8k8", out, cyc); run<M1688, 4>("m16n8k8", out, cyc);
  run<M16816, 1>("m16n8k16", out, cyc); run<M16816, 4>("m16n8k16", out, cyc);
  return 0;
}
