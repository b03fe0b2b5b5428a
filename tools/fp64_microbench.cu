// fp64 pipe micro-benchmarks for B200 (sm_100a): DFMA peak, DMMA (mma.sync f64) peak,
// both together, DFMA fed by broadcast LDS.128 operands, rcp/sqrt cost, stream copy.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_microbench fp64_microbench.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <cstdint>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

constexpr int ITER = 4096;

__global__ void k_dfma(double* out, double a, double b) {
  double acc[16];
#pragma unroll
  for (int i = 0; i < 16; i++) acc[i] = threadIdx.x + i;
  for (int it = 0; it < ITER; it++) {
#pragma unroll
    for (int i = 0; i < 16; i++) acc[i] = fma(acc[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void dmma16816(double* c, const double* a, const double* b) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};\n"
               : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
               : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]),
                 "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
}
__device__ __forceinline__ void dmma1684(double* c, const double* a, double b) {
  asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};\n"
               : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
               : "d"(a[0]), "d"(a[1]), "d"(b));
}

__global__ void k_dmma884(double* out, double a, double b) {
  double c[16];
#pragma unroll
  for (int i = 0; i < 16; i++) c[i] = threadIdx.x + i;
  for (int it = 0; it < ITER; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) dmma884(c[2 * i], c[2 * i + 1], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) s += c[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_dmma16816(double* out, double a, double b) {
  double c[16], av[8], bv[4];
#pragma unroll
  for (int i = 0; i < 16; i++) c[i] = threadIdx.x + i;
#pragma unroll
  for (int i = 0; i < 8; i++) av[i] = a + i;
#pragma unroll
  for (int i = 0; i < 4; i++) bv[i] = b + i;
  for (int it = 0; it < ITER / 4; it++) {
#pragma unroll
    for (int i = 0; i < 4; i++) dmma16816(c + 4 * i, av, bv);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) s += c[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_dmma1684(double* out, double a, double b) {
  double c[16], av[2];
#pragma unroll
  for (int i = 0; i < 16; i++) c[i] = threadIdx.x + i;
  av[0] = a; av[1] = a + 1;
  for (int it = 0; it < ITER; it++) {
#pragma unroll
    for (int i = 0; i < 4; i++) dmma1684(c + 4 * i, av, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) s += c[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// half the warps DFMA, half DMMA (warp-parity split)
__global__ void k_mixed_warps(double* out, double a, double b) {
  double c[16];
#pragma unroll
  for (int i = 0; i < 16; i++) c[i] = threadIdx.x + i;
  if ((threadIdx.x >> 5) & 1) {
    for (int it = 0; it < ITER; it++) {
#pragma unroll
      for (int i = 0; i < 8; i++) dmma884(c[2 * i], c[2 * i + 1], a, b);
    }
  } else {
    for (int it = 0; it < ITER; it++) {
#pragma unroll
      for (int i = 0; i < 16; i++) c[i] = fma(c[i], a, b);
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) s += c[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// same warp interleaves 8 DMMA (=2048 MAC) with 16 DFMA (=512 MAC)
__global__ void k_mixed_inwarp(double* out, double a, double b) {
  double c[16], d[16];
#pragma unroll
  for (int i = 0; i < 16; i++) { c[i] = threadIdx.x + i; d[i] = i; }
  for (int it = 0; it < ITER; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      dmma884(c[2 * i], c[2 * i + 1], a, b);
      d[2 * i] = fma(d[2 * i], a, b);
      d[2 * i + 1] = fma(d[2 * i + 1], a, b);
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) s += c[i] + d[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// DFMA fed by broadcast LDS.128 of operator entries: TN columns per thread, 15 rows
template <int TN>
__global__ void k_lds_dfma(double* out, const double* op_g, int kdim) {
  extern __shared__ double2 sop2[];
  double* sop = reinterpret_cast<double*>(sop2);
  for (int i = threadIdx.x; i < kdim * 16; i += blockDim.x) sop[i] = op_g[i];
  __syncthreads();
  double acc[15][TN];
#pragma unroll
  for (int r = 0; r < 15; r++)
#pragma unroll
    for (int c = 0; c < TN; c++) acc[r][c] = 0.0;
  double x[TN];
#pragma unroll
  for (int c = 0; c < TN; c++) x[c] = threadIdx.x * 0.001 + c;
  for (int it = 0; it < ITER / 64; it++) {
    for (int k = 0; k < kdim; k++) {
      const double2* row = reinterpret_cast<const double2*>(sop + k * 16);
#pragma unroll
      for (int r2 = 0; r2 < 8; r2++) {
        double2 o = row[r2];
#pragma unroll
        for (int c = 0; c < TN; c++) {
          acc[2 * r2 < 15 ? 2 * r2 : 14][c] = fma(o.x, x[c], acc[2 * r2 < 15 ? 2 * r2 : 14][c]);
          if (2 * r2 + 1 < 15) acc[2 * r2 + 1][c] = fma(o.y, x[c], acc[2 * r2 + 1][c]);
        }
      }
#pragma unroll
      for (int c = 0; c < TN; c++) x[c] += 1e-9;
    }
  }
  double s = 0;
#pragma unroll
  for (int r = 0; r < 15; r++)
#pragma unroll
    for (int c = 0; c < TN; c++) s += acc[r][c];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__constant__ double c_op[45 * 16];
// DFMA with constant-bank operands, fully unrolled K=45, 15 rows, TN columns
template <int TN>
__global__ void k_const_dfma(double* out) {
  double acc[15][TN];
#pragma unroll
  for (int r = 0; r < 15; r++)
#pragma unroll
    for (int c = 0; c < TN; c++) acc[r][c] = 0.0;
  double x[TN];
#pragma unroll
  for (int c = 0; c < TN; c++) x[c] = threadIdx.x * 0.001 + c;
  for (int it = 0; it < ITER / 64; it++) {
#pragma unroll
    for (int k = 0; k < 45; k++) {
#pragma unroll
      for (int r = 0; r < 15; r++)
#pragma unroll
        for (int c = 0; c < TN; c++) acc[r][c] = fma(c_op[k * 16 + r], x[c], acc[r][c]);
#pragma unroll
      for (int c = 0; c < TN; c++) x[c] += 1e-9;
    }
  }
  double s = 0;
#pragma unroll
  for (int r = 0; r < 15; r++)
#pragma unroll
    for (int c = 0; c < TN; c++) s += acc[r][c];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// pointwise cost: mode 0 = '/' ; 1 = sqrt ; 2 = rcp.approx+2 newton ; 3 = rsqrt.approx + newton
__device__ __forceinline__ double fast_rcp(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  double e = fma(-x, r, 1.0);
  r = fma(r, e, r);
  e = fma(-x, r, 1.0);
  r = fma(r, e, r);
  return r;
}
__device__ __forceinline__ double fast_sqrt(double x) {
  double r;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  // two Newton steps on rsqrt then one correction on sqrt
  double h = 0.5 * r;
  double s = x * r;
  double e = fma(-h, s, 0.5);
  s = fma(s, e, s);
  h = fma(h, e, h);
  e = fma(-h, s, 0.5);
  s = fma(s, e, s);
  h = fma(h, e, h);
  double d = fma(-s, s, x);
  s = fma(d, h, s);
  return s;
}
template <int MODE>
__global__ void k_pointwise(double* out, double seed) {
  double v[8];
#pragma unroll
  for (int i = 0; i < 8; i++) v[i] = seed + threadIdx.x * 0.01 + i;
  for (int it = 0; it < ITER / 8; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      if (MODE == 0) v[i] = 1.0 / v[i] + 1.5;
      if (MODE == 1) v[i] = sqrt(v[i]) + 1.5;
      if (MODE == 2) v[i] = fast_rcp(v[i]) + 1.5;
      if (MODE == 3) v[i] = fast_sqrt(v[i]) + 1.5;
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) s += v[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_copy(const double2* __restrict__ a, double2* __restrict__ b, size_t n) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) b[i] = a[i];
}

template <typename F>
float timeit(F f, int reps = 5) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  f(); CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int r = 0; r < reps; r++) {
    cudaEventRecord(e0); f(); cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
  }
  return best;
}

int main() {
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  int clk = 0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  printf("device %s SMs %d clock %d kHz\n", p.name, p.multiProcessorCount, clk);
  int nsm = p.multiProcessorCount;
  double* out; CK(cudaMalloc(&out, sizeof(double) * nsm * 8 * 1024));
  for (int wpb : {4, 8, 16}) {  // warps per block, 2 blocks per SM
    int threads = wpb * 32, blocks = nsm * 2;
    double thr = (double)threads * blocks;
    float ms = timeit([&] { k_dfma<<<blocks, threads>>>(out, 1.0000001, 1e-9); });
    printf("DFMA  warps/SM %2d: %.2f TFLOP/s\n", wpb * 2, thr * ITER * 16 * 2 / ms / 1e9);
    ms = timeit([&] { k_dmma884<<<blocks, threads>>>(out, 1.0000001, 1e-9); });
    printf("DMMA m8n8k4  warps/SM %2d: %.2f TFLOP/s\n", wpb * 2, thr / 32 * ITER * 8 * 512 / ms / 1e9);
    ms = timeit([&] { k_dmma1684<<<blocks, threads>>>(out, 1.0000001, 1e-9); });
    printf("DMMA m16n8k4 warps/SM %2d: %.2f TFLOP/s\n", wpb * 2, thr / 32 * ITER * 4 * 1024 / ms / 1e9);
    ms = timeit([&] { k_dmma16816<<<blocks, threads>>>(out, 1.0000001, 1e-9); });
    printf("DMMA m16n8k16 warps/SM %2d: %.2f TFLOP/s\n", wpb * 2, thr / 32 * (ITER / 4) * 4 * 4096 / ms / 1e9);
    ms = timeit([&] { k_mixed_warps<<<blocks, threads>>>(out, 1.0000001, 1e-9); });
    printf("MIXED warps(half DFMA half DMMA) warps/SM %2d: %.2f TFLOP/s (DFMA part %.2f + DMMA part %.2f if perfectly parallel)\n", wpb * 2,
           (thr / 2 * ITER * 16 * 2 + thr / 64 * ITER * 8 * 512) / ms / 1e9, thr / 2 * ITER * 16 * 2 / ms / 1e9, thr / 64 * ITER * 8 * 512 / ms / 1e9);
    ms = timeit([&] { k_mixed_inwarp<<<blocks, threads>>>(out, 1.0000001, 1e-9); });
    printf("MIXED in-warp warps/SM %2d: %.2f TFLOP/s (dmma %.2f dfma %.2f)\n", wpb * 2,
           (thr * ITER * 16 * 2 + thr / 32 * ITER * 8 * 512) / ms / 1e9, thr / 32 * ITER * 8 * 512 / ms / 1e9, thr * ITER * 16 * 2 / ms / 1e9);
  }
  {
    double h_op[45 * 16];
    for (int i = 0; i < 45 * 16; i++) h_op[i] = 1.0 / (i + 1);
    double* d_op; CK(cudaMalloc(&d_op, sizeof(h_op))); CK(cudaMemcpy(d_op, h_op, sizeof(h_op), cudaMemcpyHostToDevice));
    CK(cudaMemcpyToSymbol(c_op, h_op, sizeof(h_op)));
    for (int wpb : {2, 4, 8}) {
      int threads = wpb * 32, blocks = nsm * 2;
      double thr = (double)threads * blocks;
      double macs = thr * (ITER / 64) * 45 * 15;
      float ms = timeit([&] { k_lds_dfma<1><<<blocks, threads, 45 * 16 * 8>>>(out, d_op, 45); });
      printf("LDS.128-bcast DFMA TN=1 warps/SM %2d: %.2f TFLOP/s\n", wpb * 2, macs * 1 * 2 / ms / 1e9);
      ms = timeit([&] { k_lds_dfma<2><<<blocks, threads, 45 * 16 * 8>>>(out, d_op, 45); });
      printf("LDS.128-bcast DFMA TN=2 warps/SM %2d: %.2f TFLOP/s\n", wpb * 2, macs * 2 * 2 / ms / 1e9);
      ms = timeit([&] { k_lds_dfma<3><<<blocks, threads, 45 * 16 * 8>>>(out, d_op, 45); });
      printf("LDS.128-bcast DFMA TN=3 warps/SM %2d: %.2f TFLOP/s\n", wpb * 2, macs * 3 * 2 / ms / 1e9);
      ms = timeit([&] { k_const_dfma<1><<<blocks, threads>>>(out); });
      printf("const-operand DFMA TN=1 warps/SM %2d: %.2f TFLOP/s\n", wpb * 2, macs * 1 * 2 / ms / 1e9);
      ms = timeit([&] { k_const_dfma<3><<<blocks, threads>>>(out); });
      printf("const-operand DFMA TN=3 warps/SM %2d: %.2f TFLOP/s\n", wpb * 2, macs * 3 * 2 / ms / 1e9);
    }
  }
  {
    int threads = 256, blocks = nsm * 4;
    double ops = (double)threads * blocks * ITER;
    float ms = timeit([&] { k_pointwise<0><<<blocks, threads>>>(out, 1.3); });
    printf("1/x        : %.1f Gop/s  (equiv DFMA slots per op at 18.5T DFMA/s: %.1f)\n", ops / ms / 1e6, 18.5e3 / (ops / ms / 1e6));
    ms = timeit([&] { k_pointwise<1><<<blocks, threads>>>(out, 1.3); });
    printf("sqrt       : %.1f Gop/s  (%.1f)\n", ops / ms / 1e6, 18.5e3 / (ops / ms / 1e6));
    ms = timeit([&] { k_pointwise<2><<<blocks, threads>>>(out, 1.3); });
    printf("fast_rcp   : %.1f Gop/s  (%.1f)\n", ops / ms / 1e6, 18.5e3 / (ops / ms / 1e6));
    ms = timeit([&] { k_pointwise<3><<<blocks, threads>>>(out, 1.3); });
    printf("fast_sqrt  : %.1f Gop/s  (%.1f)\n", ops / ms / 1e6, 18.5e3 / (ops / ms / 1e6));
  }
  {
    size_t n = (size_t)1 << 27;  // 2 GiB per buffer of double2
    double2 *a, *b; CK(cudaMalloc(&a, n * 16)); CK(cudaMalloc(&b, n * 16));
    CK(cudaMemset(a, 1, n * 16));
    for (int bps : {4, 8, 16, 32}) {
      float ms = timeit([&] { k_copy<<<nsm * bps, 256>>>(a, b, n); });
      printf("copy kernel blocks/SM %2d: %.1f GB/s\n", bps, 2.0 * n * 16 / ms / 1e6);
    }
    float ms = timeit([&] { cudaMemcpyAsync(b, a, n * 16, cudaMemcpyDeviceToDevice); });
    printf("cudaMemcpy D2D: %.1f GB/s\n", 2.0 * n * 16 / ms / 1e6);
  }
  return 0;
}
