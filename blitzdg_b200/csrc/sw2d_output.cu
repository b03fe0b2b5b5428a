// Output path of the sw2d drivers, fed from the device.
//
// The reference downloads nothing -- it lives on the host -- but every outputInterval steps it derives eta = h - H,
// u = hu/h, v = hv/h (src/sw2d/main.cpp:199-210) and hands each to VtkOutputter::writeFieldToFile
// (include/VtkOutputter.hpp:48-101), which interpolates to the equispaced nodes of the sub-element split with
// IM = Vout * V^-1 (TriangleNodesProvisioner::splitElements, src/TriangleNodesProvisioner.cpp:1154-1227:
// resultField = sum(IM(ii,kk)*field(kk,jj),kk)).  Here one kernel does the derivation and the three interpolations
// straight from the live state buffer and leaves the result in the caller's layout ([n_out][K] row-major, the
// caller's element order); the device->host copies then run on a side stream while the time loop continues.
#include "sw2d_common.cuh"

namespace {

constexpr int kOutThreads = 128;

// EB elements per CTA.  Shared: IM [n_out][Np], derived nodal fields [3][EB][Np].
__global__ void __launch_bounds__(kOutThreads) k_output_fields(Sw2dView v, const double *__restrict__ q, const double *__restrict__ IM,
                                                               int n_out, const int32_t *__restrict__ perm, double *__restrict__ out, int EB) {
  extern __shared__ double sm[];
  const int Np = v.Np;
  const int P = Np | 1;  // odd element stride: the interpolation loop reads one node of 32 elements per wavefront
  double *sIM = sm, *sF = sm + n_out * Np;
  const long long k0 = (long long)blockIdx.x * EB;
  const int nE = (int)min((long long)EB, (long long)v.K - k0);
  for (int x = threadIdx.x; x < n_out * Np; x += blockDim.x) sIM[x] = __ldg(IM + x);
  for (int x = threadIdx.x; x < nE * Np; x += blockDim.x) {  // element-major: consecutive threads, consecutive nodes
    const long long m = k0 * Np + x;
    const double h = q[m], hu = q[v.stride + m], hv = q[2 * v.stride + m];
    const int e = x / Np, o = e * P + (x - e * Np);
    sF[o] = v.H ? h - v.H[m] : h;
    sF[EB * P + o] = hu / h;
    sF[2 * EB * P + o] = hv / h;
  }
  __syncthreads();
  const long long plane = (long long)n_out * v.K;
  for (int x = threadIdx.x; x < n_out * nE; x += blockDim.x) {  // consecutive threads, consecutive elements: row-major stores
    const int i = x / nE, e = x - i * nE;
    const double *row = sIM + i * Np, *f0 = sF + e * P, *f1 = f0 + EB * P, *f2 = f1 + EB * P;
    double a0 = 0.0, a1 = 0.0, a2 = 0.0;
    for (int j = 0; j < Np; ++j) {
      const double c = row[j];
      a0 = fma(c, f0[j], a0);
      a1 = fma(c, f1[j], a1);
      a2 = fma(c, f2[j], a2);
    }
    const long long kc = perm ? (long long)perm[k0 + e] : k0 + e;
    const long long o = (long long)i * v.K + kc;
    out[o] = a0;
    out[plane + o] = a1;
    out[2 * plane + o] = a2;
  }
}

// Positivity-preserving limiter of the Python drivers (swhelpers/limiters.py:3-21): per element
// theta = min(1, mean(h) / (mean(h) - max(min(h), 1e-3) + 1e-4)); every field is pulled towards its element mean,
// q = mean(q) + theta*(q - mean(q)).  One warp per element (Np <= 64 nodes: two per lane), shuffles for min and sums.
__global__ void __launch_bounds__(256) k_positivity_limiter(Sw2dView v, double *__restrict__ q) {
  const int lane = threadIdx.x & 31, Np = v.Np;
  const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  const double inv = 1.0 / Np;
  for (long long k = warp; k < v.K; k += nwarps) {
    const long long m0 = k * Np;
    const bool a = lane < Np, b = lane + 32 < Np;
    double x[3][2], sum[3], hmin = 1e300;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      x[i][0] = a ? q[i * v.stride + m0 + lane] : 0.0;
      x[i][1] = b ? q[i * v.stride + m0 + lane + 32] : 0.0;
      sum[i] = x[i][0] + x[i][1];
    }
    if (a) hmin = x[0][0];
    if (b) hmin = fmin(hmin, x[0][1]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      hmin = fmin(hmin, __shfl_xor_sync(0xffffffffu, hmin, o));
#pragma unroll
      for (int i = 0; i < 3; ++i) sum[i] += __shfl_xor_sync(0xffffffffu, sum[i], o);
    }
    hmin = fmax(hmin, 1e-3);
    const double hmean = sum[0] * inv;
    const double theta = fmin(1.0, hmean / (hmean - hmin + 1e-4));
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const double mean = sum[i] * inv;
      if (a) q[i * v.stride + m0 + lane] = mean + theta * (x[i][0] - mean);
      if (b) q[i * v.stride + m0 + lane + 32] = mean + theta * (x[i][1] - mean);
    }
  }
}

}  // namespace

cudaError_t launch_positivity_limiter(bdg_sw2d *s, double *q) {
  if (s->v.Np > 64) return cudaErrorInvalidValue;
  const long long warps = s->v.K;
  long long blocks = (warps + 7) / 8;
  if (blocks > 148 * 32) blocks = 148 * 32;
  k_positivity_limiter<<<(unsigned)blocks, 256, 0, s->stream>>>(s->v, q);
  s->launches++;
  return cudaGetLastError();
}

cudaError_t launch_output_fields(bdg_sw2d *s, const double *q, cudaStream_t stream) {
  const int Np = s->v.Np, n_out = s->out_n;
  int EB = 32;
  const size_t P = Np | 1;
  size_t smem = sizeof(double) * ((size_t)n_out * Np + 3 * EB * P);
  while (smem > 96 * 1024 && EB > 1) {
    EB /= 2;
    smem = sizeof(double) * ((size_t)n_out * Np + 3 * EB * P);
  }
  static size_t smem_set[16] = {0};
  if (smem > smem_set[s->device & 15]) {
    cudaError_t e = cudaFuncSetAttribute(k_output_fields, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    smem_set[s->device & 15] = smem;
  }
  const unsigned grid = (unsigned)((s->v.K + EB - 1) / EB);
  k_output_fields<<<grid, kOutThreads, smem, stream>>>(s->v, q, s->d_IM, n_out, s->d_perm, s->d_out, EB);
  s->launches++;
  return cudaGetLastError();
}
