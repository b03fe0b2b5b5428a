// Over-integrated RHS on curved triangles: sw2dComputeRHS_curved of the reference's Python
// drivers (swhelpers/rhs.py:6-176, fluxes swhelpers/flux.py:1-23, driver sw2d_curved.py:240-262) on the device state.
//
// A different kernel shape from the nodal path (sw2d_fused.cu): the fluxes are evaluated at the cubature points of the
// element and at the Gauss nodes of its faces, tested against the nodal basis (weak form), and the element's own mass
// matrix is inverted -- V V^T / J on straight elements, two triangular solves with the stored Cholesky factor on
// curved ones.  Two launches per evaluation:
//   k_curved_gauss  the four conserved fields at the Gauss face nodes of every element (the '+' side of a face node
//                   lives in the neighbour): gq[field][g + 3*NGauss*k]
//   k_curved_rhs    one CTA per element: cubature volume term, Lax-Friedrichs surface term with the maximum speed over
//                   the Gauss nodes of a face, mass-matrix solve, nodal sources
// A generic kernel (any order, any cubature), operators read through the read-only cache; it is a parity kernel for the
// Python drivers' curved workflow, not a roofline kernel: ~25 k FMAs per element at N = 4 against 1.5 KB of state.
//
// Arithmetic follows the reference function: '+' velocities of the wave speed are taken before the wall reflection
// (rhs.py:80-87), the drag enters RHS3 with the sign the reference gives it (rhs.py:171: RHS3 -= f hu - CD |u| v).
#include "sw2d_common.cuh"

// The element arithmetic is written once, as __host__ __device__ functions of (thread, element): the kernels below run it
// with 64 threads per element, tests/test_curved_rhs.py compiles the same functions for the host with ONE thread per
// element (-DBDG_CURVED_HOST_EMULATION, tools/curved_host_emulation.sh) and holds them to the oracle without a GPU.
#if defined(__CUDA_ARCH__)
#define CURVED_SYNC() __syncthreads()
#define CURVED_LDG(p) __ldg(p)
#else
#define CURVED_SYNC() ((void)0)
#define CURVED_LDG(p) (*(p))
#endif

namespace {

constexpr int kCurvedThreads = 64;

// fields at Gauss node i = g + NG3*k (rhs.py:57-71)
__host__ __device__ __forceinline__ void gauss_node(long long i, const double *__restrict__ q, long long stride, const double *__restrict__ qN,
                                                    const double *__restrict__ Interp, double *__restrict__ gq, int Np, int NG3, long long K) {
  const long long n = (long long)NG3 * K;
  const long long k = i / NG3;
  const int g = (int)(i - k * NG3);
  double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
  for (int j = 0; j < Np; ++j) {
    const double w = CURVED_LDG(Interp + g * Np + j);
    const long long m = k * Np + j;
    a0 += w * q[m];
    a1 += w * q[stride + m];
    a2 += w * q[2 * stride + m];
    if (qN) a3 += w * qN[m];
  }
  gq[i] = a0; gq[n + i] = a1; gq[2 * n + i] = a2; gq[3 * n + i] = a3;
}

__global__ void k_curved_gauss(const double *__restrict__ q, long long stride, const double *__restrict__ qN, const double *__restrict__ Interp,
                               double *__restrict__ gq, int Np, int NG3, long long K) {
  const long long n = (long long)NG3 * K;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    gauss_node(i, q, stride, qN, Interp, gq, Np, NG3, K);
}

struct Flux4 {
  double F[4], G[4];
};
__host__ __device__ __forceinline__ Flux4 fluxes(double h, double hu, double hv, double hN, double g) {  // swhelpers/flux.py:1-23
  Flux4 x;
  const double u = hu / h, v = hv / h, half = 0.5 * g * h * h;
  x.F[0] = hu;            x.G[0] = hv;
  x.F[1] = hu * u + half; x.G[1] = hu * v;
  x.F[2] = hv * u;        x.G[2] = hv * v + half;
  x.F[3] = hN * u;        x.G[3] = hN * v;
  return x;
}

// thread t of NT working on element k; sm: 12*Np + 8*Ncub + 5*NG3 doubles shared by the NT threads
template <int NT>
__host__ __device__ __forceinline__ void curved_element(const Sw2dView &v, const CurvedOps &c, const double *__restrict__ q,
                                                        const double *__restrict__ qN, const double *__restrict__ CDn, double *__restrict__ out,
                                                        double *sm, int t, long long k) {
  constexpr int kCurvedThreads = NT;
  const int Np = v.Np, Ncub = c.Ncub, NG3 = c.NG3, Ng = c.NG3 / 3;
  const long long nT = (long long)Np * v.K, nG = (long long)NG3 * v.K;
  double *sq = sm;                 // [4][Np]   nodal state of the element
  double *tr = sq + 4 * Np;        // [4][Ncub] W (rx F + ry G)
  double *ts = tr + 4 * Ncub;      // [4][Ncub] W (sx F + sy G)
  double *gf = ts + 4 * Ncub;      // [4][NG3]  W_g * numerical flux
  double *spd = gf + 4 * NG3;      // [NG3]     max(speed-, speed+) per Gauss node
  double *b = spd + NG3;           // [4][Np]   MM * RHS
  double *xs = b + 4 * Np;         // [4][Np]   RHS before the sources
  const double g = v.g;

  for (int o = t; o < 4 * Np; o += kCurvedThreads) {
    const int f = o / Np, j = o - f * Np;
    sq[o] = f < 3 ? q[f * v.stride + k * Np + j] : (qN ? qN[k * Np + j] : 0.0);
  }
  // pass 1 over the Gauss nodes: wave speeds of both traces (rhs.py:75-96)
  for (int gi = t; gi < NG3; gi += kCurvedThreads) {
    const long long i = gi + k * NG3, iM = c.mapM[i], iP = c.mapP[i];
    const double hM = c.gq[iM], huM = c.gq[nG + iM], hvM = c.gq[2 * nG + iM];
    const double hP = c.gq[iP], huP = c.gq[nG + iP], hvP = c.gq[2 * nG + iP];
    const double uM = huM / hM, vM = hvM / hM, uP = huP / hP, vP = hvP / hP;  // '+' velocities BEFORE the wall reflection
    const double sM = sqrt(uM * uM + vM * vM) + sqrt(g * hM), sP = sqrt(uP * uP + vP * vP) + sqrt(g * hP);
    spd[gi] = sM > sP ? sM : sP;
  }
  CURVED_SYNC();
  // cubature points: interpolate, fluxes, weak-form integrands (rhs.py:8-47)
  for (int p = t; p < Ncub; p += kCurvedThreads) {
    double cq[4] = {0.0, 0.0, 0.0, 0.0};
    for (int j = 0; j < Np; ++j) {
      const double w = CURVED_LDG(c.V + p * Np + j);
#pragma unroll
      for (int f = 0; f < 4; ++f) cq[f] += w * sq[f * Np + j];
    }
    const Flux4 x = fluxes(cq[0], cq[1], cq[2], cq[3], g);
    const long long cp = k * Ncub + p;
    const double W = c.cW[cp], rx = c.crx[cp], sx = c.csx[cp], ry = c.cry[cp], sy = c.csy[cp];
#pragma unroll
    for (int f = 0; f < 4; ++f) {
      tr[f * Ncub + p] = W * (rx * x.F[f] + ry * x.G[f]);
      ts[f * Ncub + p] = W * (sx * x.F[f] + sy * x.G[f]);
    }
  }
  // pass 2 over the Gauss nodes: wall reflection, local Lax-Friedrichs flux with the face maximum (rhs.py:84-132)
  for (int gi = t; gi < NG3; gi += kCurvedThreads) {
    const long long i = gi + k * NG3, iM = c.mapM[i], iP = c.mapP[i];
    const double hM = c.gq[iM], huM = c.gq[nG + iM], hvM = c.gq[2 * nG + iM], hNM = c.gq[3 * nG + iM];
    const double hP = c.gq[iP], hNP = c.gq[3 * nG + iP];
    double huP = c.gq[nG + iP], hvP = c.gq[2 * nG + iP];
    const double nx = c.gnx[i], ny = c.gny[i];
    if (c.wall[i]) {
      const double dn = huM * nx + hvM * ny;
      huP = huM - 2 * nx * dn;
      hvP = hvM - 2 * ny * dn;
    }
    const Flux4 M = fluxes(hM, huM, hvM, hNM, g), P = fluxes(hP, huP, hvP, hNP, g);
    const int face = gi / Ng;
    double lam = 0.0;
    for (int j = face * Ng; j < (face + 1) * Ng; ++j) lam = spd[j] > lam ? spd[j] : lam;
    const double dq[4] = {hM - hP, huM - huP, hvM - hvP, hNM - hNP};
    const double W = c.gW[i];
#pragma unroll
    for (int f = 0; f < 4; ++f) gf[f * NG3 + gi] = W * (0.5 * ((M.F[f] + P.F[f]) * nx + (M.G[f] + P.G[f]) * ny + lam * dq[f]));
  }
  CURVED_SYNC();
  // test against the nodal basis: b = Dr_c^T tr + Ds_c^T ts - Interp^T gf (rhs.py:26-47, 135-139)
  for (int o = t; o < 4 * Np; o += kCurvedThreads) {
    const int f = o / Np, i = o - f * Np;
    double ar = 0.0, as = 0.0, ag = 0.0;
    for (int p = 0; p < Ncub; ++p) {
      ar += CURVED_LDG(c.Dr + p * Np + i) * tr[f * Ncub + p];
      as += CURVED_LDG(c.Ds + p * Np + i) * ts[f * Ncub + p];
    }
    for (int gi = 0; gi < NG3; ++gi) ag += CURVED_LDG(c.gInterp + gi * Np + i) * gf[f * NG3 + gi];
    b[o] = (ar + as) - ag;
  }
  CURVED_SYNC();
  // mass matrix (rhs.py:141-164)
  const int ci = c.chol_idx[k];
  if (ci < 0) {
    for (int o = t; o < 4 * Np; o += kCurvedThreads) {
      const int f = o / Np, i = o - f * Np;
      double a = 0.0;
      for (int j = 0; j < Np; ++j) a += CURVED_LDG(c.mminv + i * Np + j) * (b[f * Np + j] / c.J[k * Np + j]);
      xs[o] = a;
    }
  } else {  // U^T y = b, U x = y with the element's upper Cholesky factor; one field per thread
    for (int f = t; f < 4; f += kCurvedThreads) {
    const double *U = c.chol + (long long)ci * Np * Np;
    double *y = xs + f * Np;
    const double *bb = b + f * Np;
    for (int i = 0; i < Np; ++i) {
      double a = bb[i];
      for (int j = 0; j < i; ++j) a -= U[j * Np + i] * y[j];
      y[i] = a / U[i * Np + i];
    }
    for (int i = Np - 1; i >= 0; --i) {
      double a = y[i];
      for (int j = i + 1; j < Np; ++j) a -= U[i * Np + j] * y[j];
      y[i] = a / U[i * Np + i];
    }
    }
  }
  CURVED_SYNC();
  // nodal sources (rhs.py:166-176) and store, element-major like the state
  for (int i = t; i < Np; i += kCurvedThreads) {
    const long long m = k * Np + i;
    const double h = sq[i], hu = sq[Np + i], hv = sq[2 * Np + i];
    const double u = hu / h, w = hv / h, nrm = sqrt(u * u + w * w);
    const double CD = CDn ? CDn[m] : v.CD;
    const double zx = v.Hx ? -v.Hx[m] : 0.0, zy = v.Hy ? -v.Hy[m] : 0.0;  // z = -H
    double r2 = xs[Np + i], r3 = xs[2 * Np + i];
    r2 += v.fcor * hv - CD * nrm * u;
    r3 -= v.fcor * hu - CD * nrm * w;
    r2 -= g * h * zx;
    r3 -= g * h * zy;
    out[m] = xs[i];
    out[nT + m] = r2;
    out[2 * nT + m] = r3;
    out[3 * nT + m] = xs[3 * Np + i];
  }
}

// One stage of the driver's midpoint scheme (sw2d_curved.py:244-262): dst = base + c * (Filt * R) at node (i, k), all four
// fields; Filt = NULL: no filter.  R [4][nT] element-major as k_curved_rhs wrote it; q planes are `stride` apart.
__host__ __device__ __forceinline__ void curved_update_node(long long m, int Np, long long nT, long long stride, const double *__restrict__ Filt,
                                                            const double *__restrict__ R, double c, const double *base, const double *baseN,
                                                            double *dst, double *dstN) {  // dst may be base (corrector stage)
  const long long k = m / Np;
  const int i = (int)(m - k * Np);
#pragma unroll
  for (int f = 0; f < 4; ++f) {
    if (f == 3 && !dstN) break;
    double a;
    if (Filt) {
      a = 0.0;
      for (int j = 0; j < Np; ++j) a += CURVED_LDG(Filt + i * Np + j) * R[f * nT + k * Np + j];
    } else {
      a = R[f * nT + m];
    }
    if (f < 3) dst[f * stride + m] = base[f * stride + m] + c * a;
    else dstN[m] = baseN[m] + c * a;
  }
}

__global__ void k_curved_update(int Np, long long nT, long long stride, const double *__restrict__ Filt, const double *__restrict__ R, double c,
                                const double *base, const double *baseN, double *dst, double *dstN) {
  for (long long m = blockIdx.x * (long long)blockDim.x + threadIdx.x; m < nT; m += (long long)gridDim.x * blockDim.x)
    curved_update_node(m, Np, nT, stride, Filt, R, c, base, baseN, dst, dstN);
}

__global__ void __launch_bounds__(kCurvedThreads) k_curved_rhs(Sw2dView v, CurvedOps c, const double *__restrict__ q, const double *__restrict__ qN,
                                                              const double *__restrict__ CDn, double *__restrict__ out) {
  extern __shared__ double sm[];
  curved_element<kCurvedThreads>(v, c, q, qN, CDn, out, sm, threadIdx.x, blockIdx.x);
}

}  // namespace

#ifdef BDG_CURVED_HOST_EMULATION
// Test hook of tests/test_curved_rhs.py, compiled into its own little library and never into libblitzdg_b200.so: the element
// functions above on host arrays in the device layouts (element-major, identity element order), one thread per element.
extern "C" void bdg_curved_host_emulation(int Np, int K, long long stride, double g, double fcor, double CD, const double *Hx, const double *Hy,
                                          const CurvedOps *c, const double *q, const double *qN, const double *CDn, double *sm, double *out) {
  Sw2dView v = {};
  v.Np = Np; v.K = K; v.stride = stride; v.g = g; v.fcor = fcor; v.CD = CD; v.Hx = Hx; v.Hy = Hy;
  for (long long i = 0; i < (long long)c->NG3 * K; ++i) gauss_node(i, q, stride, qN, c->gInterp, c->gq, Np, c->NG3, K);
  for (long long k = 0; k < K; ++k) curved_element<1>(v, *c, q, qN, CDn, out, sm, 0, k);
}
extern "C" void bdg_curved_update_host_emulation(int Np, int K, long long stride, const double *Filt, const double *R, double c, const double *base,
                                                 const double *baseN, double *dst, double *dstN) {
  for (long long m = 0; m < (long long)Np * K; ++m) curved_update_node(m, Np, (long long)Np * K, stride, Filt, R, c, base, baseN, dst, dstN);
}
#endif

size_t curved_smem_bytes(int Np, int Ncub, int NG3) { return sizeof(double) * (size_t)(12 * Np + 8 * Ncub + 5 * NG3); }

// q: live flow state [3][stride]; qN: tracer [Np*K] or NULL; CDn: nodal drag coefficient [Np*K] (device order) or NULL =
// prm.CD; out: [4][Np*K] element-major in device element order
cudaError_t launch_curved_rhs(bdg_sw2d *s, const double *q, const double *qN, const double *CDn, double *out) {
  const CurvedOps &c = *s->curved;
  const long long n = (long long)c.NG3 * s->v.K;
  const int grid = (int)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8);
  k_curved_gauss<<<grid, 256, 0, s->stream>>>(q, s->v.stride, qN, c.gInterp, c.gq, s->v.Np, c.NG3, s->v.K);
  s->launches++;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  const size_t smem = curved_smem_bytes(s->v.Np, c.Ncub, c.NG3);
  if (smem > 48 * 1024) {
    if ((e = cudaFuncSetAttribute(k_curved_rhs, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;
  }
  k_curved_rhs<<<(unsigned)s->v.K, kCurvedThreads, smem, s->stream>>>(s->v, c, q, qN, CDn, out);
  s->launches++;
  return cudaGetLastError();
}

// dst = base + c * (Filt * R) for the flow planes and, if dstN, the tracer (in place allowed: dst == base)
cudaError_t launch_curved_update(bdg_sw2d *s, const double *Filt, const double *R, double c, const double *base, const double *baseN, double *dst,
                                 double *dstN) {
  const long long nT = (long long)s->nT;
  const int grid = (int)((nT + 255) / 256 < 148 * 8 ? (nT + 255) / 256 : 148 * 8);
  k_curved_update<<<grid, 256, 0, s->stream>>>(s->v.Np, nT, s->v.stride, Filt, R, c, base, baseN, dst, dstN);
  s->launches++;
  return cudaGetLastError();
}
