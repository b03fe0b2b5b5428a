// Generic sw2d kernels: any element shape / order / per-node metrics / full physics / filter.
// Two launches per stage (face fluxes -> volume + lift + LSERK4 update).  This is the path of
// record for quads, non-affine meshes, the filter and orders without a fused specialisation;
// the affine-triangle fused DMMA kernel lives in sw2d_fused.cu.
//
// Arithmetic follows SURVEY.md Appendix A (reference src/sw2d/main.cpp:279-484,
// src/sw2d-simple/main.cpp:181-356) node for node; only the sweep structure differs.
#include <cstdlib>

#include "sw2d_common.cuh"

namespace {

constexpr double kPi = 3.1415926535897932;  // reference include/Constants.hpp

struct Trace {
  double hM, huM, hvM, hP, huP, hvP, nx, ny;
};

// 1/x by MUFU seed + two Newton steps (x = 0 gives NaN like the reference's 0/0)
__device__ __forceinline__ double frcp(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  double e = fma(-x, r, 1.0);
  r = fma(r, e, r);
  e = fma(-x, r, 1.0);
  return fma(r, e, r);
}

// sqrt(x), x >= 0: coupled Newton on (sqrt, 1/(2 sqrt)) from the MUFU rsqrt seed + one correction (as in sw2d_fused.cu);
// x = 0 -> 0, NaN -> NaN.  The library sqrt/divide sequences cost 2-3x the fp64-pipe slots of these.
__device__ __forceinline__ double fsqrt(double x) {
  double r;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  if (x == 0.0) r = 0.0;  // rsqrt(0) = inf would turn the products into NaN
  double hh = 0.5 * r, s = x * r;
  double e = fma(-hh, s, 0.5);
  s = fma(s, e, s);
  e = fma(-s, s, x);
  return fma(e, hh, s);
}

// A1-A4 of Appendix A: gather, wall / open BC, star states for face-node slot sl of element k
// NEED_N = false (speed-only callers): the normal is fetched for wall nodes only, 16 B per face node less.
template <bool NEED_N = true>
__device__ __forceinline__ Trace load_trace(const Sw2dView &v, const double *__restrict__ q, int k, int sl, double t) {
  const long long i = (long long)k * v.Nfn + sl;
  const long long mM = (long long)k * v.Np + v.fmask[sl];
  const long long mP = v.vmapP[i];
  const double *h = q, *hu = q + v.stride, *hv = q + 2 * v.stride;
  Trace T;
  T.hM = h[mM]; T.huM = hu[mM]; T.hvM = hv[mM];
  T.hP = h[mP]; T.huP = hu[mP]; T.hvP = hv[mP];
  const int tag = v.bctag[i];
  if (NEED_N || tag == 3) { T.nx = v.nx[i]; T.ny = v.ny[i]; } else { T.nx = T.ny = 0.0; }
  if (tag == 3) {  // Wall: reflect the normal momentum (main.cpp:340-345)
    if (v.full) T.hP = T.hM;
    const double d = T.huM * T.nx + T.hvM * T.ny;
    T.huP = T.huM - 2 * T.nx * d;
    T.hvP = T.hvM - 2 * T.ny * d;
  }
  if (v.full) {
    const double HM = v.H[mM], HP = v.H[mP];
    if (tag == 2) {  // Out: tidal free surface (main.cpp:348-353)
      const double om = 2.0 * kPi / v.tide_period;
      T.huP = T.huM; T.hvP = T.hvM;
      T.hP = HM + v.tide_amp * cos(om * t) * 0.5 * (tanh(v.tide_ramp * (t - v.tide_period)) + 1);
    }
    const double bM = -HM, bP = -HP, bmax = fmax(bP, bM);  // main.cpp:359-368
    const double hMs = fmax(0.0, T.hM + bM - bmax), hPs = fmax(0.0, T.hP + bP - bmax);
    // main.cpp:366-368: hu = h* (hu/h) is evaluated AFTER h = h* (SURVEY.md a5), an identity up to one rounding --
    // except on a dry trace (h* = 0), where 0*(hu/0) is NaN and trips the blow-up guard; four divides saved.
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    if (!(hMs > 0.0)) T.huM = T.hvM = nan;
    if (!(hPs > 0.0)) T.huP = T.hvP = nan;
    T.hM = hMs; T.hP = hPs;
  }
  return T;
}

__device__ __forceinline__ double trace_speed(const Trace &T, double g) {  // A7
  const double rM = frcp(T.hM), rP = frcp(T.hP);
  const double uM = T.huM * rM, vM = T.hvM * rM, uP = T.huP * rP, vP = T.hvP * rP;
  const double sM = fsqrt(uM * uM + vM * vM) + fsqrt(g * T.hM);
  const double sP = fsqrt(uP * uP + vP * vP) + fsqrt(g * T.hP);
  return sM > sP ? sM : sP;
}

// max of non-negative doubles through their bit patterns (speeds are >= 0; NaN compares largest and so
// propagates into the blow-up guard)
__device__ __forceinline__ void atomic_max_pos(double *addr, double val) {
  atomicMax(reinterpret_cast<unsigned long long *>(addr), (unsigned long long)__double_as_longlong(val));
}

__device__ __forceinline__ double block_max(double val) {
  __shared__ double red[32];
  for (int o = 16; o > 0; o >>= 1) {
    double other = __shfl_xor_sync(0xffffffffu, val, o);
    val = (__double_as_longlong(other) > __double_as_longlong(val)) ? other : val;
  }
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) red[w] = val;
  __syncthreads();
  if (w == 0) {
    val = (l < (int)((blockDim.x + 31) >> 5)) ? red[l] : 0.0;
    for (int o = 16; o > 0; o >>= 1) {
      double other = __shfl_xor_sync(0xffffffffu, val, o);
      val = (__double_as_longlong(other) > __double_as_longlong(val)) ? other : val;
    }
  }
  return val;  // valid in thread 0
}

// global Lax-Friedrichs speed (main.cpp:413-414): max over all face nodes into scal[slot]
__global__ void k_speed(Sw2dView v, const double *__restrict__ q, int slot) {
  const long long nF = (long long)v.Nfn * v.K;
  const double t = v.scal[S_T];
  double m = 0.0;
  // (unrolling this loop for more gathers in flight costs occupancy and was slower: 236 vs 186 us)
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nF; i += (long long)gridDim.x * blockDim.x) {
    // The '+' side of an interior face node is the '-' side of the neighbour's (same pair of values, same star states),
    // which that neighbour's thread -- or the neighbour rank, before the all-reduce -- evaluates: only BC faces need
    // both sides here.  Halves the gathers and the sqrt/reciprocal work of this launch.
    const unsigned iu = (unsigned)i;  // Nfn*K < 2^31: 32-bit division, not the 64-bit software routine
    const int k = (int)(iu / (unsigned)v.Nfn), sl = (int)(iu - (unsigned)k * (unsigned)v.Nfn);
    double s;
    if (v.bctag[i] != 0) {
      Trace T = load_trace<false>(v, q, k, sl, t);
      s = trace_speed(T, v.g);
    } else {
      const long long mM = (long long)k * v.Np + v.fmask[sl];
      double hM = q[mM];
      double huM = q[v.stride + mM], hvM = q[2 * v.stride + mM];
      if (v.full) {  // star state of the '-' side (main.cpp:359-368), as in load_trace
        const double bM = -v.H[mM], bP = -v.H[v.vmapP[i]];
        hM = fmax(0.0, hM + bM - fmax(bP, bM));
        if (!(hM > 0.0)) huM = hvM = __longlong_as_double(0x7ff8000000000000LL);
      }
      const double r = frcp(hM), uM = huM * r, vM = hvM * r;
      s = fsqrt(uM * uM + vM * vM) + fsqrt(v.g * hM);
    }
    m = (__double_as_longlong(s) > __double_as_longlong(m)) ? s : m;
  }
  m = block_max(m);
  if (threadIdx.x == 0) atomic_max_pos(&v.scal[slot], m);
}

// A5-A8: one thread per face; writes sflux[c][i] = Fscale*dFlux_c
__global__ void k_face(Sw2dView v, const double *__restrict__ q, double *__restrict__ sflux, int k0, int k1, int slot) {
  const long long nfaces = (long long)(k1 - k0) * v.Nf;
  const long long fid = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (blockIdx.x == 0 && threadIdx.x == 0 && v.lf_global) v.scal[slot ^ 1] = 0.0;  // re-arm the other slot (slots 2,3)
  if (fid >= nfaces) return;
  const int k = k0 + (int)(fid / v.Nf), f = (int)(fid % v.Nf);
  const double t = v.scal[S_T], g = v.g;
  const long long nF = (long long)v.Nfn * v.K;
  double lam;
  if (v.lf_global) {
    lam = v.scal[slot];
  } else {  // per-face max (sw2d-simple main.cpp:305-312)
    lam = 0.0;
    for (int j = 0; j < v.Nfp; ++j) {
      Trace T = load_trace(v, q, k, f * v.Nfp + j, t);
      double s = trace_speed(T, g);
      lam = (j == 0 || s > lam) ? s : lam;
    }
  }
  for (int j = 0; j < v.Nfp; ++j) {
    const int sl = f * v.Nfp + j;
    const long long i = (long long)k * v.Nfn + sl;
    Trace T = load_trace(v, q, k, sl, t);
    const double dh = T.hM - T.hP, dhu = T.huM - T.huP, dhv = T.hvM - T.hvP;
    const double F2M = (T.huM * T.huM) / T.hM + 0.5 * g * T.hM * T.hM, G2M = (T.huM * T.hvM) / T.hM,
                 G3M = (T.hvM * T.hvM) / T.hM + 0.5 * g * T.hM * T.hM;
    const double F2P = (T.huP * T.huP) / T.hP + 0.5 * g * T.hP * T.hP, G2P = (T.huP * T.hvP) / T.hP,
                 G3P = (T.hvP * T.hvP) / T.hP + 0.5 * g * T.hP * T.hP;
    const double dF1 = 0.5 * ((T.huM - T.huP) * T.nx + (T.hvM - T.hvP) * T.ny - lam * dh);
    const double dF2 = 0.5 * ((F2M - F2P) * T.nx + (G2M - G2P) * T.ny - lam * dhu);
    const double dF3 = 0.5 * ((G2M - G2P) * T.nx + (G3M - G3P) * T.ny - lam * dhv);
    const double fs = v.Fscale[i];
    sflux[i] = fs * dF1;
    sflux[nF + i] = fs * dF2;
    sflux[2 * nF + i] = fs * dF3;
  }
}

// A5-A8, one thread per face NODE (coalesced trace-map / geometry reads and sflux writes; every trace is loaded
// once): wave speeds go through shared memory for the per-face maximum (sw2d-simple main.cpp:305-312).
// EB elements per CTA, thread = e*Nfn + slot.
__global__ void k_face_nodes(Sw2dView v, const double *__restrict__ q, double *__restrict__ sflux, int k0, int k1, int slot, int EB) {
  extern __shared__ double sspd[];
  const int Nfn = v.Nfn, Nfp = v.Nfp, tid = threadIdx.x;
  const int e = tid / Nfn, sl = tid - e * Nfn;
  const long long k = k0 + (long long)blockIdx.x * EB + e;
  const bool active = e < EB && k < k1;
  if (blockIdx.x == 0 && tid == 0 && v.lf_global) v.scal[slot ^ 1] = 0.0;  // re-arm the other slot (slots 2,3)
  const double g = v.g;
  Trace T;
  double rM = 0.0, rP = 0.0;
  if (active) {
    T = load_trace(v, q, (int)k, sl, v.scal[S_T]);
    rM = frcp(T.hM);
    rP = frcp(T.hP);
    const double uM = T.huM * rM, vM = T.hvM * rM, uP = T.huP * rP, vP = T.hvP * rP;
    const double sM = fsqrt(uM * uM + vM * vM) + fsqrt(g * T.hM), sP = fsqrt(uP * uP + vP * vP) + fsqrt(g * T.hP);
    sspd[tid] = sM > sP ? sM : sP;
  }
  __syncthreads();
  if (!active) return;
  double lam;
  if (v.lf_global) {
    lam = v.scal[slot];
  } else {
    const double *sp = sspd + e * Nfn + (sl / Nfp) * Nfp;
    lam = sp[0];
    for (int j = 1; j < Nfp; ++j) lam = sp[j] > lam ? sp[j] : lam;
  }
  const long long i = k * Nfn + sl, nF = (long long)Nfn * v.K;
  const double dh = T.hM - T.hP, dhu = T.huM - T.huP, dhv = T.hvM - T.hvP;
  const double ghM = (0.5 * g * T.hM) * T.hM, ghP = (0.5 * g * T.hP) * T.hP;
  const double F2M = (T.huM * T.huM) * rM + ghM, G2M = (T.huM * T.hvM) * rM, G3M = (T.hvM * T.hvM) * rM + ghM;
  const double F2P = (T.huP * T.huP) * rP + ghP, G2P = (T.huP * T.hvP) * rP, G3P = (T.hvP * T.hvP) * rP + ghP;
  const double fs = v.Fscale[i];
  sflux[i] = fs * (0.5 * (dhu * T.nx + dhv * T.ny - lam * dh));
  sflux[nF + i] = fs * (0.5 * ((F2M - F2P) * T.nx + (G2M - G2P) * T.ny - lam * dhu));
  sflux[2 * nF + i] = fs * (0.5 * ((G2M - G2P) * T.nx + (G3M - G3P) * T.ny - lam * dhv));
}

// A9-A13: thread per (element, node); EB elements per CTA
__global__ void k_volume(Sw2dView v, double *__restrict__ q, double *__restrict__ res, const double *__restrict__ sflux,
                         double *__restrict__ out, int mode, double a, double b, int EB, double *__restrict__ qdst) {
  extern __shared__ double sm[];
  const int Np = v.Np, Nfn = v.Nfn;
  const int tid = threadIdx.x, e = tid / Np, n = tid - e * Np;
  const long long kb = (long long)blockIdx.x * EB;
  const long long k = kb + e;
  const bool active = (e < EB) && (k < v.K);
  const int nE = (int)min((long long)EB, (long long)v.K - kb);
  double *sF = sm;                       // [5][EB*Np]: hu, hv, F2, G2, G3
  double *sS = sm + 5 * EB * Np;         // [3][EB*Nfn]
  const long long nT = (long long)Np * v.K, nF = (long long)Nfn * v.K;
  const double g = v.g;
  double h = 0, hu = 0, hv = 0;
  const long long m = k * Np + n;
  if (active) {
    h = q[m]; hu = q[v.stride + m]; hv = q[2 * v.stride + m];
    const int o = e * Np + n, P = EB * Np;
    sF[o] = hu;
    sF[P + o] = hv;
    sF[2 * P + o] = (hu * hu) / h + 0.5 * g * h * h;
    sF[3 * P + o] = (hu * hv) / h;
    sF[4 * P + o] = (hv * hv) / h + 0.5 * g * h * h;
  }
  for (int x = tid; x < nE * Nfn; x += blockDim.x) {
    const long long i = kb * Nfn + x;
    sS[x] = sflux[i];
    sS[EB * Nfn + x] = sflux[nF + i];
    sS[2 * EB * Nfn + x] = sflux[2 * nF + i];
  }
  __syncthreads();
  double r1 = 0, r2 = 0, r3 = 0;
  if (active) {
    const int P = EB * Np;
    const double *F1 = sF + e * Np, *G1 = sF + P + e * Np, *F2 = sF + 2 * P + e * Np, *G2 = sF + 3 * P + e * Np,
                 *G3 = sF + 4 * P + e * Np;
    double drF1 = 0, dsF1 = 0, drG1 = 0, dsG1 = 0, drF2 = 0, dsF2 = 0, drG2 = 0, dsG2 = 0, drG3 = 0, dsG3 = 0;
    const double *Dr = v.Dr + (size_t)n * Np, *Ds = v.Ds + (size_t)n * Np;
    for (int j = 0; j < Np; ++j) {
      const double dr = __ldg(Dr + j), ds = __ldg(Ds + j);
      drF1 += dr * F1[j]; dsF1 += ds * F1[j];
      drG1 += dr * G1[j]; dsG1 += ds * G1[j];
      drF2 += dr * F2[j]; dsF2 += ds * F2[j];
      drG2 += dr * G2[j]; dsG2 += ds * G2[j];
      drG3 += dr * G3[j]; dsG3 += ds * G3[j];
    }
    const double rx = v.rx[m], sx = v.sx[m], ry = v.ry[m], sy = v.sy[m];
    r1 = -(rx * drF1 + sx * dsF1); r1 = r1 + (-(ry * drG1 + sy * dsG1));
    r2 = -(rx * drF2 + sx * dsF2); r2 = r2 + (-(ry * drG2 + sy * dsG2));
    r3 = -(rx * drG2 + sx * dsG2); r3 = r3 + (-(ry * drG3 + sy * dsG3));
    const double *L = v.Lift + (size_t)n * Nfn;
    const double *S1 = sS + e * Nfn, *S2 = sS + EB * Nfn + e * Nfn, *S3 = sS + 2 * EB * Nfn + e * Nfn;
    double l1 = 0, l2 = 0, l3 = 0;
    for (int j = 0; j < Nfn; ++j) {
      const double l = __ldg(L + j);
      l1 += l * S1[j]; l2 += l * S2[j]; l3 += l * S3[j];
    }
    r1 += l1; r2 += l2; r3 += l3;
    if (v.full) {  // A11 sources
      const double u = hu / h, w = hv / h, nu = sqrt(u * u + w * w);
      r2 = r2 + g * h * v.Hx[m]; r3 = r3 + g * h * v.Hy[m];
      r2 = r2 + (-v.CD * u * nu); r3 = r3 + (-v.CD * w * nu);
      r2 = r2 + v.fcor * hv;      r3 = r3 + (-v.fcor * hu);
    }
  }
  if (v.use_filter) {  // A12: RHS := Filt*RHS
    __syncthreads();
    const int P = EB * Np;
    if (active) { sF[e * Np + n] = r1; sF[P + e * Np + n] = r2; sF[2 * P + e * Np + n] = r3; }
    __syncthreads();
    if (active) {
      const double *Fl = v.Filt + (size_t)n * Np;
      double f1 = 0, f2 = 0, f3 = 0;
      for (int j = 0; j < Np; ++j) {
        const double c = __ldg(Fl + j);
        f1 += c * sF[e * Np + j]; f2 += c * sF[P + e * Np + j]; f3 += c * sF[2 * P + e * Np + j];
      }
      r1 = f1; r2 = f2; r3 = f3;
    }
  }
  if (!active) return;
  if (mode == 0) {
    out[m] = r1; out[nT + m] = r2; out[2 * nT + m] = r3;
  } else {  // A13 (src/advec1d/main.cpp:98,101)
    const double dt = v.scal[S_DT];
    double s1 = dt * r1, s2 = dt * r2, s3 = dt * r3;
    if (mode == 1 && a != 0.0) { s1 = a * res[m] + s1; s2 = a * res[v.stride + m] + s2; s3 = a * res[2 * v.stride + m] + s3; }
    if (mode == 1) {
      res[m] = s1; res[v.stride + m] = s2; res[2 * v.stride + m] = s3;
      q[m] = h + b * s1; q[v.stride + m] = hu + b * s2; q[2 * v.stride + m] = hv + b * s3;
    } else {  // SSP-RK2 stages + sponge of the src/sw2d driver (stepper_value, sw2d_common.cuh); qdst != q
      const double sg = v.sponge ? v.sponge[m] : 0.0;
      double o1 = 0, o2 = 0, o3 = 0;
      if (mode == 3) { o1 = qdst[m]; o2 = qdst[v.stride + m]; o3 = qdst[2 * v.stride + m]; }
      qdst[m] = stepper_value(mode, b, s1, h, o1, 0, sg);
      qdst[v.stride + m] = stepper_value(mode, b, s2, hu, o2, 1, sg);
      qdst[2 * v.stride + m] = stepper_value(mode, b, s3, hv, o3, 2, sg);
    }
  }
}

// A14: max over face nodes of Fscale*spd (src/sw2d/main.cpp:259-276) -> scal[S_DTMAX]
__global__ void k_dt_reduce(Sw2dView v, const double *__restrict__ q) {
  const long long nF = (long long)v.Nfn * v.K;
  double mx = 0.0;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nF; i += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(i / v.Nfn), sl = (int)(i % v.Nfn);
    const long long m = (long long)k * v.Np + v.fmask[sl];
    const double h = q[m], u = q[v.stride + m] / h, w = q[2 * v.stride + m] / h;
    const double val = v.Fscale[i] * (sqrt(u * u + w * w) + sqrt(v.g * h));
    mx = (__double_as_longlong(val) > __double_as_longlong(mx)) ? val : mx;
  }
  mx = block_max(mx);
  if (threadIdx.x == 0) atomic_max_pos(&v.scal[S_DTMAX], mx);
}
__global__ void k_dt_finalize(Sw2dView v, double CFL) {
  v.scal[S_DT] = CFL / ((v.N + 1) * (v.N + 1) * 0.5 * v.scal[S_DTMAX]);
  v.scal[S_DTMAX] = 0.0;
}
__global__ void k_advance_time(Sw2dView v) { v.scal[S_T] += v.scal[S_DT]; }

// a16 blow-up guard: max |h - H| and NaN flag
__global__ void k_check(Sw2dView v, const double *__restrict__ q) {
  const long long nT = (long long)v.Np * v.K;
  double mx = 0.0;
  int nan = 0;
  for (long long m = blockIdx.x * (long long)blockDim.x + threadIdx.x; m < nT; m += (long long)gridDim.x * blockDim.x) {
    const double eta = q[m] - (v.H ? v.H[m] : 0.0);
    if (eta != eta) nan = 1;
    const double ae = fabs(eta);
    if (ae == ae && ae > mx) mx = ae;
  }
  mx = block_max(mx);
  if (threadIdx.x == 0) atomic_max_pos(&v.scal[S_ETAMAX], mx);
  if (nan) v.scal[S_NANFLAG] = 1.0;
}

// halo pack: sendbuf[seg][c][j] = q[c][send_nodes[j]]; segments are contiguous so one flat gather works:
// layout per neighbour i: base = nplanes*send_off[i], plane c at base + c*cnt_i
__global__ void k_pack(const double *__restrict__ q, long long stride, const double *__restrict__ H, const int32_t *__restrict__ nodes,
                       const int32_t *__restrict__ seg_of, const int32_t *__restrict__ send_off, int n_send, int nplanes,
                       double *__restrict__ buf) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_send) return;
  const int sg = seg_of[j];
  const int o0 = send_off[sg], cnt = send_off[sg + 1] - o0;
  const long long base = (long long)nplanes * o0 + (j - o0);
  const int m = nodes[j];
  buf[base] = q[m];
  buf[base + cnt] = q[stride + m];
  buf[base + 2LL * cnt] = q[2 * stride + m];
  if (nplanes == 4) buf[base + 3LL * cnt] = H[m];
}

inline int grid_for(long long n, int block, int cap) {
  long long g = (n + block - 1) / block;
  if (g < 1) g = 1;
  if (g > cap) g = cap;
  return (int)g;
}

}  // namespace

cudaError_t launch_generic_speed(bdg_sw2d *s, const double *q, int slot) {
  k_speed<<<grid_for((long long)s->nF, 256, 148 * 8), 256, 0, s->stream>>>(s->v, q, slot);
  s->launches++;
  return cudaGetLastError();
}

cudaError_t launch_generic_face(bdg_sw2d *s, const double *q, int k0, int k1, int slot) {
  if (k1 <= k0) return cudaSuccess;
  static const bool per_face = getenv("BDG_FACE_SCALAR") != nullptr;  // study switch: the thread-per-face kernel
  if (per_face) {
    const long long nfaces = (long long)(k1 - k0) * s->v.Nf;
    k_face<<<(unsigned)((nfaces + 127) / 128), 128, 0, s->stream>>>(s->v, q, s->sflux, k0, k1, slot);
  } else {
    const int Nfn = s->v.Nfn;
    const int EB = Nfn >= 256 ? 1 : 256 / Nfn, threads = (EB * Nfn + 31) / 32 * 32;
    k_face_nodes<<<(unsigned)((k1 - k0 + EB - 1) / EB), threads, sizeof(double) * threads, s->stream>>>(s->v, q, s->sflux, k0, k1, slot, EB);
  }
  s->launches++;
  return cudaGetLastError();
}

cudaError_t launch_generic_volume(bdg_sw2d *s, double *q, double *res, double *out, int mode, double a, double b, double *qdst) {
  static const bool scalar_only = getenv("BDG_VOLUME_SCALAR") != nullptr;  // study switch: the thread-per-node kernel
  if (!scalar_only && volume_dmma_supported(s)) return launch_volume_dmma(s, q, res, out, mode, a, b, qdst);
  const int Np = s->v.Np, Nfn = s->v.Nfn;
  int EB = 256 / Np;
  if (EB < 1) EB = 1;
  const int threads = ((EB * Np + 31) / 32) * 32;
  const size_t smem = sizeof(double) * ((size_t)5 * EB * Np + (size_t)3 * EB * Nfn);
  const unsigned grid = (unsigned)((s->v.K + EB - 1) / EB);
  k_volume<<<grid, threads, smem, s->stream>>>(s->v, q, res, s->sflux, out, mode, a, b, EB, mode >= 2 ? qdst : nullptr);
  s->launches++;
  return cudaGetLastError();
}

cudaError_t launch_dt_reduce(bdg_sw2d *s, const double *q) {
  k_dt_reduce<<<grid_for((long long)s->nF, 256, 148 * 8), 256, 0, s->stream>>>(s->v, q);
  s->launches++;
  return cudaGetLastError();
}
cudaError_t launch_dt_finalize(bdg_sw2d *s) {
  k_dt_finalize<<<1, 1, 0, s->stream>>>(s->v, s->prm.CFL);
  s->launches++;
  return cudaGetLastError();
}
cudaError_t launch_advance_time(bdg_sw2d *s) {
  k_advance_time<<<1, 1, 0, s->stream>>>(s->v);
  s->launches++;
  return cudaGetLastError();
}
cudaError_t launch_check(bdg_sw2d *s, const double *q) {
  k_check<<<grid_for((long long)s->nT, 256, 148 * 8), 256, 0, s->stream>>>(s->v, q);
  s->launches++;
  return cudaGetLastError();
}

// seg_of is stored right behind send_nodes on the device: [n_send nodes][n_send seg ids][n_nbr+1 offsets]
cudaError_t launch_pack(bdg_sw2d *s, const double *q, int with_H) {
  if (s->n_send == 0) return cudaSuccess;
  const int32_t *nodes = s->d_send_nodes, *seg = s->d_send_nodes + s->n_send, *off = s->d_send_nodes + 2 * (size_t)s->n_send;
  k_pack<<<(s->n_send + 255) / 256, 256, 0, s->stream>>>(q, s->v.stride, s->d_H, nodes, seg, off, s->n_send, with_H ? 4 : 3,
                                                         s->d_sendbuf);
  s->launches++;
  return cudaGetLastError();
}
