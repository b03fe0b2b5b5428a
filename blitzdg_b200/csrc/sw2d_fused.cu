// Fused sw2d stage kernel for straight-sided (affine) triangles: ONE launch per LSERK4 stage does
//   trace gather (vmapP) + BC + Lax-Friedrichs flux            (reference src/sw2d-simple/main.cpp:246-319, src/sw2d/main.cpp:327-421)
//   volume term + LIFT*(Fscale*flux)                            (:332-355 / :434-458)
//   res = a*res + dt*RHS ; q += b*res                           (src/advec1d/main.cpp:98,101)
// reading q/res once and writing them once (B_alg of SURVEY.md §8d).
//
// Mapping.  A CTA of 4 warps owns a tile of 32 consecutive elements; the tile's h,hu,hv,res,vmapP and
// per-element geometry arrive in shared memory through 1-D TMA bulk copies (cp.async.bulk + mbarrier) and
// the results leave the same way.  Because the element is affine,
//   RHS_i = -( Dr (rx F_i + ry G_i) + Ds (sx F_i + sy G_i) ) + LIFT (Fscale dF_i)
// is one small GEMM  C[Np x cols] = A[Np x Kp] * B[Kp x cols]  with the constant operator
//   A = [ -Dr | -Ds | LIFT ]   (each block padded to a multiple of 4 columns, rows to a multiple of 8)
// and B assembled per (element, field) column from the pointwise fluxes.  The contraction runs on the fp64
// tensor pipe (mma.sync.m8n8k4.f64 = DMMA), the operator fragments live in registers (N<=5) or in shared
// memory in fragment order, and every thread computes its B fragment entries DIRECTLY in fragment layout:
//   lane = 4*c + t  ->  element e = warp + 4*c of the tile (c = 0..7), nodes / face slots j = 4*kb + t.
// The 4 lanes of a quad therefore hold all face nodes of one element and the per-face max of the wave speed
// (sw2d-simple main.cpp:305-312) is two xor-shuffles.
//
// Deviations from the reference's operation order (inside the parity tolerance, SURVEY.md §7 hard part 1):
// FMA contraction, x/h evaluated as x*(1/h) with a Newton-refined reciprocal, chain-rule factors applied
// before the contraction (exact for affine elements up to round-off).
#include <cstdio>
#include <vector>

#include "sw2d_common.cuh"

namespace {

constexpr int kWarps = 4;
constexpr int kThreads = kWarps * 32;
constexpr int kTile = kWarps * 8;  // elements per CTA tile
constexpr int kAW = 13;            // doubles per element in v.aff: rx,sx,ry,sy,(nx,ny,Fscale)x3

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  } while (!ok);
}
// 1-D TMA bulk copy global -> shared, completion counted in bytes on an mbarrier
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void bulk_s2g(void *dst, const void *src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)), "r"(bytes) : "memory");
}

__device__ __forceinline__ void dmma884(double &d0, double &d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

// 1/x: MUFU seed (20+ bits) + two Newton steps -> <= 1 ulp-ish; no denormal / inf handling (h > 0 here)
__device__ __forceinline__ double frcp(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  double e = fma(-x, r, 1.0);
  r = fma(r, e, r);
  e = fma(-x, r, 1.0);
  return fma(r, e, r);
}
// sqrt(x), x >= 0 (0 maps to 1e-150): coupled Newton on (sqrt, 1/(2 sqrt)) from the MUFU rsqrt seed
__device__ __forceinline__ double fsqrt(double x) {
  x = fmax(x, 1e-300);
  double r;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  double hh = 0.5 * r, s = x * r;
  double e = fma(-hh, s, 0.5);
  s = fma(s, e, s);
  hh = fma(hh, e, hh);
  e = fma(-hh, s, 0.5);
  s = fma(s, e, s);
  hh = fma(hh, e, hh);
  e = fma(-s, s, x);
  return fma(e, hh, s);
}

template <int NP, int NFP>
struct Cfg {
  static constexpr int NFN = 3 * NFP;
  static constexpr int NPB = (NP + 3) / 4;    // k-blocks of the Dr part (and of the Ds part)
  static constexpr int NFB = (NFN + 3) / 4;   // k-blocks of the LIFT part
  static constexpr int KB = 2 * NPB + NFB;
  static constexpr int MB = (NP + 7) / 8;     // row blocks
  static constexpr bool A_REGS = (MB * KB <= 36);
  static constexpr int TNP = kTile * NP;      // doubles per field per tile
  // one tile buffer (bytes, every section 16-B aligned): q, res, geometry, vmapP, face tags
  static constexpr size_t bQ = 0;
  static constexpr size_t bR = bQ + sizeof(double) * 3 * TNP;
  static constexpr size_t bAff = bR + sizeof(double) * 3 * TNP;
  static constexpr size_t bVm = bAff + sizeof(double) * kTile * kAW;
  static constexpr size_t bTag = bVm + sizeof(int32_t) * kTile * NFN;
  static constexpr size_t BUF = bTag + ((kTile * 3 + 15) / 16) * 16;
  static constexpr uint32_t txQ = sizeof(double) * TNP;  // one field plane of a tile
  static constexpr uint32_t txMeta = sizeof(int32_t) * kTile * NFN + sizeof(double) * kTile * kAW + kTile * 3;
  // CTA-wide sections behind NBUF tile buffers
  static constexpr size_t sizeFl = sizeof(double) * 4 * TNP;
  static constexpr size_t sizeA = A_REGS ? 0 : sizeof(double) * MB * KB * 32;
  static constexpr size_t sizeFm = ((sizeof(int32_t) * NFN + 15) / 16) * 16;
  static constexpr size_t total(int nbuf) { return nbuf * BUF + sizeFl + sizeA + sizeFm + 64; }
};

struct StageArgs {
  const double *q_in;
  double *q_out;
  double *res;
  double *rhs_out;      // write_rhs mode: [3][nT]
  const double *Afrag;  // [MB][KB][32] operator in DMMA A-fragment order
  int k0, k1;           // active element range
  int tile0, ntiles;    // tiles [tile0, tile0 + ntiles) of this launch
  int write_rhs;
  int slot;             // scal slot of the global LF speed
  double a, b;          // LSERK4 stage coefficients
};

template <int NP, int NFP>
struct Smem {
  using C = Cfg<NP, NFP>;
  unsigned char *base;
  int nbuf;
  __device__ double *Q(int b) const { return reinterpret_cast<double *>(base + b * C::BUF + C::bQ); }
  __device__ double *R(int b) const { return reinterpret_cast<double *>(base + b * C::BUF + C::bR); }
  __device__ double *Aff(int b) const { return reinterpret_cast<double *>(base + b * C::BUF + C::bAff); }
  __device__ int32_t *Vm(int b) const { return reinterpret_cast<int32_t *>(base + b * C::BUF + C::bVm); }
  __device__ uint8_t *Tag(int b) const { return base + b * C::BUF + C::bTag; }
  __device__ double *Fl() const { return reinterpret_cast<double *>(base + nbuf * C::BUF); }
  __device__ double *A() const { return reinterpret_cast<double *>(base + nbuf * C::BUF + C::sizeFl); }
  __device__ int32_t *Fm() const { return reinterpret_cast<int32_t *>(base + nbuf * C::BUF + C::sizeFl + C::sizeA); }
  __device__ uint64_t *Bar() const { return reinterpret_cast<uint64_t *>(base + nbuf * C::BUF + C::sizeFl + C::sizeA + C::sizeFm); }
};

// thread 0: TMA bulk loads of one full tile into buffer b, completion on bar
template <int NP, int NFP>
__device__ __forceinline__ void issue_tile_load(const Sw2dView &v, const StageArgs &p, const Smem<NP, NFP> &S, int b, long long kt0,
                                                bool need_res, uint64_t *bar) {
  using C = Cfg<NP, NFP>;
  const long long base = kt0 * NP;
  mbar_expect_tx(bar, (need_res ? 6 : 3) * C::txQ + C::txMeta);
  bulk_g2s(S.Vm(b), v.vmapP + kt0 * C::NFN, sizeof(int32_t) * kTile * C::NFN, bar);
  bulk_g2s(S.Tag(b), v.facetag + kt0 * 3, kTile * 3, bar);
  bulk_g2s(S.Aff(b), v.aff + kt0 * kAW, sizeof(double) * kTile * kAW, bar);
  for (int f = 0; f < 3; ++f) bulk_g2s(S.Q(b) + f * C::TNP, p.q_in + f * v.stride + base, C::txQ, bar);
  if (need_res)
    for (int f = 0; f < 3; ++f) bulk_g2s(S.R(b) + f * C::TNP, p.res + f * v.stride + base, C::txQ, bar);
}

// The arithmetic of one tile that sits in shared memory (A1-A13 of SURVEY.md Appendix A).  On return the
// tile's q and res planes hold the updated values (stage mode) -- or, in write_rhs mode, rhs_out has been
// written and the tile is untouched.  Contains two __syncthreads(); the caller makes the tile visible first.
template <int NP, int NFP>
__device__ __forceinline__ void compute_tile(const Sw2dView &v, const StageArgs &p, double *__restrict__ sQ, double *__restrict__ sR,
                                             const double *__restrict__ sAff, const int32_t *__restrict__ sVm,
                                             const uint8_t *__restrict__ sTag, double *__restrict__ sFl, const double *__restrict__ sA,
                                             const int32_t *__restrict__ sFm, const double *Areg, long long kt0, int nload,
                                             bool need_res) {
  using C = Cfg<NP, NFP>;
  constexpr int NFN = C::NFN, NPB = C::NPB, NFB = C::NFB, KB = C::KB, MB = C::MB, TNP = C::TNP;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int c = lane >> 2, t = lane & 3;
  const int e = w + 4 * c;  // my element of the tile (B-fragment column c)
  const int eo = e * NP;
  const long long stride = v.stride;
  const long long base = kt0 * NP;  // first volume node of the tile
  const int tile_nodes = nload * NP;
  const double g = v.g, half_g = 0.5 * v.g;
  const double *af = sAff + e * kAW;

  // ---- '+' traces that live outside the tile: issue the L2 gathers now, consume them in phase 2 ---------
  double gP[NFB][3];
  int rel[NFB];  // >= 0: tile-local node of the '+' trace; -1: gathered into gP; -2: wall; -3: no slot
#pragma unroll
  for (int kb = 0; kb < NFB; ++kb) {
    const int sl = 4 * kb + t;
    rel[kb] = -3;
    gP[kb][0] = gP[kb][1] = gP[kb][2] = 0.0;
    if (sl < NFN) {
      const int tag = sTag[e * 3 + sl / NFP];
      const long long mP = sVm[e * NFN + sl];
      const long long r = mP - base;
      if (tag == 3) {
        rel[kb] = -2;
      } else if (r >= 0 && r < tile_nodes) {
        rel[kb] = (int)r;
      } else {
        rel[kb] = -1;
        gP[kb][0] = __ldg(p.q_in + mP);
        gP[kb][1] = __ldg(p.q_in + stride + mP);
        gP[kb][2] = __ldg(p.q_in + 2 * stride + mP);
      }
    }
  }

  // ---- phase 1: nodal fluxes and wave speed at my volume nodes (A6, A7 '-' side) ----------------------
#pragma unroll
  for (int kb = 0; kb < NPB; ++kb) {
    const int n = 4 * kb + t;
    if (n < NP) {
      const double h = sQ[eo + n], hu = sQ[TNP + eo + n], hv = sQ[2 * TNP + eo + n];
      const double r = frcp(h), u = hu * r, vv = hv * r, gh2 = (half_g * h) * h;
      sFl[eo + n] = fma(hu, u, gh2);            // F2
      sFl[TNP + eo + n] = hu * vv;              // G2 = F3
      sFl[2 * TNP + eo + n] = fma(hv, vv, gh2); // G3
      sFl[3 * TNP + eo + n] = fsqrt(fma(u, u, vv * vv)) + fsqrt(g * h);
    }
  }
  __syncthreads();

  // ---- phase 2: face slots of my element: '+' trace, BC, speed, flux jump (A1-A8) ---------------------
  double fn[NFB][3], dq[NFB][3];  // normal flux differences and jumps per slot
  double lam0 = 0.0, lam1 = 0.0, lam2 = 0.0;  // per-face max speed (scalars: a runtime face index must not spill)
#pragma unroll
  for (int kb = 0; kb < NFB; ++kb) {
    const int sl = 4 * kb + t;
    fn[kb][0] = fn[kb][1] = fn[kb][2] = 0.0;
    dq[kb][0] = dq[kb][1] = dq[kb][2] = 0.0;
    if (sl < NFN) {
      const int f = sl / NFP;
      const int n = sFm[sl];
      const double nx = af[4 + 3 * f], ny = af[5 + 3 * f];
      const double hM = sQ[eo + n], huM = sQ[TNP + eo + n], hvM = sQ[2 * TNP + eo + n];
      const double F2M = sFl[eo + n], G2M = sFl[TNP + eo + n], G3M = sFl[2 * TNP + eo + n], spdM = sFl[3 * TNP + eo + n];
      double hP, huP, hvP, F2P, G2P, G3P, spdP;
      const int r = rel[kb];
      if (r >= 0) {  // neighbour inside the tile: everything is already in shared memory
        hP = sQ[r]; huP = sQ[TNP + r]; hvP = sQ[2 * TNP + r];
        F2P = sFl[r]; G2P = sFl[TNP + r]; G3P = sFl[2 * TNP + r]; spdP = sFl[3 * TNP + r];
      } else {
        if (r == -2) {  // Wall (main.cpp:340-345)
          const double d = huM * nx + hvM * ny;
          hP = hM;
          huP = fma(-2.0 * nx, d, huM);
          hvP = fma(-2.0 * ny, d, hvM);
        } else {  // neighbour tile or ghost slot
          hP = gP[kb][0]; huP = gP[kb][1]; hvP = gP[kb][2];
        }
        const double rr = frcp(hP), u = huP * rr, vv = hvP * rr, gh2 = (half_g * hP) * hP;
        F2P = fma(huP, u, gh2); G2P = huP * vv; G3P = fma(hvP, vv, gh2);
        spdP = fsqrt(fma(u, u, vv * vv)) + fsqrt(g * hP);
      }
      const double sp = fmax(spdM, spdP);
      lam0 = f == 0 ? fmax(lam0, sp) : lam0;
      lam1 = f == 1 ? fmax(lam1, sp) : lam1;
      lam2 = f == 2 ? fmax(lam2, sp) : lam2;
      dq[kb][0] = hM - hP; dq[kb][1] = huM - huP; dq[kb][2] = hvM - hvP;
      fn[kb][0] = fma(dq[kb][1], nx, dq[kb][2] * ny);
      fn[kb][1] = fma(F2M - F2P, nx, (G2M - G2P) * ny);
      fn[kb][2] = fma(G2M - G2P, nx, (G3M - G3P) * ny);
    }
  }
  if (v.lf_global) {
    lam0 = lam1 = lam2 = v.scal[p.slot];
  } else {  // per-face max over the quad (the 4 lanes of a quad hold all face nodes of element e)
    lam0 = fmax(lam0, __shfl_xor_sync(0xffffffffu, lam0, 1));
    lam1 = fmax(lam1, __shfl_xor_sync(0xffffffffu, lam1, 1));
    lam2 = fmax(lam2, __shfl_xor_sync(0xffffffffu, lam2, 1));
    lam0 = fmax(lam0, __shfl_xor_sync(0xffffffffu, lam0, 2));
    lam1 = fmax(lam1, __shfl_xor_sync(0xffffffffu, lam1, 2));
    lam2 = fmax(lam2, __shfl_xor_sync(0xffffffffu, lam2, 2));
  }
  double sfl[NFB][3];  // Fscale * dFlux_i at my slots = LIFT-part B fragments
#pragma unroll
  for (int kb = 0; kb < NFB; ++kb) {
    const int sl = 4 * kb + t;
    const int f = sl < NFN ? sl / NFP : 0;
    const double l = f == 0 ? lam0 : (f == 1 ? lam1 : lam2);
    const double hfs = 0.5 * af[6 + 3 * f];
#pragma unroll
    for (int i = 0; i < 3; ++i) sfl[kb][i] = hfs * fma(-l, dq[kb][i], fn[kb][i]);
  }
  __syncthreads();  // all '+' lookups into sQ/sFl are done: the tile may now be updated in place

  // ---- phase 3: per field, B fragments -> DMMA -> stage update ------------------------------------------
  const double rx = af[0], sx = af[1], ry = af[2], sy = af[3];
  const double dt = p.write_rhs ? 0.0 : v.scal[S_DT];
  const int gr = lane >> 2;                       // C fragment row within a row block
  const int e0 = w + 4 * (2 * t), e1 = e0 + 4;    // C fragment columns 2t, 2t+1
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const double *X = i == 0 ? sQ + TNP : (i == 1 ? sFl : sFl + TNP);                // F_i: hu, F2, G2
    const double *Y = i == 0 ? sQ + 2 * TNP : (i == 1 ? sFl + TNP : sFl + 2 * TNP);  // G_i: hv, G2, G3
    double acc[MB][2];
#pragma unroll
    for (int rb = 0; rb < MB; ++rb) acc[rb][0] = acc[rb][1] = 0.0;
#pragma unroll
    for (int kb = 0; kb < NPB; ++kb) {
      const int n = 4 * kb + t;
      double ba = 0.0, bb = 0.0;
      if (n < NP) {
        const double xx = X[eo + n], yy = Y[eo + n];
        ba = fma(rx, xx, ry * yy);
        bb = fma(sx, xx, sy * yy);
      }
#pragma unroll
      for (int rb = 0; rb < MB; ++rb) {
        const double a1 = C::A_REGS ? Areg[rb * KB + kb] : sA[(rb * KB + kb) * 32 + lane];
        const double a2 = C::A_REGS ? Areg[rb * KB + NPB + kb] : sA[(rb * KB + NPB + kb) * 32 + lane];
        dmma884(acc[rb][0], acc[rb][1], a1, ba);
        dmma884(acc[rb][0], acc[rb][1], a2, bb);
      }
    }
#pragma unroll
    for (int kb = 0; kb < NFB; ++kb) {
#pragma unroll
      for (int rb = 0; rb < MB; ++rb) {
        const double a3 = C::A_REGS ? Areg[rb * KB + 2 * NPB + kb] : sA[(rb * KB + 2 * NPB + kb) * 32 + lane];
        dmma884(acc[rb][0], acc[rb][1], a3, sfl[kb][i]);
      }
    }
    // epilogue: rows n = 8*rb + gr of elements e0, e1
#pragma unroll
    for (int rb = 0; rb < MB; ++rb) {
      const int n = 8 * rb + gr;
      if (n < NP) {
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
          const int o = (cc ? e1 : e0) * NP + n;
          const double rhs = acc[rb][cc];
          if (p.write_rhs) {
            const long long k = kt0 + (cc ? e1 : e0);
            if (k >= p.k0 && k < p.k1) p.rhs_out[(long long)i * NP * v.K + k * NP + n] = rhs;
          } else {
            double rs = dt * rhs;
            if (need_res) rs = fma(p.a, sR[i * TNP + o], rs);
            sR[i * TNP + o] = rs;
            sQ[i * TNP + o] = fma(p.b, rs, sQ[i * TNP + o]);
          }
        }
      }
    }
  }
}

template <int NP, int NFP>
__device__ __forceinline__ void load_operator(const StageArgs &p, double *Areg, double *sA, int32_t *sFm, const Sw2dView &v) {
  using C = Cfg<NP, NFP>;
  const int tid = threadIdx.x, lane = tid & 31;
  if (C::A_REGS) {
#pragma unroll
    for (int i = 0; i < C::MB * C::KB; ++i) Areg[i] = __ldg(p.Afrag + i * 32 + lane);
  } else {
    for (int x = tid; x < C::MB * C::KB * 32; x += kThreads) sA[x] = __ldg(p.Afrag + x);
  }
  if (tid < C::NFN) sFm[tid] = v.fmask[tid];
}

// ---- edge kernel: one tile per CTA, any tile (ragged mesh end, partially active range, write_rhs) -------
template <int NP, int NFP>
__global__ void __launch_bounds__(kThreads, 3) k_fused_edge(Sw2dView v, StageArgs p) {
  using C = Cfg<NP, NFP>;
  constexpr int NFN = C::NFN, TNP = C::TNP;
  extern __shared__ __align__(128) unsigned char smem[];
  const Smem<NP, NFP> S{smem, 1};
  double *sQ = S.Q(0), *sR = S.R(0), *sAff = S.Aff(0);
  int32_t *sVm = S.Vm(0);
  uint8_t *sTag = S.Tag(0);
  uint64_t *bar = S.Bar();

  const int tid = threadIdx.x;
  const long long kt0 = (long long)(p.tile0 + blockIdx.x) * kTile;
  const int nload = (int)min((long long)kTile, (long long)v.K - kt0);
  const bool full_tile = nload == kTile;
  const bool all_active = full_tile && kt0 >= p.k0 && kt0 + kTile <= p.k1;
  const bool need_res = !p.write_rhs && p.a != 0.0;
  const long long stride = v.stride;
  const long long base = kt0 * NP;

  if (full_tile) {
    if (tid == 0) {
      mbar_init(&bar[0], 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      issue_tile_load<NP, NFP>(v, p, S, 0, kt0, need_res, &bar[0]);
    }
  } else {  // ragged last tile: plain loads, neutral fill behind the mesh end
    for (int x = tid; x < 3 * TNP; x += kThreads) {
      const int f = x / TNP, r = x - f * TNP;
      const bool in = r < nload * NP;
      sQ[x] = in ? p.q_in[f * stride + base + r] : (f == 0 ? 1.0 : 0.0);
      sR[x] = (in && need_res) ? p.res[f * stride + base + r] : 0.0;
    }
    for (int x = tid; x < kTile * NFN; x += kThreads) {
      const int el = x / NFN, sl = x - el * NFN;
      sVm[x] = el < nload ? v.vmapP[kt0 * NFN + x] : (int32_t)(base + el * NP + v.fmask[sl]);
    }
    for (int x = tid; x < kTile * kAW; x += kThreads) sAff[x] = x < nload * kAW ? v.aff[kt0 * kAW + x] : 0.0;
    for (int x = tid; x < kTile * 3; x += kThreads) sTag[x] = x < nload * 3 ? v.facetag[kt0 * 3 + x] : 0;
  }
  double Areg[C::A_REGS ? C::MB * C::KB : 1];
  load_operator<NP, NFP>(p, Areg, S.A(), S.Fm(), v);
  __syncthreads();  // mbarrier init + plain stores visible
  if (full_tile) mbar_wait(&bar[0], 0);

  compute_tile<NP, NFP>(v, p, sQ, sR, sAff, sVm, sTag, S.Fl(), S.A(), S.Fm(), Areg, kt0, nload, need_res);
  if (p.write_rhs) return;

  if (all_active) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (tid == 0) {
      for (int f = 0; f < 3; ++f) {
        bulk_s2g(p.q_out + f * stride + base, sQ + f * TNP, C::txQ);
        bulk_s2g(p.res + f * stride + base, sR + f * TNP, C::txQ);
      }
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    }
  } else {
    __syncthreads();
    for (int x = tid; x < 3 * TNP; x += kThreads) {
      const int f = x / TNP, r = x - f * TNP;
      const long long k = kt0 + r / NP;
      if (k >= p.k0 && k < p.k1) {
        p.q_out[f * stride + base + r] = sQ[x];
        p.res[f * stride + base + r] = sR[x];
      }
    }
  }
}

// ---- persistent kernel: full, fully active tiles; two tile buffers, the next tile's TMA loads fly while
// the current tile is computed; results leave through TMA bulk stores from the same buffer -----------------
template <int NP, int NFP>
__global__ void __launch_bounds__(kThreads, 3) k_fused_persistent(Sw2dView v, StageArgs p) {
  using C = Cfg<NP, NFP>;
  constexpr int TNP = C::TNP;
  extern __shared__ __align__(128) unsigned char smem[];
  const Smem<NP, NFP> S{smem, 2};
  uint64_t *bar = S.Bar();
  const int tid = threadIdx.x;
  const bool need_res = p.a != 0.0;
  const long long stride = v.stride;
  const int nmine = (p.ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;  // tiles blockIdx.x + it*gridDim.x

  if (tid == 0) {
    mbar_init(&bar[0], 1);
    mbar_init(&bar[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    for (int it = 0; it < 2 && it < nmine; ++it)
      issue_tile_load<NP, NFP>(v, p, S, it, (long long)(p.tile0 + blockIdx.x + it * gridDim.x) * kTile, need_res, &bar[it]);
  }
  double Areg[C::A_REGS ? C::MB * C::KB : 1];
  load_operator<NP, NFP>(p, Areg, S.A(), S.Fm(), v);
  __syncthreads();

  for (int it = 0; it < nmine; ++it) {
    const int b = it & 1;
    const long long kt0 = (long long)(p.tile0 + blockIdx.x + it * gridDim.x) * kTile;
    mbar_wait(&bar[b], (it >> 1) & 1);
    compute_tile<NP, NFP>(v, p, S.Q(b), S.R(b), S.Aff(b), S.Vm(b), S.Tag(b), S.Fl(), S.A(), S.Fm(), Areg, kt0, kTile, need_res);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (tid == 0) {
      const long long base = kt0 * NP;
      for (int f = 0; f < 3; ++f) {
        bulk_s2g(p.q_out + f * stride + base, S.Q(b) + f * TNP, C::txQ);
        bulk_s2g(p.res + f * stride + base, S.R(b) + f * TNP, C::txQ);
      }
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // buffer b drained -> refill it
      if (it + 2 < nmine)
        issue_tile_load<NP, NFP>(v, p, S, b, (long long)(p.tile0 + blockIdx.x + (it + 2) * gridDim.x) * kTile, need_res, &bar[b]);
    }
  }
}

struct FusedOps {
  double *Afrag;
  int Np, Nfp;
};

int g_sm_count[16] = {};

template <int NP, int NFP>
cudaError_t launch_t(bdg_sw2d *s, StageArgs a) {
  using C = Cfg<NP, NFP>;
  static bool attr_set[16] = {};
  static int ctas_per_sm[16] = {};
  const int dev = s->device & 15;
  const size_t smem_edge = C::total(1), smem_pers = C::total(2);
  cudaError_t e;
  if (!attr_set[dev]) {
    if ((e = cudaFuncSetAttribute(k_fused_edge<NP, NFP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_edge)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_fused_persistent<NP, NFP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_pers)) != cudaSuccess) return e;
    cudaFuncSetAttribute(k_fused_edge<NP, NFP>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaFuncSetAttribute(k_fused_persistent<NP, NFP>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm[dev], k_fused_persistent<NP, NFP>, kThreads, smem_pers)) != cudaSuccess) return e;
    if ((e = cudaDeviceGetAttribute(&g_sm_count[dev], cudaDevAttrMultiProcessorCount, s->device)) != cudaSuccess) return e;
    if (ctas_per_sm[dev] < 1) return cudaErrorLaunchOutOfResources;
    attr_set[dev] = true;
  }
  // tiles [t0,t1) cover [k0,k1); the fully active ones [p0,p1) go to the persistent kernel (k1 <= K, so
  // they are also full), ragged / partially active head and tail tiles to the edge kernel
  const int t0 = a.k0 / kTile, t1 = (a.k1 - 1) / kTile + 1;
  int p0 = (a.k0 + kTile - 1) / kTile, p1 = a.k1 / kTile;
  if (a.write_rhs || p1 <= p0) p0 = p1 = t1;
  auto edge = [&](int first, int last) {
    if (last <= first) return;
    StageArgs h = a;
    h.tile0 = first; h.ntiles = last - first;
    k_fused_edge<NP, NFP><<<h.ntiles, kThreads, smem_edge, s->stream>>>(s->v, h);
    s->launches++;
  };
  edge(t0, p0);
  if (p1 > p0) {
    StageArgs m = a;
    m.tile0 = p0; m.ntiles = p1 - p0;
    int grid = g_sm_count[dev] * ctas_per_sm[dev];
    if (grid > m.ntiles) grid = m.ntiles;
    k_fused_persistent<NP, NFP><<<grid, kThreads, smem_pers, s->stream>>>(s->v, m);
    s->launches++;
  }
  edge(p1, t1);
  return cudaGetLastError();
}

template <int NP, int NFP>
void build_afrag(const double *Dr, const double *Ds, const double *Lift, std::vector<double> &out) {
  using C = Cfg<NP, NFP>;
  constexpr int NPB = C::NPB, KB = C::KB, MB = C::MB, NFN = C::NFN;
  out.assign((size_t)MB * KB * 32, 0.0);
  for (int rb = 0; rb < MB; ++rb)
    for (int kb = 0; kb < KB; ++kb)
      for (int lane = 0; lane < 32; ++lane) {
        const int row = 8 * rb + lane / 4;  // A fragment of m8n8k4: row = lane/4, col = lane%4
        double val = 0.0;
        if (row < NP) {
          if (kb < NPB) {
            const int j = 4 * kb + lane % 4;
            if (j < NP) val = -Dr[row * NP + j];
          } else if (kb < 2 * NPB) {
            const int j = 4 * (kb - NPB) + lane % 4;
            if (j < NP) val = -Ds[row * NP + j];
          } else {
            const int j = 4 * (kb - 2 * NPB) + lane % 4;
            if (j < NFN) val = Lift[row * NFN + j];
          }
        }
        out[((size_t)rb * KB + kb) * 32 + lane] = val;
      }
}

#define BDG_FUSED_ORDERS(X) X(3, 2) X(6, 3) X(10, 4) X(15, 5) X(21, 6)

}  // namespace

int fused_supported(const bdg_sw2d *s) {
  if (!s->is_affine || s->v.Nf != 3 || s->v.full || s->v.use_filter) return 0;
#define X(np, nfp) \
  if (s->v.Np == np && s->v.Nfp == nfp) return 1;
  BDG_FUSED_ORDERS(X)
#undef X
  return 0;
}

cudaError_t fused_prepare(bdg_sw2d *s) {
  const int Np = s->v.Np, Nfn = s->v.Nfn;
  std::vector<double> Dr((size_t)Np * Np), Ds((size_t)Np * Np), Lift((size_t)Np * Nfn), frag;
  cudaError_t e;
  if ((e = cudaStreamSynchronize(s->stream)) != cudaSuccess) return e;
  if ((e = cudaMemcpy(Dr.data(), s->d_Dr, sizeof(double) * Dr.size(), cudaMemcpyDeviceToHost)) != cudaSuccess) return e;
  if ((e = cudaMemcpy(Ds.data(), s->d_Ds, sizeof(double) * Ds.size(), cudaMemcpyDeviceToHost)) != cudaSuccess) return e;
  if ((e = cudaMemcpy(Lift.data(), s->d_Lift, sizeof(double) * Lift.size(), cudaMemcpyDeviceToHost)) != cudaSuccess) return e;
#define X(np, nfp) \
  if (Np == np && s->v.Nfp == nfp) build_afrag<np, nfp>(Dr.data(), Ds.data(), Lift.data(), frag);
  BDG_FUSED_ORDERS(X)
#undef X
  if (frag.empty()) return cudaErrorNotSupported;
  FusedOps *ops = new FusedOps();
  ops->Np = Np;
  ops->Nfp = s->v.Nfp;
  if ((e = cudaMalloc(&ops->Afrag, sizeof(double) * frag.size())) != cudaSuccess) { delete ops; return e; }
  if ((e = cudaMemcpy(ops->Afrag, frag.data(), sizeof(double) * frag.size(), cudaMemcpyHostToDevice)) != cudaSuccess) {
    cudaFree(ops->Afrag);
    delete ops;
    return e;
  }
  s->fused_ops = ops;
  return cudaSuccess;
}

void fused_release(bdg_sw2d *s) {
  if (!s->fused_ops) return;
  FusedOps *ops = static_cast<FusedOps *>(s->fused_ops);
  cudaFree(ops->Afrag);
  delete ops;
  s->fused_ops = nullptr;
}

cudaError_t launch_fused_stage(bdg_sw2d *s, const double *q_in, double *q_out, double *res, int k0, int k1, double a, double b,
                               int write_rhs, double *rhs_out, int slot) {
  if (k1 <= k0) return cudaSuccess;
  FusedOps *ops = static_cast<FusedOps *>(s->fused_ops);
  if (!ops) return cudaErrorNotSupported;
  StageArgs args;
  args.q_in = q_in; args.q_out = q_out; args.res = res; args.rhs_out = rhs_out; args.Afrag = ops->Afrag;
  args.k0 = k0; args.k1 = k1; args.tile0 = 0; args.ntiles = 0;
  args.write_rhs = write_rhs; args.slot = slot; args.a = a; args.b = b;
#define X(np, nfp) \
  if (ops->Np == np && ops->Nfp == nfp) return launch_t<np, nfp>(s, args);
  BDG_FUSED_ORDERS(X)
#undef X
  return cudaErrorNotSupported;
}
