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
  // shared memory carve-up (bytes, every section 16-B aligned)
  static constexpr size_t oQ = 0;
  static constexpr size_t oR = oQ + sizeof(double) * 3 * TNP;
  static constexpr size_t oFl = oR + sizeof(double) * 3 * TNP;
  static constexpr size_t oAff = oFl + sizeof(double) * 4 * TNP;
  static constexpr size_t oA = oAff + sizeof(double) * kTile * kAW;
  static constexpr size_t oVm = oA + (A_REGS ? 0 : sizeof(double) * MB * KB * 32);
  static constexpr size_t oTag = oVm + sizeof(int32_t) * kTile * NFN;
  static constexpr size_t oFm = oTag + ((kTile * 3 + 15) / 16) * 16;
  static constexpr size_t oBar = oFm + ((sizeof(int32_t) * NFN + 15) / 16) * 16;
  static constexpr size_t bytes = oBar + 32;
};

struct StageArgs {
  const double *q_in;
  double *q_out;
  double *res;
  double *rhs_out;      // write_rhs mode: [3][nT]
  const double *Afrag;  // [MB][KB][32] operator in DMMA A-fragment order
  int k0, k1;           // active element range
  int tile0;            // first tile index of this launch
  int write_rhs;
  int slot;             // scal slot of the global LF speed
  double a, b;          // LSERK4 stage coefficients
};

template <int NP, int NFP>
__global__ void __launch_bounds__(kThreads) k_fused_stage(Sw2dView v, StageArgs p) {
  using C = Cfg<NP, NFP>;
  constexpr int NFN = C::NFN, NPB = C::NPB, NFB = C::NFB, KB = C::KB, MB = C::MB, TNP = C::TNP;
  extern __shared__ __align__(128) unsigned char smem[];
  double *sQ = reinterpret_cast<double *>(smem + C::oQ);
  double *sR = reinterpret_cast<double *>(smem + C::oR);
  double *sFl = reinterpret_cast<double *>(smem + C::oFl);
  double *sAff = reinterpret_cast<double *>(smem + C::oAff);
  double *sA = reinterpret_cast<double *>(smem + C::oA);
  int32_t *sVm = reinterpret_cast<int32_t *>(smem + C::oVm);
  uint8_t *sTag = reinterpret_cast<uint8_t *>(smem + C::oTag);
  int32_t *sFm = reinterpret_cast<int32_t *>(smem + C::oFm);
  uint64_t *bar = reinterpret_cast<uint64_t *>(smem + C::oBar);

  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int c = lane >> 2, t = lane & 3;
  const int e = w + 4 * c;  // my element of the tile (B-fragment column c)
  const long long kt0 = (long long)(p.tile0 + blockIdx.x) * kTile;
  const int nload = (int)min((long long)kTile, (long long)v.K - kt0);
  const bool full_tile = nload == kTile;
  const bool all_active = full_tile && kt0 >= p.k0 && kt0 + kTile <= p.k1;
  const bool need_res = !p.write_rhs && p.a != 0.0;
  const long long stride = v.stride;
  const long long base = kt0 * NP;  // first volume node of the tile

  // ---- stage the tile ------------------------------------------------------------------------------
  if (full_tile) {
    if (tid == 0) {
      mbar_init(&bar[0], 1);
      mbar_init(&bar[1], 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid == 0) {
      constexpr uint32_t bq = sizeof(double) * TNP;
      constexpr uint32_t b0 = 3 * bq + sizeof(int32_t) * kTile * NFN + sizeof(double) * kTile * kAW + kTile * 3;
      mbar_expect_tx(&bar[0], b0);
      for (int f = 0; f < 3; ++f) bulk_g2s(sQ + f * TNP, p.q_in + f * stride + base, bq, &bar[0]);
      bulk_g2s(sVm, v.vmapP + kt0 * NFN, sizeof(int32_t) * kTile * NFN, &bar[0]);
      bulk_g2s(sAff, v.aff + kt0 * kAW, sizeof(double) * kTile * kAW, &bar[0]);
      bulk_g2s(sTag, v.facetag + kt0 * 3, kTile * 3, &bar[0]);
      if (need_res) {
        mbar_expect_tx(&bar[1], 3 * bq);
        for (int f = 0; f < 3; ++f) bulk_g2s(sR + f * TNP, p.res + f * stride + base, bq, &bar[1]);
      }
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
  if (tid < NFN) sFm[tid] = v.fmask[tid];

  // operator fragments: registers (small orders) or shared memory in fragment order
  double Areg[C::A_REGS ? MB * KB : 1];
  if (C::A_REGS) {
#pragma unroll
    for (int i = 0; i < MB * KB; ++i) Areg[i] = __ldg(p.Afrag + i * 32 + lane);
  } else {
    for (int x = tid; x < MB * KB * 32; x += kThreads) sA[x] = __ldg(p.Afrag + x);
  }
  const double g = v.g, half_g = 0.5 * v.g;

  if (full_tile) mbar_wait(&bar[0], 0);
  __syncthreads();

  // ---- phase 1: nodal fluxes and wave speed at my volume nodes (A6, A7 '-' side) ----------------------
  const int eo = e * NP;
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
  const double *af = sAff + e * kAW;
  double fn[NFB][3], dq[NFB][3];  // normal flux differences and jumps per slot
  double lam0 = 0.0, lam1 = 0.0, lam2 = 0.0;  // per-face max speed (scalars: a runtime face index must not spill)
  const int tile_nodes = nload * NP;
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
      const int tag = sTag[e * 3 + f];
      const long long mP = sVm[e * NFN + sl];
      const long long rel = mP - base;
      const bool in_tile = rel >= 0 && rel < tile_nodes;
      double hP, huP, hvP, F2P, G2P, G3P, spdP;
      if (in_tile && tag != 3) {  // neighbour inside the tile: everything is already in shared memory
        const int r = (int)rel;
        hP = sQ[r]; huP = sQ[TNP + r]; hvP = sQ[2 * TNP + r];
        F2P = sFl[r]; G2P = sFl[TNP + r]; G3P = sFl[2 * TNP + r]; spdP = sFl[3 * TNP + r];
      } else {
        if (tag == 3) {  // Wall (main.cpp:340-345)
          const double d = huM * nx + hvM * ny;
          hP = hM;
          huP = fma(-2.0 * nx, d, huM);
          hvP = fma(-2.0 * ny, d, hvM);
        } else {  // neighbour tile or ghost slot: L2 gather
          hP = __ldg(p.q_in + mP); huP = __ldg(p.q_in + stride + mP); hvP = __ldg(p.q_in + 2 * stride + mP);
        }
        const double r = frcp(hP), u = huP * r, vv = hvP * r, gh2 = (half_g * hP) * hP;
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
  if (full_tile && need_res) mbar_wait(&bar[1], 0);
  const int gr = lane >> 2;                       // C fragment row within a row block
  const int e0 = w + 4 * (2 * t), e1 = e0 + 4;    // C fragment columns 2t, 2t+1
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const double *X = i == 0 ? sQ + TNP : (i == 1 ? sFl : sFl + TNP);          // F_i: hu, F2, G2
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
  if (p.write_rhs) return;

  // ---- write the tile back ----------------------------------------------------------------------------
  if (all_active) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (tid == 0) {
      constexpr uint32_t bq = sizeof(double) * TNP;
      for (int f = 0; f < 3; ++f) {
        bulk_s2g(p.q_out + f * stride + base, sQ + f * TNP, bq);
        bulk_s2g(p.res + f * stride + base, sR + f * TNP, bq);
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

struct FusedOps {
  double *Afrag;
  int Np, Nfp;
};

template <int NP, int NFP>
cudaError_t launch_t(bdg_sw2d *s, const StageArgs &a, int ntiles) {
  using C = Cfg<NP, NFP>;
  static bool attr_set[16] = {};
  if (!attr_set[s->device & 15]) {
    cudaError_t e = cudaFuncSetAttribute(k_fused_stage<NP, NFP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::bytes);
    if (e != cudaSuccess) return e;
    attr_set[s->device & 15] = true;
  }
  k_fused_stage<NP, NFP><<<ntiles, kThreads, C::bytes, s->stream>>>(s->v, a);
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
  args.k0 = k0; args.k1 = k1; args.tile0 = k0 / kTile;
  args.write_rhs = write_rhs; args.slot = slot; args.a = a; args.b = b;
  const int ntiles = (k1 - 1) / kTile - args.tile0 + 1;
  s->launches++;
#define X(np, nfp) \
  if (ops->Np == np && ops->Nfp == nfp) return launch_t<np, nfp>(s, args, ntiles);
  BDG_FUSED_ORDERS(X)
#undef X
  return cudaErrorNotSupported;
}
