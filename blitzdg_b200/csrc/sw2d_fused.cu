// Fused sw2d stage kernel for straight-sided (affine) triangles: ONE launch per LSERK4 stage does
//   trace gather (vmapP) + BC + Lax-Friedrichs flux            (reference src/sw2d-simple/main.cpp:246-319, src/sw2d/main.cpp:327-421)
//   volume term + LIFT*(Fscale*flux)                            (:332-355 / :434-458)
//   res = a*res + dt*RHS ; q += b*res                           (src/advec1d/main.cpp:98,101)
// reading q/res once and writing them once (B_alg of SURVEY.md §8d).
//
// Tiles.  A CTA of 4 warps works on tiles of 32 consecutive elements.  At create time every tile gets
//   * a LOCAL trace map: for each face node the index of its '+' node inside the tile's node array, either one
//     of the tile's own 32*Np nodes or one of the tile's GHOST nodes (the distinct '+' nodes that live in other
//     tiles or in off-rank halo slots), packed with the face's BC tag in one 32-bit word, and
//   * the ghost source list (global node ids).
// A tile in shared memory is therefore a closed problem: h,hu,hv of own + ghost nodes, res, 13 geometry doubles
// per element, the local map.  Own nodes / res / geometry / map arrive through 1-D TMA bulk copies
// (cp.async.bulk + mbarrier), ghost nodes through 8-byte cp.async gathers from L2 that signal the same mbarrier;
// results leave through TMA bulk stores.  The persistent kernel double-buffers tiles so the next tile's copies
// and gathers fly while the current one is computed.
//
// Arithmetic.  Because the element is affine,
//   RHS_i = -( Dr (rx F_i + ry G_i) + Ds (sx F_i + sy G_i) ) + LIFT (Fscale dF_i)
// is one small GEMM  C[Np x cols] = A[Np x Kp] * B[Kp x cols]  with the constant operator
//   A = [ -Dr | -Ds | LIFT ]   (each block padded to a multiple of 4 columns, rows to a multiple of 8)
// run on the fp64 tensor pipe (mma.sync.m8n8k4.f64 = DMMA) with A fragments in registers (N<=4) or in shared
// memory in fragment order.  Phase 1 evaluates nodal fluxes and wave speeds for own AND ghost nodes with a flat,
// branch-free thread->node mapping; phase 2 builds the LIFT-part B fragments directly in fragment layout
//   lane = 4*c + t  ->  element E(c) of the tile, face slots j = 4*kb + t
// so the 4 lanes of a quad hold all face nodes of one element and the per-face max of the wave speed
// (sw2d-simple main.cpp:305-312) is two xor-shuffles; phase 3 builds the volume B fragments, issues the DMMAs and
// applies the stage update in place in shared memory.
//
// Deviations from the reference's operation order (inside the parity tolerance, SURVEY.md §7 hard part 1):
// FMA contraction, x/h evaluated as x*(1/h) with a Newton-refined reciprocal, chain-rule factors applied
// before the contraction (exact for affine elements up to round-off).
#include <algorithm>
#include <cstdio>
#include <type_traits>
#include <vector>

#include "sw2d_common.cuh"

namespace {

constexpr int kWarps = 4;
constexpr int kThreads = kWarps * 32;
constexpr int kTile = kWarps * 8;  // elements per CTA tile
constexpr int kAW = 13;            // doubles per element in v.aff: rx,sx,ry,sy,(nx,ny,Fscale)x3
constexpr int kGI = 5;             // ghost indices a thread can carry: Gmax <= kGI*kThreads

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
// 8-byte asynchronous gather global -> shared (LDGSTS); completion through cp_async_arrive on an mbarrier
__device__ __forceinline__ void cp_async8(void *dst, const void *src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_arrive(uint64_t *bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void dmma884(double &d0, double &d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

__device__ __forceinline__ double dmax(double a, double b) { return a > b ? a : b; }

// 1/x: MUFU seed (20+ bits) + two Newton steps; no denormal / inf handling (h > 0 here)
__device__ __forceinline__ double frcp(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  double e = fma(-x, r, 1.0);
  r = fma(r, e, r);
  e = fma(-x, r, 1.0);
  return fma(r, e, r);
}
// sqrt(x), x > 0: coupled Newton on (sqrt, 1/(2 sqrt)) from the MUFU rsqrt seed, two final corrections
__device__ __forceinline__ double fsqrt(double x) {
  double r;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  double hh = 0.5 * r, s = x * r;
  double e = fma(-hh, s, 0.5);
  s = fma(s, e, s);
  hh = fma(hh, e, hh);
  e = fma(-s, s, x);
  s = fma(e, hh, s);
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
  static constexpr int TNP = kTile * NP;      // own nodes per tile
  static constexpr uint32_t txQ = sizeof(double) * TNP;  // one field plane of a tile
  static constexpr uint32_t txMeta = sizeof(uint32_t) * kTile * NFN + sizeof(double) * kTile * kAW;
  static constexpr size_t sizeA = A_REGS ? 0 : sizeof(double) * MB * KB * 32;
  // per tile buffer: Q[3][PS], Aff[kTile][13], Lm[kTile][NFN]; CTA-wide: R[3][TNP], Fl[4][PS], A, barriers
  static size_t buf_bytes(int G) { return sizeof(double) * 3 * (TNP + G) + sizeof(double) * kTile * kAW + sizeof(uint32_t) * kTile * NFN; }
  static size_t total(int nbuf, int G) { return nbuf * buf_bytes(G) + sizeof(double) * 3 * TNP + sizeof(double) * 4 * (TNP + G) + sizeA + 64; }
};

struct StageArgs {
  const double *q_in;
  double *q_out;
  double *res;
  double *rhs_out;        // write_rhs mode: [3][nT]
  const double *Afrag;    // [MB][KB][32] operator in DMMA A-fragment order
  const uint32_t *lmap;   // [ntiles_all*kTile][NFN]: local '+' node (low 16 bits) | face tag << 16
  const int32_t *gflat;   // ghost node sources (global node ids) of all tiles back to back
  const long long *goff;  // [ntiles_all + 1]: tile T owns gflat[goff[T] .. goff[T+1])
  const int4 *rec;        // persistent kernel: regular tiles {tile id, ghost count, goff low / high 32 bits}
  const int32_t *biglist; // edge kernel in list mode: tile ids with more than Gcap ghosts
  int G;                  // ghost slots per tile in shared memory (Gcap: persistent kernel, Gmax: edge kernel), multiple of 8
  int list_first, use_list;  // first list position of this launch (rec / biglist); edge kernel: 0 = tiles tile0 + blockIdx.x
  int k0, k1;             // active element range
  int tile0, ntiles;      // tiles [tile0, tile0 + ntiles) of this launch
  int write_rhs;
  int slot;               // scal slot of the global LF speed
  double a, b;            // LSERK4 stage coefficients
};

template <int NP, int NFP>
struct Smem {
  using C = Cfg<NP, NFP>;
  unsigned char *base;
  int nbuf, G;
  __device__ int PS() const { return C::TNP + G; }
  __device__ size_t bufb() const { return sizeof(double) * 3 * PS() + sizeof(double) * kTile * kAW + sizeof(uint32_t) * kTile * C::NFN; }
  __device__ double *Q(int b) const { return reinterpret_cast<double *>(base + b * bufb()); }
  __device__ double *Aff(int b) const { return Q(b) + 3 * PS(); }
  __device__ uint32_t *Lm(int b) const { return reinterpret_cast<uint32_t *>(Aff(b) + kTile * kAW); }
  __device__ double *R() const { return reinterpret_cast<double *>(base + nbuf * bufb()); }
  __device__ double *Fl() const { return R() + 3 * C::TNP; }
  __device__ double *A() const { return Fl() + 4 * PS(); }
  __device__ uint64_t *Bar() const { return reinterpret_cast<uint64_t *>(reinterpret_cast<unsigned char *>(A()) + C::sizeA); }
};

// thread 0: TMA bulk loads of one full tile's own nodes, geometry and local map into buffer b
template <int NP, int NFP>
__device__ __forceinline__ void issue_tile_load(const Sw2dView &v, const StageArgs &p, const Smem<NP, NFP> &S, int b, long long tile,
                                                uint64_t *bar) {
  using C = Cfg<NP, NFP>;
  const long long kt0 = tile * kTile, base = kt0 * NP;
  mbar_expect_tx(bar, 3 * C::txQ + C::txMeta);
  bulk_g2s(S.Lm(b), p.lmap + kt0 * C::NFN, sizeof(uint32_t) * kTile * C::NFN, bar);
  bulk_g2s(S.Aff(b), v.aff + kt0 * kAW, sizeof(double) * kTile * kAW, bar);
  for (int f = 0; f < 3; ++f) bulk_g2s(S.Q(b) + f * S.PS(), p.q_in + f * v.stride + base, C::txQ, bar);
}
template <int NP, int NFP>
__device__ __forceinline__ void issue_res_load(const Sw2dView &v, const StageArgs &p, const Smem<NP, NFP> &S, long long tile, uint64_t *bar) {
  using C = Cfg<NP, NFP>;
  const long long base = tile * kTile * NP;
  mbar_expect_tx(bar, 3 * C::txQ);
  for (int f = 0; f < 3; ++f) bulk_g2s(S.R() + f * C::TNP, p.res + f * v.stride + base, C::txQ, bar);
}

// The arithmetic of one tile that sits in shared memory (A1-A13 of SURVEY.md Appendix A).  sQ: [3][PS] own +
// ghost nodes, sR: [3][TNP].  On return the tile's q and res planes hold the updated values (stage mode) -- or,
// in write_rhs mode, rhs_out has been written and the tile is untouched.  Contains two __syncthreads() and an
// optional wait on the res barrier before phase 3; the caller makes the tile visible first.
template <int NP, int NFP>
__device__ __forceinline__ void compute_tile(const Sw2dView &v, const StageArgs &p, double *__restrict__ sQ, double *__restrict__ sR,
                                             const double *__restrict__ sAff, const uint32_t *__restrict__ sLm, double *__restrict__ sFl,
                                             const double *__restrict__ sA, const double *Areg, const int *fm, int PS, int nnodes, long long kt0,
                                             bool need_res, uint64_t *bar_res, uint32_t res_parity) {
  using C = Cfg<NP, NFP>;
  constexpr int NFN = C::NFN, NPB = C::NPB, NFB = C::NFB, KB = C::KB, MB = C::MB, TNP = C::TNP;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int c = lane >> 2, t = lane & 3;
  const int e = w + 4 * (c < 4 ? c : (c ^ 1));  // my element (B-fragment column c); E(c) keeps B and C accesses conflict-free
  const int eo = e * NP;
  const double g = v.g, half_g = 0.5 * v.g;
  const double *af = sAff + e * kAW;

  // ---- phase 1: nodal fluxes and wave speed at own and ghost nodes (A6, A7), flat mapping ------------------
#pragma unroll 1
  for (int j0 = threadIdx.x; j0 < nnodes; j0 += 3 * kThreads) {  // nnodes = own + this tile's ghost nodes
#pragma unroll
    for (int u3 = 0; u3 < 3; ++u3) {
      const int j = j0 + u3 * kThreads;
      if (j < nnodes) {
        const double h = sQ[j], hu = sQ[PS + j], hv = sQ[2 * PS + j];
        const double r = frcp(h), u = hu * r, vv = hv * r, gh2 = (half_g * h) * h;
        sFl[j] = fma(hu, u, gh2);            // F2
        sFl[PS + j] = hu * vv;               // G2 = F3
        sFl[2 * PS + j] = fma(hv, vv, gh2);  // G3
        sFl[3 * PS + j] = fsqrt(fma(u, u, fma(vv, vv, 1e-300))) + fsqrt(g * h);
      }
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
      const int m = eo + fm[kb];
      const uint32_t word = sLm[e * NFN + sl];
      const int r = word & 0xffffu;
      const double nx = af[4 + 3 * f], ny = af[5 + 3 * f];
      const double hM = sQ[m], huM = sQ[PS + m], hvM = sQ[2 * PS + m];
      const double F2M = sFl[m], G2M = sFl[PS + m], G3M = sFl[2 * PS + m], spdM = sFl[3 * PS + m];
      double hP = sQ[r], huP = sQ[PS + r], hvP = sQ[2 * PS + r];
      double F2P = sFl[r], G2P = sFl[PS + r], G3P = sFl[2 * PS + r], spdP = sFl[3 * PS + r];
      if ((word >> 16) == 3u) {  // Wall (main.cpp:340-345): reflect the normal momentum of the '-' trace
        const double d = huM * nx + hvM * ny;
        hP = hM;
        huP = fma(-2.0 * nx, d, huM);
        hvP = fma(-2.0 * ny, d, hvM);
        const double rr = frcp(hP), u = huP * rr, vv = hvP * rr, gh2 = (half_g * hP) * hP;
        F2P = fma(huP, u, gh2); G2P = huP * vv; G3P = fma(hvP, vv, gh2);
        spdP = fsqrt(fma(u, u, fma(vv, vv, 1e-300))) + fsqrt(g * hP);
      }
      const double sp = dmax(spdM, spdP);
      lam0 = f == 0 ? dmax(lam0, sp) : lam0;
      lam1 = f == 1 ? dmax(lam1, sp) : lam1;
      lam2 = f == 2 ? dmax(lam2, sp) : lam2;
      dq[kb][0] = hM - hP; dq[kb][1] = huM - huP; dq[kb][2] = hvM - hvP;
      const double dG2 = G2M - G2P;
      fn[kb][0] = fma(dq[kb][1], nx, dq[kb][2] * ny);
      fn[kb][1] = fma(F2M - F2P, nx, dG2 * ny);
      fn[kb][2] = fma(dG2, nx, (G3M - G3P) * ny);
    }
  }
  if (v.lf_global) {
    lam0 = lam1 = lam2 = v.scal[p.slot];
  } else {  // per-face max over the quad (the 4 lanes of a quad hold all face nodes of element e)
    lam0 = dmax(lam0, __shfl_xor_sync(0xffffffffu, lam0, 1));
    lam1 = dmax(lam1, __shfl_xor_sync(0xffffffffu, lam1, 1));
    lam2 = dmax(lam2, __shfl_xor_sync(0xffffffffu, lam2, 1));
    lam0 = dmax(lam0, __shfl_xor_sync(0xffffffffu, lam0, 2));
    lam1 = dmax(lam1, __shfl_xor_sync(0xffffffffu, lam1, 2));
    lam2 = dmax(lam2, __shfl_xor_sync(0xffffffffu, lam2, 2));
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
  if (bar_res) mbar_wait(bar_res, res_parity);
  const int gr = lane >> 2;  // C fragment row within a row block
  const int c0 = 2 * t, c1 = 2 * t + 1;
  const int e0 = w + 4 * (c0 < 4 ? c0 : (c0 ^ 1));  // C fragment column 2t   -> element E(2t)
  const int e1 = w + 4 * (c1 < 4 ? c1 : (c1 ^ 1));  // C fragment column 2t+1 -> element E(2t+1)
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const double *X = i == 0 ? sQ + PS : (i == 1 ? sFl : sFl + PS);               // F_i: hu, F2, G2
    const double *Y = i == 0 ? sQ + 2 * PS : (i == 1 ? sFl + PS : sFl + 2 * PS);  // G_i: hv, G2, G3
    double acc[MB][2][2];  // two independent DMMA chains per row block
#pragma unroll
    for (int rb = 0; rb < MB; ++rb) acc[rb][0][0] = acc[rb][0][1] = acc[rb][1][0] = acc[rb][1][1] = 0.0;
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
        dmma884(acc[rb][0][0], acc[rb][0][1], a1, ba);
        dmma884(acc[rb][1][0], acc[rb][1][1], a2, bb);
      }
    }
#pragma unroll
    for (int kb = 0; kb < NFB; ++kb) {
#pragma unroll
      for (int rb = 0; rb < MB; ++rb) {
        const double a3 = C::A_REGS ? Areg[rb * KB + 2 * NPB + kb] : sA[(rb * KB + 2 * NPB + kb) * 32 + lane];
        dmma884(acc[rb][kb & 1][0], acc[rb][kb & 1][1], a3, sfl[kb][i]);
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
          const double rhs = acc[rb][0][cc] + acc[rb][1][cc];
          if (p.write_rhs) {
            const long long k = kt0 + (cc ? e1 : e0);
            if (k >= p.k0 && k < p.k1) p.rhs_out[(long long)i * NP * v.K + k * NP + n] = rhs;
          } else {
            double rs = dt * rhs;
            if (need_res) rs = fma(p.a, sR[i * TNP + o], rs);
            sR[i * TNP + o] = rs;
            sQ[i * PS + o] = fma(p.b, rs, sQ[i * PS + o]);
          }
        }
      }
    }
  }
}

// per-thread constants: operator fragments, '-' node offsets of my face slots
template <int NP, int NFP>
__device__ __forceinline__ void load_constants(const StageArgs &p, const Sw2dView &v, double *Areg, double *sA, int *fm) {
  using C = Cfg<NP, NFP>;
  const int tid = threadIdx.x, lane = tid & 31;
  if (C::A_REGS) {
#pragma unroll
    for (int i = 0; i < C::MB * C::KB; ++i) Areg[i] = __ldg(p.Afrag + i * 32 + lane);
  } else {
    for (int x = tid; x < C::MB * C::KB * 32; x += kThreads) sA[x] = __ldg(p.Afrag + x);
  }
#pragma unroll
  for (int kb = 0; kb < C::NFB; ++kb) {
    const int sl = 4 * kb + (lane & 3);
    fm[kb] = sl < C::NFN ? __ldg(v.fmask + sl) : 0;
  }
}

// ---- edge kernel: one tile per CTA, any tile (ragged mesh end, partially active range, write_rhs, tiles with
// more ghost nodes than the persistent kernel's buffers hold) --------------------------------------------------
template <int NP, int NFP>
__global__ void __launch_bounds__(kThreads, 3) k_fused_edge(Sw2dView v, StageArgs p) {
  using C = Cfg<NP, NFP>;
  constexpr int NFN = C::NFN, TNP = C::TNP;
  extern __shared__ __align__(128) unsigned char smem[];
  const Smem<NP, NFP> S{smem, 1, p.G};
  const int PS = S.PS();
  double *sQ = S.Q(0), *sR = S.R(), *sAff = S.Aff(0);
  uint32_t *sLm = S.Lm(0);

  const int tid = threadIdx.x;
  const long long tile = p.use_list ? p.biglist[p.list_first + blockIdx.x] : p.tile0 + blockIdx.x;
  const long long kt0 = tile * kTile;
  const int nload = (int)min((long long)kTile, (long long)v.K - kt0);
  const bool all_active = nload == kTile && kt0 >= p.k0 && kt0 + kTile <= p.k1;
  const bool need_res = !p.write_rhs && p.a != 0.0;
  const long long stride = v.stride;
  const long long base = kt0 * NP;
  const long long g0 = p.goff[tile];
  const int gc = (int)(p.goff[tile + 1] - g0);

  // plain loads; neutral fill (h = 1) behind the mesh end; lmap is padded to whole tiles at create
  for (int x = tid; x < 3 * TNP; x += kThreads) {
    const int f = x / TNP, r = x - f * TNP;
    const bool in = r < nload * NP;
    sQ[f * PS + r] = in ? p.q_in[f * stride + base + r] : (f == 0 ? 1.0 : 0.0);
    sR[x] = (in && need_res) ? p.res[f * stride + base + r] : 0.0;
  }
  for (int x = tid; x < gc; x += kThreads) {
    const long long m = p.gflat[g0 + x];
    for (int f = 0; f < 3; ++f) sQ[f * PS + TNP + x] = p.q_in[f * stride + m];
  }
  for (int x = tid; x < kTile * NFN; x += kThreads) sLm[x] = p.lmap[kt0 * NFN + x];
  for (int x = tid; x < kTile * kAW; x += kThreads) sAff[x] = x < nload * kAW ? v.aff[kt0 * kAW + x] : 0.0;
  double Areg[C::A_REGS ? C::MB * C::KB : 1];
  int fm[C::NFB];
  load_constants<NP, NFP>(p, v, Areg, S.A(), fm);
  __syncthreads();

  compute_tile<NP, NFP>(v, p, sQ, sR, sAff, sLm, S.Fl(), S.A(), Areg, fm, PS, TNP + gc, kt0, need_res, nullptr, 0);
  if (p.write_rhs) return;

  if (all_active) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (tid == 0) {
      for (int f = 0; f < 3; ++f) {
        bulk_s2g(p.q_out + f * stride + base, sQ + f * PS, C::txQ);
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
        p.q_out[f * stride + base + r] = sQ[f * PS + r];
        p.res[f * stride + base + r] = sR[x];
      }
    }
  }
}

// ---- persistent kernel: full, fully active tiles with at most Gcap ghosts, addressed through the record list
// rec[] = {tile id, ghost count, ghost offset}.  Two tile buffers: while tile `it` is computed, the TMA copies of
// tile it+1 (issued at the end of iteration it-1) and its ghost gathers (issued at the top of iteration it) are in
// flight; results leave through TMA bulk stores from the same buffer. ---------------------------------------------
template <int NP, int NFP>
__global__ void __launch_bounds__(kThreads, 3) k_fused_persistent(Sw2dView v, StageArgs p) {
  using C = Cfg<NP, NFP>;
  constexpr int TNP = C::TNP;
  extern __shared__ __align__(128) unsigned char smem[];
  const Smem<NP, NFP> S{smem, 2, p.G};
  const int PS = S.PS();
  uint64_t *bar = S.Bar();  // [0],[1]: tile buffers (1 expect_tx arrival + kThreads cp.async arrivals); [2]: res
  const int tid = threadIdx.x;
  const bool need_res = p.a != 0.0;
  const long long stride = v.stride;
  const int nmine = (p.ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;  // list positions blockIdx.x + it*gridDim.x
  auto rec_of = [&](int it) { return __ldg(p.rec + p.list_first + blockIdx.x + (long long)it * gridDim.x); };
  // all threads: 8-byte gathers of a tile's ghost nodes into buffer b, then one arrival on its barrier
  auto gather = [&](int b, int gcount, const int *ids) {
    double *dst = S.Q(b) + TNP;
#pragma unroll
    for (int u = 0; u < kGI; ++u) {
      const int x = tid + u * kThreads;
      if (x < gcount)
        for (int f = 0; f < 3; ++f) cp_async8(dst + f * PS + x, p.q_in + f * stride + ids[u]);
    }
    cp_async_arrive(&bar[b]);
  };
  auto load_ids = [&](const int4 &r, int *ids) {
    const int32_t *gs = p.gflat + (((long long)r.w << 32) | (unsigned)r.z);
#pragma unroll
    for (int u = 0; u < kGI; ++u) ids[u] = (tid + u * kThreads < r.y) ? __ldg(gs + tid + u * kThreads) : 0;
  };

  if (tid == 0) {
    mbar_init(&bar[0], 1 + kThreads);
    mbar_init(&bar[1], 1 + kThreads);
    mbar_init(&bar[2], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  int4 rcur = rec_of(0);                              // tile being computed
  int4 rnext = nmine > 1 ? rec_of(1) : rcur;          // tile whose copies are in flight
  int4 rnext2 = nmine > 2 ? rec_of(2) : rcur;         // tile after that
  if (tid == 0) {
    issue_tile_load<NP, NFP>(v, p, S, 0, rcur.x, &bar[0]);
    if (nmine > 1) issue_tile_load<NP, NFP>(v, p, S, 1, rnext.x, &bar[1]);
    if (need_res) issue_res_load<NP, NFP>(v, p, S, rcur.x, &bar[2]);
  }
  int gi[kGI];  // ghost ids of tile `rnext`
  load_ids(rcur, gi);
  gather(0, rcur.y, gi);
  if (nmine > 1) load_ids(rnext, gi);
  double Areg[C::A_REGS ? C::MB * C::KB : 1];
  int fm[C::NFB];
  load_constants<NP, NFP>(p, v, Areg, S.A(), fm);
  __syncthreads();

  for (int it = 0; it < nmine; ++it) {
    const int b = it & 1;
    const long long kt0 = (long long)rcur.x * kTile;
    if (it + 1 < nmine) {  // ghost gathers of the next tile into the other buffer, ids of the one after into registers
      gather(b ^ 1, rnext.y, gi);
      if (it + 2 < nmine) load_ids(rnext2, gi);
    }
    mbar_wait(&bar[b], (it >> 1) & 1);
    compute_tile<NP, NFP>(v, p, S.Q(b), S.R(), S.Aff(b), S.Lm(b), S.Fl(), S.A(), Areg, fm, PS, TNP + rcur.y, kt0, need_res,
                          need_res ? &bar[2] : nullptr, it & 1);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (tid == 0) {
      const long long base = kt0 * NP;
      for (int f = 0; f < 3; ++f) {
        bulk_s2g(p.q_out + f * stride + base, S.Q(b) + f * PS, C::txQ);
        bulk_s2g(p.res + f * stride + base, S.R() + f * TNP, C::txQ);
      }
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // buffer b and R drained -> refill them
      if (it + 2 < nmine) issue_tile_load<NP, NFP>(v, p, S, b, rnext2.x, &bar[b]);
      if (need_res && it + 1 < nmine) issue_res_load<NP, NFP>(v, p, S, rnext.x, &bar[2]);
    }
    rcur = rnext;
    rnext = rnext2;
    if (it + 3 < nmine) rnext2 = rec_of(it + 3);
  }
}

struct FusedOps {
  double *Afrag;
  uint32_t *lmap;
  int32_t *gflat;
  long long *goff;
  int4 *rec;
  int32_t *biglist;
  std::vector<int> reg_tiles, big_tiles;  // host copies (sorted tile ids) to cut launch ranges
  int Gcap, Gmax;                         // ghost slots of the persistent / edge kernel buffers (multiples of 8)
  int Np, Nfp;
};

int g_sm_count[16] = {};

template <int NP, int NFP>
cudaError_t launch_t(bdg_sw2d *s, StageArgs a, const FusedOps *ops) {
  using C = Cfg<NP, NFP>;
  static int ctas_per_sm[16] = {};
  static int attr_key[16] = {};
  const int dev = s->device & 15;
  const size_t smem_edge = C::total(1, ops->Gmax), smem_pers = C::total(2, ops->Gcap);
  cudaError_t e;
  if (attr_key[dev] != ops->Gmax * 4096 + ops->Gcap + 1) {
    if ((e = cudaFuncSetAttribute(k_fused_edge<NP, NFP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_edge)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_fused_persistent<NP, NFP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_pers)) != cudaSuccess) return e;
    cudaFuncSetAttribute(k_fused_edge<NP, NFP>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaFuncSetAttribute(k_fused_persistent<NP, NFP>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm[dev], k_fused_persistent<NP, NFP>, kThreads, smem_pers)) != cudaSuccess) return e;
    if ((e = cudaDeviceGetAttribute(&g_sm_count[dev], cudaDevAttrMultiProcessorCount, s->device)) != cudaSuccess) return e;
    if (ctas_per_sm[dev] < 1) return cudaErrorLaunchOutOfResources;
    attr_key[dev] = ops->Gmax * 4096 + ops->Gcap + 1;
  }
  // tiles [t0,t1) cover [k0,k1); the fully active ones [p0,p1) (k1 <= K, so they are also full) go to the
  // persistent kernel when they are regular (<= Gcap ghosts); ragged / partially active head and tail tiles and
  // the big tiles inside [p0,p1) go to the edge kernel
  const int t0 = a.k0 / kTile, t1 = (a.k1 - 1) / kTile + 1;
  int p0 = (a.k0 + kTile - 1) / kTile, p1 = a.k1 / kTile;
  if (a.write_rhs || p1 <= p0) p0 = p1 = t1;
  auto edge = [&](int first, int count, int use_list) {
    if (count <= 0) return;
    StageArgs h = a;
    h.G = ops->Gmax;
    h.use_list = use_list;
    h.tile0 = use_list ? 0 : first;
    h.list_first = use_list ? first : 0;
    h.ntiles = count;
    k_fused_edge<NP, NFP><<<count, kThreads, smem_edge, s->stream>>>(s->v, h);
    s->launches++;
  };
  edge(t0, p0 - t0, 0);
  if (p1 > p0) {
    const int jr0 = (int)(std::lower_bound(ops->reg_tiles.begin(), ops->reg_tiles.end(), p0) - ops->reg_tiles.begin());
    const int jr1 = (int)(std::lower_bound(ops->reg_tiles.begin(), ops->reg_tiles.end(), p1) - ops->reg_tiles.begin());
    if (jr1 > jr0) {
      StageArgs m = a;
      m.G = ops->Gcap;
      m.list_first = jr0; m.ntiles = jr1 - jr0; m.use_list = 1; m.tile0 = 0;
      int grid = g_sm_count[dev] * ctas_per_sm[dev];
      // while a halo exchange is in flight leave a few SMs one CTA short, so the NCCL send/recv kernel finds room
      if (s->comm && !s->peer_ok && s->n_nbr > 0 && a.k0 == 0 && a.k1 < s->v.K) grid -= 16;
      if (grid > m.ntiles) grid = m.ntiles;
      k_fused_persistent<NP, NFP><<<grid, kThreads, smem_pers, s->stream>>>(s->v, m);
      s->launches++;
    }
    const int jb0 = (int)(std::lower_bound(ops->big_tiles.begin(), ops->big_tiles.end(), p0) - ops->big_tiles.begin());
    const int jb1 = (int)(std::lower_bound(ops->big_tiles.begin(), ops->big_tiles.end(), p1) - ops->big_tiles.begin());
    edge(jb0, jb1 - jb0, 1);
  }
  edge(p1, t1 - p1, 0);
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

// Builds the operator fragments and the per-tile local trace maps / ghost lists.  vmapP and facetag are the
// host arrays of create (vmapP entries >= Np*K address halo slots and are always ghosts).  Sets fused_ops on
// success; leaves it NULL (generic path) when a tile needs more ghost slots than the kernel can carry.
cudaError_t fused_prepare(bdg_sw2d *s, const double *Dr, const double *Ds, const double *Lift, const int32_t *vmapP,
                          const uint8_t *facetag) {
  const int Np = s->v.Np, Nfp = s->v.Nfp, Nfn = s->v.Nfn, K = s->v.K;
  std::vector<double> frag;
#define X(np, nfp) \
  if (Np == np && Nfp == nfp) build_afrag<np, nfp>(Dr, Ds, Lift, frag);
  BDG_FUSED_ORDERS(X)
#undef X
  if (frag.empty()) return cudaErrorNotSupported;

  const int ntiles = (K + kTile - 1) / kTile;
  const int tnp = kTile * Np;
  std::vector<uint32_t> lmap((size_t)ntiles * kTile * Nfn);
  std::vector<int32_t> gflat;           // ghosts of all tiles back to back
  std::vector<size_t> goff(ntiles + 1, 0);
  int Gmax = 0;
  std::vector<int32_t> ext;
  for (int T = 0; T < ntiles; ++T) {
    const long long kt0 = (long long)T * kTile, base = kt0 * Np;
    ext.clear();
    for (int el = 0; el < kTile && kt0 + el < K; ++el)
      for (int sl = 0; sl < Nfn; ++sl) {
        const int32_t mP = vmapP[(kt0 + el) * Nfn + sl];
        const long long r = (long long)mP - base;
        if (r < 0 || r >= tnp) ext.push_back(mP);
      }
    std::sort(ext.begin(), ext.end());
    ext.erase(std::unique(ext.begin(), ext.end()), ext.end());
    gflat.insert(gflat.end(), ext.begin(), ext.end());
    goff[T + 1] = gflat.size();
    Gmax = std::max(Gmax, (int)ext.size());
    for (int el = 0; el < kTile; ++el)
      for (int sl = 0; sl < Nfn; ++sl) {
        uint32_t word;
        if (kt0 + el >= K) {
          word = (uint32_t)(el * Np);  // padding element behind the mesh end: any valid own node
        } else {
          const int32_t mP = vmapP[(kt0 + el) * Nfn + sl];
          const long long r = (long long)mP - base;
          uint32_t loc;
          if (r >= 0 && r < tnp)
            loc = (uint32_t)r;
          else
            loc = (uint32_t)(tnp + (std::lower_bound(ext.begin(), ext.end(), mP) - ext.begin()));
          word = loc | ((uint32_t)facetag[(kt0 + el) * 3 + sl / Nfp] << 16);
        }
        lmap[((size_t)kt0 + el) * Nfn + sl] = word;
      }
  }
  // Buffer sizing: the persistent kernel's tile buffers hold Gcap ghosts, chosen so that (almost) every tile
  // fits -- a few outliers (e.g. tiles whose elements are not neighbours of each other after a partition's
  // interior-first renumbering) must not dictate the shared memory and phase-1 work of all tiles.
  std::vector<int> cnt(ntiles);
  for (int T = 0; T < ntiles; ++T) cnt[T] = (int)(goff[T + 1] - goff[T]);
  std::vector<int> sorted = cnt;
  std::sort(sorted.begin(), sorted.end());
  int cap = sorted[(size_t)((ntiles - 1) * 0.97)];
  if (Gmax - cap <= 16) cap = Gmax;
  const int Gcap = std::max(8, (cap + 7) / 8 * 8), Gmax8 = std::max(8, (Gmax + 7) / 8 * 8);
  if (Gmax8 > kGI * kThreads || tnp + Gmax8 > 65535) return cudaSuccess;  // not representable: stay on the generic path
  std::vector<int4> rec;
  FusedOps *ops = new FusedOps();
  for (int T = 0; T < ntiles; ++T) {
    if (cnt[T] <= Gcap) {
      ops->reg_tiles.push_back(T);
      rec.push_back(make_int4(T, cnt[T], (int)(goff[T] & 0xffffffffLL), (int)(goff[T] >> 32)));
    } else {
      ops->big_tiles.push_back(T);
    }
  }
  std::vector<long long> goff64(goff.begin(), goff.end());
  if (gflat.empty()) gflat.push_back(0);
  std::vector<int32_t> big(ops->big_tiles.begin(), ops->big_tiles.end());
  if (big.empty()) big.push_back(0);
  if (rec.empty()) rec.push_back(make_int4(0, 0, 0, 0));

  ops->Np = Np; ops->Nfp = Nfp; ops->Gcap = Gcap; ops->Gmax = Gmax8;
  ops->Afrag = nullptr; ops->lmap = nullptr; ops->gflat = nullptr; ops->goff = nullptr; ops->rec = nullptr; ops->biglist = nullptr;
  cudaError_t e;
  auto up = [&](auto **dst, const auto &vec) -> cudaError_t {
    using T = typename std::remove_reference<decltype(vec[0])>::type;
    cudaError_t r = cudaMalloc((void **)dst, sizeof(T) * vec.size());
    if (r != cudaSuccess) return r;
    return cudaMemcpy((void *)*dst, vec.data(), sizeof(T) * vec.size(), cudaMemcpyHostToDevice);
  };
  if ((e = up(&ops->Afrag, frag)) != cudaSuccess || (e = up(&ops->lmap, lmap)) != cudaSuccess ||
      (e = up(&ops->gflat, gflat)) != cudaSuccess || (e = up(&ops->goff, goff64)) != cudaSuccess ||
      (e = up(&ops->rec, rec)) != cudaSuccess || (e = up(&ops->biglist, big)) != cudaSuccess) {
    s->fused_ops = ops;
    fused_release(s);
    return e;
  }
  s->fused_ops = ops;
  return cudaSuccess;
}

void fused_release(bdg_sw2d *s) {
  if (!s->fused_ops) return;
  FusedOps *ops = static_cast<FusedOps *>(s->fused_ops);
  cudaFree(ops->Afrag); cudaFree(ops->lmap); cudaFree(ops->gflat); cudaFree(ops->goff); cudaFree(ops->rec); cudaFree(ops->biglist);
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
  args.lmap = ops->lmap; args.gflat = ops->gflat; args.goff = ops->goff; args.rec = ops->rec; args.biglist = ops->biglist;
  args.G = ops->Gmax; args.list_first = 0; args.use_list = 0;
  args.k0 = k0; args.k1 = k1; args.tile0 = 0; args.ntiles = 0;
  args.write_rhs = write_rhs; args.slot = slot; args.a = a; args.b = b;
#define X(np, nfp) \
  if (ops->Np == np && ops->Nfp == nfp) return launch_t<np, nfp>(s, args, ops);
  BDG_FUSED_ORDERS(X)
#undef X
  return cudaErrorNotSupported;
}
