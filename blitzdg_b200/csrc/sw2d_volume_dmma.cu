// Volume + LIFT + stage-update kernel of the generic path on the fp64 tensor pipe (DMMA), for any element shape
// and order with Np <= 64: the second launch of a generic stage (after k_face wrote Fscale*dFlux).
//   reference: volume term src/sw2d/main.cpp:434-442 [sw2d-simple :332-339], surface term :444-458 [:341-355],
//   sources :460-483, stage update include/LSERK4.hpp:16-27 + src/advec1d/main.cpp:98,101.
//
// At N = 8 triangles (Np = 45) or N = 6 quads (Np = 49) the contraction is 52-61 kflop per element and stage against
// 4.5-7 KB of compulsory traffic: the path is bound by the fp64 pipe, not by HBM, which is the
// case north_star reserves for tensor cores.  A warp owns 8 elements (the n = 8 columns of mma.m8n8k4.f64); the
// constant operator A = [-Dr | -Ds | LIFT], padded to 8-row / 4-column blocks, sits in shared memory in A-fragment
// order (one conflict-free 8-byte load per lane and DMMA); B fragments are built in registers straight from the
// state (lane = 4*column + t holds node 4*kb + t of its element: every (element, node) flux is evaluated exactly once).
//   AFFINE (straight-sided triangles): chain-rule factors are element constants and are folded in before the
//     contraction, RHS_i = A * [rx F_i + ry G_i ; sx F_i + sy G_i ; Fscale dF_i]: 3 accumulator sets.
//   otherwise (quads, per-node metrics): Dr F, Ds F, Dr G, Ds G and LIFT*flux are accumulated separately, one field
//     at a time, and combined with the nodal rx, sx, ry, sy in the epilogue (reference operation order).
// The RHS filter (a15) is not handled here: launch_generic_volume keeps the scalar kernel for it.
#include <cmath>
#include <cstdlib>
#include <vector>

#include "sw2d_common.cuh"

namespace {

constexpr int kVW = 4;  // warps per CTA

__device__ __forceinline__ void dmma884(double &d0, double &d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
// 1/x: MUFU seed + two Newton steps (h > 0)
__device__ __forceinline__ double frcp(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  double e = fma(-x, r, 1.0);
  r = fma(r, e, r);
  e = fma(-x, r, 1.0);
  return fma(r, e, r);
}

struct VolArgs {
  const double *Afrag;  // [MB][KB][32]
  int NPB, NFB, KB;     // k-blocks of Dr (= of Ds), of LIFT, total
  double *q, *res;
  const double *sflux;
  double *out;
  int mode;             // 0: write RHS to out, 1: LSERK4 stage update, 2/3: SSP-RK2 stages (stepper_value, sw2d_common.cuh)
  double a, b;
  double *qdst;         // where the new state goes: q itself in mode 1, the other state buffer in modes 2, 3
  int ngroups;          // ceil(K / 8)
  const double *Dq;     // quads with a tensor-product basis: the 1-D differentiation matrix [nq][nq], else NULL
  int nq;
};

struct Flux {
  double hu, hv, F2, G2, G3;
};
struct State {
  double h, hu, hv;
};
__device__ __forceinline__ State load_state(const Sw2dView &v, const double *q, long long m) {
  return State{q[m], q[v.stride + m], q[2 * v.stride + m]};
}
__device__ __forceinline__ Flux nodal_flux(const Sw2dView &v, const State &s) {
  const double r = frcp(s.h), u = s.hu * r, w = s.hv * r, gh2 = (0.5 * v.g * s.h) * s.h;
  return Flux{s.hu, s.hv, fma(s.hu, u, gh2), s.hu * w, fma(s.hv, w, gh2)};
}

// RHS_i at node m of the C fragment -> sources (full mode) -> out / stage update.  q is NOT written here when
// `defer_q` (the non-affine path still needs the old state for the next field).
__device__ __forceinline__ void finish_node(const Sw2dView &v, const VolArgs &p, long long m, int i, double rhs, double dt, bool defer_q) {
  if (v.full && i > 0) {  // A11 sources (main.cpp:460-483)
    const double h = p.q[m], hu = p.q[v.stride + m], hv = p.q[2 * v.stride + m];
    const double r = frcp(h), u = hu * r, w = hv * r, nu = sqrt(u * u + w * w);
    if (i == 1) rhs = rhs + v.g * h * v.Hx[m] + (-v.CD * u * nu) + v.fcor * hv;
    else rhs = rhs + v.g * h * v.Hy[m] + (-v.CD * w * nu) + (-v.fcor * hu);
  }
  if (p.mode == 0) {
    p.out[(long long)i * v.Np * v.K + m] = rhs;
  } else {
    double s = dt * rhs;
    if (p.a != 0.0) s = fma(p.a, p.res[i * v.stride + m], s);
    p.res[i * v.stride + m] = s;
    if (!defer_q) p.q[i * v.stride + m] = fma(p.b, s, p.q[i * v.stride + m]);
  }
}

// Epilogue in the C-fragment layout (row n = 8*rb + lane/4, columns = elements kbase + 2*(lane%4) + {0,1}):
// acc = RHS_i without sources -> sources (full mode, A11) -> out (mode 0) / LSERK4 stage update in place (mode 1).
// RK2 instances (modes 2, 3) also batch the destination's old values and the sponge; they are separate kernel
// instantiations so that the LSERK4 instances keep their register budget.
template <int MB, bool RK2>
__device__ __forceinline__ void stage_epilogue(const Sw2dView &v, const VolArgs &p, double (&acc)[3][MB][2], long long kbase, int lane, double dt) {
  const int gr = lane >> 2, t = lane & 3, Np = v.Np;
  const double *q = p.q;
#pragma unroll
  for (int rb = 0; rb < MB; ++rb) {
    const int n = 8 * rb + gr;
    const bool ok0 = n < Np && kbase + 2 * t < v.K, ok1 = n < Np && kbase + 2 * t + 1 < v.K;
    const long long m0 = (kbase + 2 * t) * Np + n, m1 = m0 + Np;
    // all loads of the row block first (the stores below may alias them as far as the compiler knows)
    double qo[2][3], ro[2][3], hx[2] = {0.0, 0.0}, hy[2] = {0.0, 0.0};
    double qd[RK2 ? 2 : 1][3], sg[2] = {0.0, 0.0};
#pragma unroll
    for (int cc = 0; cc < 2; ++cc) {
      const bool ok = cc ? ok1 : ok0;
      const long long m = cc ? m1 : m0;
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        qo[cc][i] = (ok && (p.mode >= 1 || v.full)) ? q[i * v.stride + m] : 1.0;
        ro[cc][i] = (!RK2 && ok && p.mode == 1 && p.a != 0.0) ? p.res[i * v.stride + m] : 0.0;
        if (RK2) qd[cc][i] = (ok && p.mode == 3) ? p.qdst[i * v.stride + m] : 0.0;
      }
      if (ok && v.full) { hx[cc] = v.Hx[m]; hy[cc] = v.Hy[m]; }
      if (RK2 && ok && v.sponge) sg[cc] = __ldg(v.sponge + m);
    }
#pragma unroll
    for (int cc = 0; cc < 2; ++cc) {
      const bool ok = cc ? ok1 : ok0;
      const long long m = cc ? m1 : m0;
      double rr[3] = {acc[0][rb][cc], acc[1][rb][cc], acc[2][rb][cc]};
      if (v.full) {  // A11 sources (main.cpp:460-483) from the old state
        const double h = qo[cc][0], hu = qo[cc][1], hv = qo[cc][2];
        const double r = frcp(h), u = hu * r, ww = hv * r, nu = sqrt(u * u + ww * ww);
        rr[1] = rr[1] + v.g * h * hx[cc] + (-v.CD * u * nu) + v.fcor * hv;
        rr[2] = rr[2] + v.g * h * hy[cc] + (-v.CD * ww * nu) + (-v.fcor * hu);
      }
      if (ok) {
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          if (p.mode == 0) {
            p.out[(long long)i * Np * v.K + m] = rr[i];
          } else {
            const double sres = fma(p.a, ro[cc][i], dt * rr[i]);  // a = 0 in modes 2, 3
            if (!RK2) {
              p.res[i * v.stride + m] = sres;
              p.q[i * v.stride + m] = fma(p.b, sres, qo[cc][i]);
            } else {  // SSP-RK2 forms (stepper_value, sw2d_common.cuh)
              p.qdst[i * v.stride + m] = stepper_value(p.mode, p.b, sres, qo[cc][i], qd[RK2 ? cc : 0][i], i, sg[cc]);
            }
          }
        }
      }
    }
  }
}

#ifndef BDG_VOL_MINB_SMALL
#define BDG_VOL_MINB_SMALL 5  // N <= 4 affine: 96 registers, 5 CTAs/SM measured +1 % over 3 (tools/vol_occupancy_sweep.sh)
#endif
template <int MB, bool AFFINE, bool RK2>
__global__ void __launch_bounds__(kVW * 32, (MB <= 2 && AFFINE) ? BDG_VOL_MINB_SMALL : 3) k_volume_dmma(Sw2dView v, VolArgs p) {
  extern __shared__ __align__(16) double sA[];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int c = lane >> 2, t = lane & 3, gr = lane >> 2;
  const int Np = v.Np, Nfn = v.Nfn, NPB = p.NPB, NFB = p.NFB, KB = p.KB;
  const long long nF = (long long)Nfn * v.K;
  for (int x = threadIdx.x; x < MB * KB * 32; x += blockDim.x) sA[x] = __ldg(p.Afrag + x);
  __syncthreads();
  const double dt = p.mode ? v.scal[S_DT] : 0.0;
  const double *q = p.q;

  for (int g = blockIdx.x * kVW + w; g < p.ngroups; g += gridDim.x * kVW) {
    const long long kbase = 8LL * g;
    const long long kB = min(kbase + c, (long long)v.K - 1);  // my B-fragment column's element (ragged end: duplicate)
    if (AFFINE) {
      const double rx = v.rx[kB * Np], sx = v.sx[kB * Np], ry = v.ry[kB * Np], sy = v.sy[kB * Np];
      double acc[3][MB][2];
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int rb = 0; rb < MB; ++rb) acc[i][rb][0] = acc[i][rb][1] = 0.0;
      // the state loads run two k-blocks ahead of the DMMAs that consume them (a k-block is MB*6 DMMAs ~ 600 cycles)
      State st0 = load_state(v, q, kB * Np + min(t, Np - 1)), st1 = load_state(v, q, kB * Np + min(4 + t, Np - 1));
#pragma unroll 2
      for (int kb = 0; kb < NPB; ++kb) {
        const State st2 = load_state(v, q, kB * Np + min(4 * (kb + 2) + t, Np - 1));
        const Flux f = nodal_flux(v, st0);  // padding rows of B meet zero columns of A
        st0 = st1;
        st1 = st2;
        double ba[3], bb[3];
        ba[0] = fma(rx, f.hu, ry * f.hv); bb[0] = fma(sx, f.hu, sy * f.hv);
        ba[1] = fma(rx, f.F2, ry * f.G2); bb[1] = fma(sx, f.F2, sy * f.G2);
        ba[2] = fma(rx, f.G2, ry * f.G3); bb[2] = fma(sx, f.G2, sy * f.G3);
#pragma unroll
        for (int rb = 0; rb < MB; ++rb) {
          const double a1 = sA[(rb * KB + kb) * 32 + lane], a2 = sA[(rb * KB + NPB + kb) * 32 + lane];
#pragma unroll
          for (int i = 0; i < 3; ++i) dmma884(acc[i][rb][0], acc[i][rb][1], a1, ba[i]);
#pragma unroll
          for (int i = 0; i < 3; ++i) dmma884(acc[i][rb][0], acc[i][rb][1], a2, bb[i]);
        }
      }
      long long i0 = kB * Nfn + min(t, Nfn - 1);
      double n0 = p.sflux[i0], n1 = p.sflux[nF + i0], n2 = p.sflux[2 * nF + i0];
      for (int kb = 0; kb < NFB; ++kb) {
        const double s0 = n0, s1 = n1, s2 = n2;
        i0 = kB * Nfn + min(4 * (kb + 1) + t, Nfn - 1);
        n0 = p.sflux[i0]; n1 = p.sflux[nF + i0]; n2 = p.sflux[2 * nF + i0];
#pragma unroll
        for (int rb = 0; rb < MB; ++rb) {
          const double a3 = sA[(rb * KB + 2 * NPB + kb) * 32 + lane];
          dmma884(acc[0][rb][0], acc[0][rb][1], a3, s0);
          dmma884(acc[1][rb][0], acc[1][rb][1], a3, s1);
          dmma884(acc[2][rb][0], acc[2][rb][1], a3, s2);
        }
      }
      __syncwarp();  // every lane's state loads of these 8 elements precede the in-place update
      stage_epilogue<MB, RK2>(v, p, acc, kbase, lane, dt);
    } else {
      // per-node metrics: one field at a time; C = {Dr X, Ds X, Dr Y, Ds Y, LIFT s} with X,Y = (F_i, G_i)
#pragma unroll 1
      for (int i = 0; i < 3; ++i) {
        double aRX[MB][2], aSX[MB][2], aRY[MB][2], aSY[MB][2], aL[MB][2];
#pragma unroll
        for (int rb = 0; rb < MB; ++rb)
          aRX[rb][0] = aRX[rb][1] = aSX[rb][0] = aSX[rb][1] = aRY[rb][0] = aRY[rb][1] = aSY[rb][0] = aSY[rb][1] = aL[rb][0] = aL[rb][1] = 0.0;
        State st0 = load_state(v, q, kB * Np + min(t, Np - 1)), st1 = load_state(v, q, kB * Np + min(4 + t, Np - 1));
#pragma unroll 2
        for (int kb = 0; kb < NPB; ++kb) {
          const State st2 = load_state(v, q, kB * Np + min(4 * (kb + 2) + t, Np - 1));
          const Flux f = nodal_flux(v, st0);
          st0 = st1;
          st1 = st2;
          const double X = i == 0 ? f.hu : (i == 1 ? f.F2 : f.G2), Y = i == 0 ? f.hv : (i == 1 ? f.G2 : f.G3);
#pragma unroll
          for (int rb = 0; rb < MB; ++rb) {
            const double a1 = sA[(rb * KB + kb) * 32 + lane], a2 = sA[(rb * KB + NPB + kb) * 32 + lane];
            dmma884(aRX[rb][0], aRX[rb][1], a1, X);
            dmma884(aSX[rb][0], aSX[rb][1], a2, X);
            dmma884(aRY[rb][0], aRY[rb][1], a1, Y);
            dmma884(aSY[rb][0], aSY[rb][1], a2, Y);
          }
        }
        for (int kb = 0; kb < NFB; ++kb) {
          const double s = p.sflux[i * nF + kB * Nfn + min(4 * kb + t, Nfn - 1)];
#pragma unroll
          for (int rb = 0; rb < MB; ++rb) dmma884(aL[rb][0], aL[rb][1], sA[(rb * KB + 2 * NPB + kb) * 32 + lane], s);
        }
#pragma unroll
        for (int rb = 0; rb < MB; ++rb) {
          const int n = 8 * rb + gr;
#pragma unroll
          for (int cc = 0; cc < 2; ++cc) {
            const long long k = kbase + 2 * t + cc;
            if (n < Np && k < v.K) {
              const long long m = k * Np + n;
              // A holds -Dr, -Ds: RHS_i = -(rx Dr F + sx Ds F) - (ry Dr G + sy Ds G) + LIFT s   (main.cpp:434-458)
              double rhs = fma(v.rx[m], aRX[rb][cc], v.sx[m] * aSX[rb][cc]);
              rhs = rhs + fma(v.ry[m], aRY[rb][cc], v.sy[m] * aSY[rb][cc]);
              rhs = rhs + aL[rb][cc];
              finish_node(v, p, m, i, rhs, dt, true);
            }
          }
        }
      }
      if (p.mode >= 1) {  // all three fields have read the old state: q_i += b * res_i  (advec1d/main.cpp:101), or the SSP-RK2 forms
        __syncwarp();
#pragma unroll
        for (int rb = 0; rb < MB; ++rb) {
          const int n = 8 * rb + gr;
#pragma unroll
          for (int cc = 0; cc < 2; ++cc) {
            const long long k = kbase + 2 * t + cc;
            if (n < Np && k < v.K) {
              const long long m = k * Np + n;
#pragma unroll
              for (int i = 0; i < 3; ++i) {
                const double qold = p.mode == 3 ? p.qdst[i * v.stride + m] : 0.0;
                const double sig = (p.mode >= 2 && v.sponge) ? v.sponge[m] : 0.0;
                p.qdst[i * v.stride + m] = stepper_value(p.mode, p.b, p.res[i * v.stride + m], p.q[i * v.stride + m], qold, i, sig);
              }
            }
          }
        }
      }
    }
  }
}

// ---- quads with a tensor-product nodal basis (QuadNodesProvisioner: node = i1*n + i2, Dr = D (x) I, Ds = I (x) D,
// reference src/QuadNodesProvisioner.cpp:204-260): sum factorisation.  Per element and flux component U (an n x n
// matrix) Dr U = D U and Ds U = U D^T are two 8x8x8 products instead of two Np x Np ones: 24 DMMAs per element for
// the volume term instead of ~137.  Both products are laid out so that their C fragments coincide (row i1, columns
// i2 = 2t, 2t+1 of ONE element): Dr U takes -D as the A operand and U as B, Ds U takes U as A and -D^T as B -- the
// same register serves both.  The warp stages an element's nodal fluxes in shared memory (every flux evaluated once),
// combines the four products with the nodal rx, sx, ry, sy, parks the volume RHS of its 8 elements in shared memory
// and then runs the dense LIFT contraction across the 8 elements (columns = elements) from there.
template <int MB, bool RK2>
__global__ void __launch_bounds__(kVW * 32, 3) k_volume_quad_sf(Sw2dView v, VolArgs p) {
  extern __shared__ __align__(16) double sm[];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int c = lane >> 2, t = lane & 3, gr = lane >> 2;
  const int Np = v.Np, Nfn = v.Nfn, n = p.nq, NPB = p.NPB, NFB = p.NFB, KB = p.KB;
  constexpr int VS = MB * 8;  // padded nodes per element in the parked volume RHS
  const long long nF = (long long)Nfn * v.K;
  double *sL = sm;                                                      // LIFT fragments [MB][NFB][32]
  double *sF = sL + MB * NFB * 32 + w * (5 * 64 + 3 * 8 * VS);          // my warp: flux planes hu,hv,F2,G2,G3 [5][8x8]
  double *sV = sF + 5 * 64;                                             // my warp: volume RHS [3][8 elements][VS]
  for (int x = threadIdx.x; x < MB * NFB * 32; x += blockDim.x) {
    const int rb = x / (NFB * 32), r = x - rb * NFB * 32;
    sL[x] = __ldg(p.Afrag + (rb * KB + 2 * NPB) * 32 + r);
  }
  for (int x = lane; x < 3 * 8 * VS; x += 32) sV[x] = 0.0;
  double aD[2];  // -D as A fragment (row gr, column 4*kb+t) == -D^T as B fragment (k = 4*kb+t, column c = gr)
#pragma unroll
  for (int kb = 0; kb < 2; ++kb) aD[kb] = (gr < n && 4 * kb + t < n) ? -__ldg(p.Dq + gr * n + 4 * kb + t) : 0.0;
  __syncthreads();
  const double dt = p.mode ? v.scal[S_DT] : 0.0;
  const double *q = p.q;

  for (int g = blockIdx.x * kVW + w; g < p.ngroups; g += gridDim.x * kVW) {
    const long long kbase = 8LL * g;
    for (int j = 0; j < 8; ++j) {
      const long long e = kbase + j;
      if (e >= v.K) break;  // warp-uniform (ragged last group: the epilogue guards its stores)
      // phase A: nodal fluxes of element e into the 8x8 planes (padding = 0)
#pragma unroll
      for (int rep = 0; rep < 2; ++rep) {
        const int idx = lane + 32 * rep, i1 = idx >> 3, i2 = idx & 7;
        Flux f{0.0, 0.0, 0.0, 0.0, 0.0};
        if (i1 < n && i2 < n) f = nodal_flux(v, load_state(v, q, e * Np + i1 * n + i2));
        sF[idx] = f.hu; sF[64 + idx] = f.hv; sF[128 + idx] = f.F2; sF[192 + idx] = f.G2; sF[256 + idx] = f.G3;
      }
      // nodal metrics at my two C-fragment nodes (i1 = gr, i2 = 2t + cc)
      double rxv[2], sxv[2], ryv[2], syv[2];
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
        const bool ok = gr < n && 2 * t + cc < n;
        const long long m = e * Np + (ok ? gr * n + 2 * t + cc : 0);
        rxv[cc] = v.rx[m]; sxv[cc] = v.sx[m]; ryv[cc] = v.ry[m]; syv[cc] = v.sy[m];
      }
      __syncwarp();
      // phase B: per field the four 8x8x8 products, combined with the metrics (main.cpp:434-442)
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        const double *X = sF + (i == 0 ? 0 : (i == 1 ? 128 : 192)), *Y = sF + (i == 0 ? 64 : (i == 1 ? 192 : 256));
        double drx[2] = {0.0, 0.0}, dsx[2] = {0.0, 0.0}, dry[2] = {0.0, 0.0}, dsy[2] = {0.0, 0.0};
#pragma unroll
        for (int kb = 0; kb < 2; ++kb) {
          dmma884(drx[0], drx[1], aD[kb], X[(4 * kb + t) * 8 + c]);
          dmma884(dry[0], dry[1], aD[kb], Y[(4 * kb + t) * 8 + c]);
          dmma884(dsx[0], dsx[1], X[gr * 8 + 4 * kb + t], aD[kb]);
          dmma884(dsy[0], dsy[1], Y[gr * 8 + 4 * kb + t], aD[kb]);
        }
#pragma unroll
        for (int cc = 0; cc < 2; ++cc)
          if (gr < n && 2 * t + cc < n) {
            double r = fma(rxv[cc], drx[cc], sxv[cc] * dsx[cc]);
            r = r + fma(ryv[cc], dry[cc], syv[cc] * dsy[cc]);
            sV[(i * 8 + j) * VS + gr * n + 2 * t + cc] = r;
          }
      }
      __syncwarp();  // the planes are rewritten for the next element
    }
    // LIFT across the 8 elements (columns = elements), accumulators start from the parked volume RHS
    double acc[3][MB][2];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int rb = 0; rb < MB; ++rb) {
        acc[i][rb][0] = sV[(i * 8 + 2 * t) * VS + 8 * rb + gr];
        acc[i][rb][1] = sV[(i * 8 + 2 * t + 1) * VS + 8 * rb + gr];
      }
    const long long kB = min(kbase + c, (long long)v.K - 1);
    long long i0 = kB * Nfn + min(t, Nfn - 1);
    double n0 = p.sflux[i0], n1 = p.sflux[nF + i0], n2 = p.sflux[2 * nF + i0];
    for (int kb = 0; kb < NFB; ++kb) {
      const double s0 = n0, s1 = n1, s2 = n2;
      i0 = kB * Nfn + min(4 * (kb + 1) + t, Nfn - 1);
      n0 = p.sflux[i0]; n1 = p.sflux[nF + i0]; n2 = p.sflux[2 * nF + i0];
#pragma unroll
      for (int rb = 0; rb < MB; ++rb) {
        const double a3 = sL[(rb * NFB + kb) * 32 + lane];
        dmma884(acc[0][rb][0], acc[0][rb][1], a3, s0);
        dmma884(acc[1][rb][0], acc[1][rb][1], a3, s1);
        dmma884(acc[2][rb][0], acc[2][rb][1], a3, s2);
      }
    }
    __syncwarp();
    stage_epilogue<MB, RK2>(v, p, acc, kbase, lane, dt);
    __syncwarp();
  }
}

template <int MB>
cudaError_t launch_mb(bdg_sw2d *s, const VolArgs &a, size_t smem) {
  static int ctas[16][4] = {};
  static size_t smem_set[16][4] = {};
  const int dev = s->device & 15, aff = (s->is_affine ? 1 : 0) + (a.mode >= 2 ? 2 : 0);
  auto kern = a.mode >= 2 ? (s->is_affine ? k_volume_dmma<MB, true, true> : k_volume_dmma<MB, false, true>)
                          : (s->is_affine ? k_volume_dmma<MB, true, false> : k_volume_dmma<MB, false, false>);
  if (smem != smem_set[dev][aff]) {  // one instance serves every order with this row-block count: the operator size varies
    cudaError_t e;
    if ((e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas[dev][aff], kern, kVW * 32, smem)) != cudaSuccess) return e;
    if (ctas[dev][aff] < 1) return cudaErrorLaunchOutOfResources;
    smem_set[dev][aff] = smem;
  }
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, s->device);
  int grid = sms * ctas[dev][aff];
  const int need = (a.ngroups + kVW - 1) / kVW;
  if (grid > need) grid = need;
  kern<<<grid, kVW * 32, smem, s->stream>>>(s->v, a);
  s->launches++;
  return cudaGetLastError();
}

template <int MB>
cudaError_t launch_quad(bdg_sw2d *s, const VolArgs &a) {
  static int ctas[16][2] = {};
  static size_t smem_set[16][2] = {};
  const int dev = s->device & 15, rk = a.mode >= 2 ? 1 : 0;
  auto kern = rk ? k_volume_quad_sf<MB, true> : k_volume_quad_sf<MB, false>;
  const size_t smem = sizeof(double) * ((size_t)MB * a.NFB * 32 + (size_t)kVW * (5 * 64 + 3 * 8 * MB * 8));
  if (smem != smem_set[dev][rk]) {
    cudaError_t e;
    if ((e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas[dev][rk], kern, kVW * 32, smem)) != cudaSuccess) return e;
    if (ctas[dev][rk] < 1) return cudaErrorLaunchOutOfResources;
    smem_set[dev][rk] = smem;
  }
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, s->device);
  int grid = sms * ctas[dev][rk];
  const int need = (a.ngroups + kVW - 1) / kVW;
  if (grid > need) grid = need;
  kern<<<grid, kVW * 32, smem, s->stream>>>(s->v, a);
  s->launches++;
  return cudaGetLastError();
}

}  // namespace

// Operator fragments of the generic DMMA volume kernel: A = [-Dr | -Ds | LIFT] in m8n8k4 A-fragment order
// [row block][k block][lane] (row = lane/4, column = lane%4).  Leaves s->d_vfrag NULL when Np > 64.
cudaError_t volume_dmma_prepare(bdg_sw2d *s, const double *Dr, const double *Ds, const double *Lift) {
  const int Np = s->v.Np, Nfn = s->v.Nfn;
  s->d_vfrag = nullptr;
  if (Np > 64) return cudaSuccess;
  const int MB = (Np + 7) / 8, NPB = (Np + 3) / 4, NFB = (Nfn + 3) / 4, KB = 2 * NPB + NFB;
  std::vector<double> frag((size_t)MB * KB * 32, 0.0);
  for (int rb = 0; rb < MB; ++rb)
    for (int kb = 0; kb < KB; ++kb)
      for (int lane = 0; lane < 32; ++lane) {
        const int row = 8 * rb + lane / 4;
        double val = 0.0;
        if (row < Np) {
          if (kb < NPB) {
            const int j = 4 * kb + lane % 4;
            if (j < Np) val = -Dr[row * Np + j];
          } else if (kb < 2 * NPB) {
            const int j = 4 * (kb - NPB) + lane % 4;
            if (j < Np) val = -Ds[row * Np + j];
          } else {
            const int j = 4 * (kb - 2 * NPB) + lane % 4;
            if (j < Nfn) val = Lift[row * Nfn + j];
          }
        }
        frag[((size_t)rb * KB + kb) * 32 + lane] = val;
      }
  cudaError_t e = cudaMalloc((void **)&s->d_vfrag, sizeof(double) * frag.size());
  if (e != cudaSuccess) return e;
  if ((e = cudaMemcpy(s->d_vfrag, frag.data(), sizeof(double) * frag.size(), cudaMemcpyHostToDevice)) != cudaSuccess) return e;
  // quads: is the basis a tensor product with node = i1*n + i2, Dr = D (x) I, Ds = I (x) D ?  (sum-factorised kernel)
  s->d_Dq = nullptr;
  s->quad_n = 0;
  if (s->v.Nf == 4) {
    int n = 1;
    while (n * n < Np) ++n;
    if (n * n == Np && n >= 2 && n <= 8 && s->v.Nfp == n) {
      std::vector<double> D((size_t)n * n);
      double scale = 0.0, err = 0.0;
      for (int a = 0; a < n; ++a)
        for (int b = 0; b < n; ++b) { D[a * n + b] = Dr[(a * n) * Np + b * n]; scale = fmax(scale, fabs(D[a * n + b])); }
      for (int i1 = 0; i1 < n; ++i1)
        for (int i2 = 0; i2 < n; ++i2)
          for (int j1 = 0; j1 < n; ++j1)
            for (int j2 = 0; j2 < n; ++j2) {
              const int r = i1 * n + i2, cidx = j1 * n + j2;
              err = fmax(err, fabs(Dr[r * Np + cidx] - (i2 == j2 ? D[i1 * n + j1] : 0.0)));
              err = fmax(err, fabs(Ds[r * Np + cidx] - (i1 == j1 ? D[i2 * n + j2] : 0.0)));
            }
      if (err <= 1e-12 * scale) {
        if ((e = cudaMalloc((void **)&s->d_Dq, sizeof(double) * D.size())) != cudaSuccess) return e;
        if ((e = cudaMemcpy(s->d_Dq, D.data(), sizeof(double) * D.size(), cudaMemcpyHostToDevice)) != cudaSuccess) return e;
        s->quad_n = n;
      }
    }
  }
  return cudaSuccess;
}

int volume_dmma_supported(const bdg_sw2d *s) { return s->d_vfrag != nullptr && !s->v.use_filter; }

cudaError_t launch_volume_dmma(bdg_sw2d *s, double *q, double *res, double *out, int mode, double a, double b, double *qdst) {
  const int Np = s->v.Np, Nfn = s->v.Nfn;
  VolArgs p;
  p.Afrag = s->d_vfrag;
  p.NPB = (Np + 3) / 4; p.NFB = (Nfn + 3) / 4; p.KB = 2 * p.NPB + p.NFB;
  p.q = q; p.res = res; p.sflux = s->sflux; p.out = out; p.mode = mode; p.a = mode >= 2 ? 0.0 : a; p.b = b; p.qdst = mode >= 2 ? qdst : q;
  p.ngroups = (s->v.K + 7) / 8;
  p.Dq = s->d_Dq; p.nq = s->quad_n;
  const int MB = (Np + 7) / 8;
  static const bool quad_dense = getenv("BDG_QUAD_DENSE") != nullptr;  // study switch: dense Dr/Ds contraction for quads
  if (s->d_Dq && !quad_dense) {
    switch (MB) {
      case 1: return launch_quad<1>(s, p);
      case 2: return launch_quad<2>(s, p);
      case 4: return launch_quad<4>(s, p);
      case 5: return launch_quad<5>(s, p);
      case 7: return launch_quad<7>(s, p);
      case 8: return launch_quad<8>(s, p);
    }
  }
  const size_t smem = sizeof(double) * MB * p.KB * 32;
  switch (MB) {
    case 1: return launch_mb<1>(s, p, smem);
    case 2: return launch_mb<2>(s, p, smem);
    case 3: return launch_mb<3>(s, p, smem);
    case 4: return launch_mb<4>(s, p, smem);
    case 5: return launch_mb<5>(s, p, smem);
    case 6: return launch_mb<6>(s, p, smem);
    case 7: return launch_mb<7>(s, p, smem);
    case 8: return launch_mb<8>(s, p, smem);
  }
  return cudaErrorNotSupported;
}
