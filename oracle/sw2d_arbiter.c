/* TEST INFRASTRUCTURE -- extended-precision ARBITER for the sw2d RHS.
 *
 * Same mathematics as sw2d::computeRHS (reference src/sw2d/main.cpp:279-484; per-face LF variant
 * src/sw2d-simple/main.cpp:289-319), evaluated element by element in `long double` (x87, 64-bit mantissa,
 * eps = 5.4e-20) from the SAME double-precision inputs.  It is not the oracle (oracle/sw2d_oracle.c restates the
 * reference's fp64 operation order and stays the parity definition): it is the yardstick that says how far the
 * reference's own fp64 arithmetic and the CUDA kernels each are from the exact value of the discrete operator on a
 * given state -- the volume term differentiates g h^2/2 with metric factors ~ 1/h_cell, so on the BASELINE-size
 * meshes a few ulps of cancellation error exceed 1e-12 of the result for ANY fp64 summation order.
 * Its own error is ~2000x below fp64's (observed agreement with a __float128 run of the same code: < 3e-18 relative).
 *
 * Only tests/ and bench.py's parity leg call it.  An element range [k0,k1) can be evaluated on its own (the inputs
 * are the full arrays), so callers thread over ranges and sample the 16 M-element mesh.
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/blitzdg_b200.h"

#ifdef ARB_QUAD
#include <quadmath.h>
typedef __float128 real;
#define SQRT sqrtq
#define COS cosq
#define TANH tanhq
#else
typedef long double real;
#define SQRT sqrtl
#define COS cosl
#define TANH tanhl
#endif

static const real ARB_PI = 3.14159265358979323846264338327950288L;

static inline real rmax(real a, real b) { return a > b ? a : b; }

typedef struct {
  real h, hu, hv, H;
} trace;

static inline real wave_speed(real g, real h, real hu, real hv) {
  const real u = hu / h, v = hv / h;
  return SQRT(u * u + v * v) + SQRT(g * h);
}

/* traces of face node i = k*Nfn + sl after BC and (full mode) star states; returns per-node max speed */
static real face_node(const bdg_dgcontext2d *dg, const bdg_sw2d_params *prm, const uint8_t *bctag, const double *H,
                      const double *h, const double *hu, const double *hv, double t, int n_ghost, const double *ghost,
                      long long k, int sl, trace *M, trace *P, real *nx, real *ny) {
  const int Np = dg->Np, Nfn = dg->Nfaces * dg->Nfp;
  const long long K = dg->K, nT = (long long)Np * K;
  const int full = prm->sources != 0;
  const long long i = k * Nfn + sl;
  const long long mM = dg->vmapM[i], mP = dg->vmapP[i];
  const long long nM = mM % Np, kM = mM / Np;
  M->h = h[nM * K + kM]; M->hu = hu[nM * K + kM]; M->hv = hv[nM * K + kM];
  M->H = full ? (real)H[nM * K + kM] : 0;
  if (mP >= nT) {
    const long long gi = mP - nT;
    P->h = ghost[gi]; P->hu = ghost[n_ghost + gi]; P->hv = ghost[2LL * n_ghost + gi];
    P->H = full ? (real)ghost[3LL * n_ghost + gi] : 0;
  } else {
    const long long nP = mP % Np, kP = mP / Np;
    P->h = h[nP * K + kP]; P->hu = hu[nP * K + kP]; P->hv = hv[nP * K + kP];
    P->H = full ? (real)H[nP * K + kP] : 0;
  }
  *nx = dg->nx[(long long)sl * K + k];
  *ny = dg->ny[(long long)sl * K + k];
  const int tag = bctag ? bctag[i] : 0;
  if (tag == 3) { /* wall, main.cpp:340-345 */
    const real d = M->hu * *nx + M->hv * *ny;
    if (full) P->h = M->h;
    P->hu = M->hu - 2 * *nx * d;
    P->hv = M->hv - 2 * *ny * d;
  } else if (tag == 2 && full) { /* open boundary with tide, main.cpp:348-353 */
    const real T = prm->tide_period, om = 2 * ARB_PI / T;
    P->hu = M->hu; P->hv = M->hv;
    P->h = M->H + (real)prm->tide_amp * COS(om * t) * 0.5L * (TANH((real)prm->tide_ramp * (t - T)) + 1);
  }
  if (full) { /* star states, main.cpp:355-368 */
    const real bM = -M->H, bP = -P->H, bmax = rmax(bP, bM);
    const real hMs = rmax(0, M->h + bM - bmax), hPs = rmax(0, P->h + bP - bmax);
    /* the reference overwrites hM with hM* before rescaling the momenta (:366-368): hu* = hM* (hu / hM*) = hu */
    M->h = hMs; P->h = hPs;
  }
  return rmax(wave_speed(prm->g, M->h, M->hu, M->hv), wave_speed(prm->g, P->h, P->hu, P->hv));
}

/* global LF speed over the face nodes of elements [k0,k1) (main.cpp:397-414); callers take the max over ranges */
double arbiter_sw2d_max_speed(const bdg_dgcontext2d *dg, const bdg_sw2d_params *prm, const uint8_t *bctag, const double *H,
                              const double *h, const double *hu, const double *hv, double t, int n_ghost,
                              const double *ghost, int k0, int k1) {
  const int Nfn = dg->Nfaces * dg->Nfp;
  real mx = 0;
  for (long long k = k0; k < k1; ++k)
    for (int sl = 0; sl < Nfn; ++sl) {
      trace M, P;
      real nx, ny;
      mx = rmax(mx, face_node(dg, prm, bctag, H, h, hu, hv, t, n_ghost, ghost, k, sl, &M, &P, &nx, &ny));
    }
  return (double)mx;
}

/* RHS1..3 [Np][K] row-major: only the columns k0 <= k < k1 are written.  bctag [Nfn*K] (BCTag per face node, 0 =
 * interior; NULL = none).  lam_global: the global LF speed when prm->lf_mode == 1 (from arbiter_sw2d_max_speed). */
int arbiter_sw2d_rhs(const bdg_dgcontext2d *dg, const bdg_sw2d_params *prm, const uint8_t *bctag, const double *H,
                     const double *Hx, const double *Hy, const double *h, const double *hu, const double *hv, double t,
                     int n_ghost, const double *ghost, double lam_global, int k0, int k1, double *RHS1, double *RHS2,
                     double *RHS3) {
  const int Np = dg->Np, Nfp = dg->Nfp, Nf = dg->Nfaces, Nfn = Nf * Nfp;
  const long long K = dg->K;
  const int full = prm->sources != 0;
  const real g = prm->g;
  real *F = (real *)malloc(sizeof(real) * (6 * (size_t)Np + 4 * (size_t)Nfn + 3 * (size_t)Np));
  if (!F) return 1;
  real *q0 = F, *q1 = F + Np, *q2 = q1 + Np, *F2 = q2 + Np, *G2 = F2 + Np, *G3 = G2 + Np;
  real *s0 = G3 + Np, *s1 = s0 + Nfn, *s2 = s1 + Nfn, *lam = s2 + Nfn, *R = lam + Nfn;
  double *out[3] = {RHS1, RHS2, RHS3};
  for (long long k = k0; k < k1; ++k) {
    for (int n = 0; n < Np; ++n) {
      const real hh = h[n * K + k], a = hu[n * K + k], b = hv[n * K + k];
      q0[n] = hh; q1[n] = a; q2[n] = b;
      F2[n] = a * a / hh + 0.5L * g * hh * hh;
      G2[n] = a * b / hh;
      G3[n] = b * b / hh + 0.5L * g * hh * hh;
    }
    /* face nodes: traces, speeds, strong-form flux jump times Fscale (main.cpp:376-421, 451-454) */
    trace M[64], P[64];
    real nx[64], ny[64];
    for (int f = 0; f < Nf; ++f) {
      real fmaxs = 0;
      for (int j = 0; j < Nfp; ++j) {
        const int sl = f * Nfp + j;
        lam[sl] = face_node(dg, prm, bctag, H, h, hu, hv, t, n_ghost, ghost, k, sl, &M[j], &P[j], &nx[j], &ny[j]);
        fmaxs = rmax(fmaxs, lam[sl]);
      }
      for (int j = 0; j < Nfp; ++j) {
        const int sl = f * Nfp + j;
        const real l = prm->lf_mode == 1 ? (real)lam_global : fmaxs; /* sw2d-simple main.cpp:305-312: max over the face */
        const trace *m = &M[j], *p = &P[j];
        const real F2M = m->hu * m->hu / m->h + 0.5L * g * m->h * m->h, F2P = p->hu * p->hu / p->h + 0.5L * g * p->h * p->h;
        const real G2M = m->hu * m->hv / m->h, G2P = p->hu * p->hv / p->h;
        const real G3M = m->hv * m->hv / m->h + 0.5L * g * m->h * m->h, G3P = p->hv * p->hv / p->h + 0.5L * g * p->h * p->h;
        const real fs = dg->Fscale[(long long)sl * K + k];
        s0[sl] = fs * 0.5L * ((m->hu - p->hu) * nx[j] + (m->hv - p->hv) * ny[j] - l * (m->h - p->h));
        s1[sl] = fs * 0.5L * ((F2M - F2P) * nx[j] + (G2M - G2P) * ny[j] - l * (m->hu - p->hu));
        s2[sl] = fs * 0.5L * ((G2M - G2P) * nx[j] + (G3M - G3P) * ny[j] - l * (m->hv - p->hv));
      }
    }
    /* volume + surface terms (main.cpp:434-458), sources (:460-483) */
    const real *Fx[3] = {q1, F2, G2}, *Gy[3] = {q2, G2, G3}, *S[3] = {s0, s1, s2};
    for (int c = 0; c < 3; ++c)
      for (int n = 0; n < Np; ++n) {
        real drF = 0, dsF = 0, drG = 0, dsG = 0, lf = 0;
        for (int j = 0; j < Np; ++j) {
          const real dr = dg->Dr[n * Np + j], ds = dg->Ds[n * Np + j];
          drF += dr * Fx[c][j]; dsF += ds * Fx[c][j];
          drG += dr * Gy[c][j]; dsG += ds * Gy[c][j];
        }
        for (int j = 0; j < Nfn; ++j) lf += (real)dg->Lift[n * Nfn + j] * S[c][j];
        const real rx = dg->rx[n * K + k], sx = dg->sx[n * K + k], ry = dg->ry[n * K + k], sy = dg->sy[n * K + k];
        real r = -(rx * drF + sx * dsF) - (ry * drG + sy * dsG) + lf;
        if (full && c > 0) {
          const real u = q1[n] / q0[n], v = q2[n] / q0[n], nu = SQRT(u * u + v * v);
          if (c == 1) r += g * q0[n] * (real)Hx[n * K + k] - (real)prm->CD * u * nu + (real)prm->f * q2[n];
          if (c == 2) r += g * q0[n] * (real)Hy[n * K + k] - (real)prm->CD * v * nu - (real)prm->f * q1[n];
        }
        R[c * Np + n] = r;
      }
    for (int c = 0; c < 3; ++c)
      for (int n = 0; n < Np; ++n) {
        real r = R[c * Np + n];
        if (prm->filter && dg->Filt) { /* sw2d-simple main.cpp:134-136 */
          r = 0;
          for (int j = 0; j < Np; ++j) r += (real)dg->Filt[n * Np + j] * R[c * Np + j];
        }
        out[c][n * K + k] = (double)r;
      }
  }
  free(F);
  return 0;
}

int arbiter_real_bytes(void) { return (int)sizeof(real); }
int arbiter_mantissa_bits(void) {
#ifdef ARB_QUAD
  return 113;
#else
  return LDBL_MANT_DIG;
#endif
}
