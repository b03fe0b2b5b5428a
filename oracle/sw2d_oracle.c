/* TEST INFRASTRUCTURE -- CPU oracle for the sw2d nodal-DG RHS + LSERK4 hot path.
 *
 * A plain-C restatement of the reference's arithmetic, operation order and sweep structure
 * (one pass + one temporary per blitz expression, row-major [Np][K] storage, sequential
 * non-FMA dot products).  It is the parity checker for tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py; the product path never links or calls it.
 *
 * PARITY PINNING: the reference's C++ specs hold no golden vector for computeRHS or any time
 * step; what they do pin -- every input array of the path and the
 * fullToVector/applyIndexMap semantics -- is checked in tests/test_setup_golden.py and
 * tests/test_oracle.py.  The RHS itself is pinned against outputs of the reference's own
 * Python RHS functions run in the build container (sw2dquads.py:10-123, swhelpers/rhs.py:178-311,
 * swhelpers/limiters.py:3-21 executed unmodified by tests/golden/make_python_reference_vectors.py;
 * fixture tests/golden/python_reference_rhs.npz; tests/test_python_reference_golden.py):
 * per-face LF, wall BC, Coriolis/drag/bed-slope sources, tri N=4,7 and quad N=3,6, to 1e-12.
 * The lines only the C++ src/sw2d driver has (global LF maximum, star states, Out BC with tide,
 * SSP-RK2 + sponge) have no runnable reference here and stay pinned by transcription, the second
 * NumPy derivation (oracle/sw2d_numpy.py) and the self-checks of SURVEY.md §8c.
 *
 * Followed sources (reference = WQCG/blitzdg @ d61b584):
 *   full RHS     src/sw2d/main.cpp:279-484        dt   src/sw2d/main.cpp:253-277
 *   simple RHS   src/sw2d-simple/main.cpp:181-356 (per-face LF speed :305-312)
 *   stage loop   include/LSERK4.hpp:16-27 + src/advec1d/main.cpp:92-102
 *   helpers      include/BlitzHelpers.hpp:215-297
 * Compile with -O2 -ffp-contract=off (the reference's x86-64 build emits no FMAs).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/blitzdg_b200.h"

#ifdef _OPENMP
#include <omp.h>
#define PFOR _Pragma("omp parallel for schedule(static)")
#else
#define PFOR
#endif

static const double ORACLE_PI = 3.1415926535897932; /* include/Constants.hpp */

int oracle_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* include/BlitzHelpers.hpp:215-233 with byRows=false: out[i + rows*j] = M(i,j), M row-major [rows][cols] */
void oracle_full_to_vector(const double *M, double *v, int rows, int cols) {
  PFOR
  for (int j = 0; j < cols; ++j)
    for (int i = 0; i < rows; ++i) v[i + (size_t)rows * j] = M[(size_t)i * cols + j];
}
/* include/BlitzHelpers.hpp:283-286: M(i,j) = v[i + rows*j] */
void oracle_vector_to_full(const double *v, double *M, int rows, int cols) {
  PFOR
  for (int j = 0; j < cols; ++j)
    for (int i = 0; i < rows; ++i) M[(size_t)i * cols + j] = v[i + (size_t)rows * j];
}
/* include/BlitzHelpers.hpp:270-275: out[k] = vec[map[k]] */
void oracle_apply_index_map(const double *vec, const int32_t *map, double *out, int n) {
  PFOR
  for (int k = 0; k < n; ++k) out[k] = vec[map[k]];
}

static double *dalloc(size_t n) { return (double *)malloc(sizeof(double) * (n ? n : 1)); }

/* sum(D(ii,kk)*F(kk,jj),kk): sequential k-sum from 0.0, mul then add (blitz reduction semantics) */
static void contract(const double *D, int rows, int inner, const double *F, int K, double *out) {
  PFOR
  for (int i = 0; i < rows; ++i)
    for (int k = 0; k < K; ++k) {
      double acc = 0.0;
      for (int j = 0; j < inner; ++j) acc = acc + D[(size_t)i * inner + j] * F[(size_t)j * K + k];
      out[(size_t)i * K + k] = acc;
    }
}

/* One evaluation of sw2d::computeRHS.  ghost (may be NULL): n_ghost values each of h, hu, hv, H
 * laid out [4][n_ghost]; vmapP entries >= Np*K address them (multi-rank halo). */
int oracle_sw2d_rhs(const bdg_dgcontext2d *dg, const bdg_sw2d_params *prm, const double *H,
                    const double *Hx, const double *Hy, const double *h, const double *hu,
                    const double *hv, double t, int n_ghost, const double *ghost, double *RHS1,
                    double *RHS2, double *RHS3) {
  const int Np = dg->Np, Nfp = dg->Nfp, Nf = dg->Nfaces, K = dg->K;
  const int Nfn = Nf * Nfp;
  const size_t nF = (size_t)Nfn * K, nT = (size_t)Np * K;
  const double g = prm->g;
  const int full = prm->sources != 0;

  double *hVec = dalloc(nT + n_ghost), *huVec = dalloc(nT + n_ghost), *hvVec = dalloc(nT + n_ghost),
         *HVec = dalloc(nT + n_ghost);
  double *nxV = dalloc(nF), *nyV = dalloc(nF);
  double *hM = dalloc(nF), *hP = dalloc(nF), *huM = dalloc(nF), *huP = dalloc(nF), *hvM = dalloc(nF),
         *hvP = dalloc(nF), *HM = dalloc(nF), *HP = dalloc(nF);

  oracle_full_to_vector(dg->nx, nxV, Nfn, K);
  oracle_full_to_vector(dg->ny, nyV, Nfn, K);
  oracle_full_to_vector(h, hVec, Np, K);
  oracle_full_to_vector(hu, huVec, Np, K);
  oracle_full_to_vector(hv, hvVec, Np, K);
  if (full) oracle_full_to_vector(H, HVec, Np, K);
  if (n_ghost > 0) {
    memcpy(hVec + nT, ghost, sizeof(double) * n_ghost);
    memcpy(huVec + nT, ghost + n_ghost, sizeof(double) * n_ghost);
    memcpy(hvVec + nT, ghost + 2 * (size_t)n_ghost, sizeof(double) * n_ghost);
    if (full) memcpy(HVec + nT, ghost + 3 * (size_t)n_ghost, sizeof(double) * n_ghost);
  }
  oracle_apply_index_map(hVec, dg->vmapM, hM, (int)nF);
  oracle_apply_index_map(hVec, dg->vmapP, hP, (int)nF);
  oracle_apply_index_map(huVec, dg->vmapM, huM, (int)nF);
  oracle_apply_index_map(huVec, dg->vmapP, huP, (int)nF);
  oracle_apply_index_map(hvVec, dg->vmapM, hvM, (int)nF);
  oracle_apply_index_map(hvVec, dg->vmapP, hvP, (int)nF);
  if (full) {
    oracle_apply_index_map(HVec, dg->vmapM, HM, (int)nF);
    oracle_apply_index_map(HVec, dg->vmapP, HP, (int)nF);
  }

  /* wall BC (main.cpp:340-345; simple :256-260 leaves hP alone, vmapP==vmapM there anyway) */
  {
    const int32_t *mapW = dg->bc_face_nodes[3];
    for (int i = 0; i < dg->bc_count[3]; ++i) {
      int w = mapW[i];
      if (full) hP[w] = hM[w];
      huP[w] = huM[w] - 2 * nxV[w] * (huM[w] * nxV[w] + hvM[w] * nyV[w]);
      hvP[w] = hvM[w] - 2 * nyV[w] * (huM[w] * nxV[w] + hvM[w] * nyV[w]);
    }
  }
  /* open BC (main.cpp:348-353); an absent Out list is skipped (the reference would throw, F3) */
  if (full) {
    const double T_tide = prm->tide_period, om_tide = 2.0 * ORACLE_PI / T_tide, amp_tide = prm->tide_amp;
    const int32_t *mapO = dg->bc_face_nodes[2];
    for (int i = 0; i < dg->bc_count[2]; ++i) {
      int o = mapO[i];
      huP[o] = huM[o];
      hvP[o] = hvM[o];
      hP[o] = HM[o] + amp_tide * cos(om_tide * t) * 0.5 * (tanh(prm->tide_ramp * (t - T_tide)) + 1);
    }
  }
  /* star states (main.cpp:355-368) */
  if (full) {
    double *hMs = dalloc(nF), *hPs = dalloc(nF);
    PFOR
    for (size_t i = 0; i < nF; ++i) {
      double bM = -HM[i], bP = -HP[i];
      hMs[i] = fmax(0.0, hM[i] + bM - fmax(bP, bM));
      hPs[i] = fmax(0.0, hP[i] + bP - fmax(bP, bM));
    }
    PFOR
    for (size_t i = 0; i < nF; ++i) { hM[i] = hMs[i]; hP[i] = hPs[i]; }
    PFOR
    for (size_t i = 0; i < nF; ++i) { huM[i] = hMs[i] * (huM[i] / hM[i]); huP[i] = hPs[i] * (huP[i] / hP[i]); }
    PFOR
    for (size_t i = 0; i < nF; ++i) { hvM[i] = hMs[i] * (hvM[i] / hM[i]); hvP[i] = hPs[i] * (hvP[i] / hP[i]); }
    free(hMs); free(hPs);  /* (0.5*g*hM*hM - 0.5*g*hMstar*hMstar) is identically 0 after hM=hMstar */
  }

  double *dh = dalloc(nF), *dhu = dalloc(nF), *dhv = dalloc(nF);
  PFOR
  for (size_t i = 0; i < nF; ++i) dh[i] = hM[i] - hP[i];
  PFOR
  for (size_t i = 0; i < nF; ++i) dhu[i] = huM[i] - huP[i];
  PFOR
  for (size_t i = 0; i < nF; ++i) dhv[i] = hvM[i] - hvP[i];

  /* flux tensor at the traces (main.cpp:376-387): F1=hu, G1=hv */
  double *F2M = dalloc(nF), *G2M = dalloc(nF), *G3M = dalloc(nF), *F2P = dalloc(nF), *G2P = dalloc(nF),
         *G3P = dalloc(nF);
  PFOR
  for (size_t i = 0; i < nF; ++i) F2M[i] = (huM[i] * huM[i]) / hM[i] + 0.5 * g * hM[i] * hM[i];
  PFOR
  for (size_t i = 0; i < nF; ++i) G2M[i] = (huM[i] * hvM[i]) / hM[i];
  PFOR
  for (size_t i = 0; i < nF; ++i) G3M[i] = (hvM[i] * hvM[i]) / hM[i] + 0.5 * g * hM[i] * hM[i];
  PFOR
  for (size_t i = 0; i < nF; ++i) F2P[i] = (huP[i] * huP[i]) / hP[i] + 0.5 * g * hP[i] * hP[i];
  PFOR
  for (size_t i = 0; i < nF; ++i) G2P[i] = (huP[i] * hvP[i]) / hP[i];
  PFOR
  for (size_t i = 0; i < nF; ++i) G3P[i] = (hvP[i] * hvP[i]) / hP[i] + 0.5 * g * hP[i] * hP[i];

  /* volume fluxes (main.cpp:389-394) */
  double *F2 = dalloc(nT), *G2 = dalloc(nT), *G3 = dalloc(nT);
  PFOR
  for (size_t m = 0; m < nT; ++m) F2[m] = (hu[m] * hu[m]) / h[m] + 0.5 * g * h[m] * h[m];
  PFOR
  for (size_t m = 0; m < nT; ++m) G2[m] = (hu[m] * hv[m]) / h[m];
  PFOR
  for (size_t m = 0; m < nT; ++m) G3[m] = (hv[m] * hv[m]) / h[m] + 0.5 * g * h[m] * h[m];

  /* wave speeds (main.cpp:397-414 / simple :289-312) */
  double *spdMax = dalloc(nF);
  {
    double *uM = dalloc(nF), *vM = dalloc(nF), *uP = dalloc(nF), *vP = dalloc(nF), *spdM = dalloc(nF),
           *spdP = dalloc(nF);
    PFOR
    for (size_t i = 0; i < nF; ++i) { uM[i] = huM[i] / hM[i]; vM[i] = hvM[i] / hM[i]; }
    PFOR
    for (size_t i = 0; i < nF; ++i) { uP[i] = huP[i] / hP[i]; vP[i] = hvP[i] / hP[i]; }
    PFOR
    for (size_t i = 0; i < nF; ++i) spdM[i] = sqrt(uM[i] * uM[i] + vM[i] * vM[i]) + sqrt(g * hM[i]);
    PFOR
    for (size_t i = 0; i < nF; ++i) spdP[i] = sqrt(uP[i] * uP[i] + vP[i] * vP[i]) + sqrt(g * hP[i]);
    PFOR
    for (size_t i = 0; i < nF; ++i) spdMax[i] = spdM[i] > spdP[i] ? spdM[i] : spdP[i]; /* std::max(a,b) */
    free(uM); free(vM); free(uP); free(vP); free(spdM); free(spdP);
  }
  if (prm->lf_mode == 1) {
    double gmax = spdMax[0]; /* blitz::max: sequential, NaN never wins a '>' comparison */
    for (size_t i = 1; i < nF; ++i)
      if (spdMax[i] > gmax) gmax = spdMax[i];
    PFOR
    for (size_t i = 0; i < nF; ++i) spdMax[i] = gmax;
  } else {
    const size_t nfaces = (size_t)Nf * K;
    PFOR
    for (size_t c = 0; c < nfaces; ++c) {
      double m = spdMax[c * Nfp];
      for (int j = 1; j < Nfp; ++j)
        if (spdMax[c * Nfp + j] > m) m = spdMax[c * Nfp + j];
      for (int j = 0; j < Nfp; ++j) spdMax[c * Nfp + j] = m;
    }
  }

  /* strong-form flux jump (main.cpp:416-421): F3 == G2 */
  double *dF1 = dalloc(nF), *dF2 = dalloc(nF), *dF3 = dalloc(nF);
  PFOR
  for (size_t i = 0; i < nF; ++i)
    dF1[i] = 0.5 * ((huM[i] - huP[i]) * nxV[i] + (hvM[i] - hvP[i]) * nyV[i] - spdMax[i] * dh[i]);
  PFOR
  for (size_t i = 0; i < nF; ++i)
    dF2[i] = 0.5 * ((F2M[i] - F2P[i]) * nxV[i] + (G2M[i] - G2P[i]) * nyV[i] - spdMax[i] * dhu[i]);
  PFOR
  for (size_t i = 0; i < nF; ++i)
    dF3[i] = 0.5 * ((G2M[i] - G2P[i]) * nxV[i] + (G3M[i] - G3P[i]) * nyV[i] - spdMax[i] * dhv[i]);

  double *dF1m = dalloc(nF), *dF2m = dalloc(nF), *dF3m = dalloc(nF);
  oracle_vector_to_full(dF1, dF1m, Nfn, K);
  oracle_vector_to_full(dF2, dF2m, Nfn, K);
  oracle_vector_to_full(dF3, dF3m, Nfn, K);

  /* flux divergence (main.cpp:434-442) */
  {
    double *a = dalloc(nT), *b = dalloc(nT);
    const double *F[3] = {hu, F2, G2}, *G[3] = {hv, G2, G3};
    double *R[3] = {RHS1, RHS2, RHS3};
    for (int c = 0; c < 3; ++c) {
      contract(dg->Dr, Np, Np, F[c], K, a);
      contract(dg->Ds, Np, Np, F[c], K, b);
      PFOR
      for (size_t m = 0; m < nT; ++m) R[c][m] = -(dg->rx[m] * a[m] + dg->sx[m] * b[m]);
      contract(dg->Dr, Np, Np, G[c], K, a);
      contract(dg->Ds, Np, Np, G[c], K, b);
      PFOR
      for (size_t m = 0; m < nT; ++m) R[c][m] = R[c][m] + (-(dg->ry[m] * a[m] + dg->sy[m] * b[m]));
    }
    /* surface term (main.cpp:444-458) */
    double *sr = dalloc(nF);
    const double *dFm[3] = {dF1m, dF2m, dF3m};
    for (int c = 0; c < 3; ++c) {
      PFOR
      for (size_t i = 0; i < nF; ++i) sr[i] = dg->Fscale[i] * dFm[c][i];
      contract(dg->Lift, Np, Nfn, sr, K, a);
      PFOR
      for (size_t m = 0; m < nT; ++m) R[c][m] = R[c][m] + a[m];
    }
    free(sr);
    /* sources (main.cpp:460-483) */
    if (full) {
      double *u = dalloc(nT), *v = dalloc(nT), *nu = dalloc(nT);
      PFOR
      for (size_t m = 0; m < nT; ++m) { u[m] = hu[m] / h[m]; v[m] = hv[m] / h[m]; }
      PFOR
      for (size_t m = 0; m < nT; ++m) { RHS2[m] = RHS2[m] + g * h[m] * Hx[m]; RHS3[m] = RHS3[m] + g * h[m] * Hy[m]; }
      PFOR
      for (size_t m = 0; m < nT; ++m) nu[m] = sqrt(u[m] * u[m] + v[m] * v[m]);
      PFOR
      for (size_t m = 0; m < nT; ++m) { RHS2[m] = RHS2[m] + (-prm->CD * u[m] * nu[m]); RHS3[m] = RHS3[m] + (-prm->CD * v[m] * nu[m]); }
      PFOR
      for (size_t m = 0; m < nT; ++m) { RHS2[m] = RHS2[m] + prm->f * hv[m]; RHS3[m] = RHS3[m] + (-prm->f * hu[m]); }
      free(u); free(v); free(nu);
    }
    /* optional RHS filter (sw2d-simple main.cpp:134-136) */
    if (prm->filter && dg->Filt) {
      for (int c = 0; c < 3; ++c) {
        contract(dg->Filt, Np, Np, R[c], K, a);
        memcpy(R[c], a, sizeof(double) * nT);
      }
    }
    free(a); free(b);
  }

  free(hVec); free(huVec); free(hvVec); free(HVec); free(nxV); free(nyV);
  free(hM); free(hP); free(huM); free(huP); free(hvM); free(hvP); free(HM); free(HP);
  free(dh); free(dhu); free(dhv);
  free(F2M); free(G2M); free(G3M); free(F2P); free(G2P); free(G3P); free(F2); free(G2); free(G3);
  free(spdMax); free(dF1); free(dF2); free(dF3); free(dF1m); free(dF2m); free(dF3m);
  return 0;
}

/* 4th (tracer) equation of the reference's Python drivers, swhelpers/rhs.py:206-297 with swhelpers/flux.py:18-20:
 * F4 = hN*u, G4 = hN*v, per-face LF speed of the flow (rhs.py:259-268), reflected momentum on walls (rhs.py:236-238;
 * hN+ = hN- there because vmapP = vmapM).  Sweep structure of the Python code: whole-array statements, np.dot
 * contractions (sequential here).  RHS4 [Np][K] row-major. */
int oracle_sw2d_rhs_tracer(const bdg_dgcontext2d *dg, const bdg_sw2d_params *prm, const double *h, const double *hu,
                           const double *hv, const double *hN, double *RHS4) {
  const int Np = dg->Np, Nfp = dg->Nfp, Nf = dg->Nfaces, K = dg->K, Nfn = Nf * Nfp;
  const size_t nF = (size_t)Nfn * K, nT = (size_t)Np * K;
  const double g = prm->g;
  double *hC = dalloc(nT), *huC = dalloc(nT), *hvC = dalloc(nT), *hNC = dalloc(nT), *nxC = dalloc(nF), *nyC = dalloc(nF);
  oracle_full_to_vector(h, hC, Np, K);
  oracle_full_to_vector(hu, huC, Np, K);
  oracle_full_to_vector(hv, hvC, Np, K);
  oracle_full_to_vector(hN, hNC, Np, K);
  oracle_full_to_vector(dg->nx, nxC, Nfn, K);
  oracle_full_to_vector(dg->ny, nyC, Nfn, K);
  double *hM = dalloc(nF), *hP = dalloc(nF), *uM = dalloc(nF), *uP = dalloc(nF), *vM = dalloc(nF), *vP = dalloc(nF),
         *nM = dalloc(nF), *nP = dalloc(nF), *huM = dalloc(nF), *huP = dalloc(nF), *hvM = dalloc(nF), *hvP = dalloc(nF);
  for (size_t i = 0; i < nF; ++i) {
    const int mM = dg->vmapM[i], mP = dg->vmapP[i];
    hM[i] = hC[mM]; hP[i] = hC[mP];
    uM[i] = huC[mM] / hC[mM]; uP[i] = huC[mP] / hC[mP];
    vM[i] = hvC[mM] / hC[mM]; vP[i] = hvC[mP] / hC[mP];
    nM[i] = hNC[mM]; nP[i] = hNC[mP];
  }
  for (size_t i = 0; i < nF; ++i) { huM[i] = hM[i] * uM[i]; hvM[i] = hM[i] * vM[i]; huP[i] = hP[i] * uP[i]; hvP[i] = hP[i] * vP[i]; } /* rhs.py:228-232 */
  for (int j = 0; j < dg->bc_count[3]; ++j) { /* rhs.py:236-238 */
    const int w = dg->bc_face_nodes[3][j];
    const double d = huM[w] * nxC[w] + hvM[w] * nyC[w];
    huP[w] = huM[w] - 2 * nxC[w] * d;
    hvP[w] = hvM[w] - 2 * nyC[w] * d;
  }
  double *F4M = dalloc(nF), *G4M = dalloc(nF), *F4P = dalloc(nF), *G4P = dalloc(nF), *spd = dalloc(nF), *dF = dalloc(nF);
  for (size_t i = 0; i < nF; ++i) { /* flux.py:6-20 at the traces */
    const double um = huM[i] / hM[i], vm = hvM[i] / hM[i], up = huP[i] / hP[i], vp = hvP[i] / hP[i];
    F4M[i] = nM[i] * um; G4M[i] = nM[i] * vm;
    F4P[i] = nP[i] * up; G4P[i] = nP[i] * vp;
    const double sM = sqrt(um * um + vm * vm) + sqrt(g * hM[i]), sP = sqrt(up * up + vp * vp) + sqrt(g * hP[i]);
    spd[i] = sM > sP ? sM : sP;
  }
  for (size_t c = 0; c < (size_t)Nf * K; ++c) { /* rhs.py:266-268 */
    double m = spd[c * Nfp];
    for (int j = 1; j < Nfp; ++j)
      if (spd[c * Nfp + j] > m) m = spd[c * Nfp + j];
    for (int j = 0; j < Nfp; ++j) spd[c * Nfp + j] = m;
  }
  for (size_t i = 0; i < nF; ++i) dF[i] = 0.5 * ((F4M[i] - F4P[i]) * nxC[i] + (G4M[i] - G4P[i]) * nyC[i] - spd[i] * (nM[i] - nP[i]));
  double *dFm = dalloc(nF), *F4 = dalloc(nT), *G4 = dalloc(nT), *a = dalloc(nT), *b = dalloc(nT), *sr = dalloc(nF);
  oracle_vector_to_full(dF, dFm, Nfn, K);
  for (size_t m = 0; m < nT; ++m) { const double u = hu[m] / h[m], v = hv[m] / h[m]; F4[m] = hN[m] * u; G4[m] = hN[m] * v; }
  contract(dg->Dr, Np, Np, F4, K, a);
  contract(dg->Ds, Np, Np, F4, K, b);
  for (size_t m = 0; m < nT; ++m) RHS4[m] = -(dg->rx[m] * a[m] + dg->sx[m] * b[m]);
  contract(dg->Dr, Np, Np, G4, K, a);
  contract(dg->Ds, Np, Np, G4, K, b);
  for (size_t m = 0; m < nT; ++m) RHS4[m] = RHS4[m] + (-(dg->ry[m] * a[m] + dg->sy[m] * b[m]));
  for (size_t i = 0; i < nF; ++i) sr[i] = dg->Fscale[i] * dFm[i];
  contract(dg->Lift, Np, Nfn, sr, K, a);
  for (size_t m = 0; m < nT; ++m) RHS4[m] = RHS4[m] + a[m];
  free(hC); free(huC); free(hvC); free(hNC); free(nxC); free(nyC); free(hM); free(hP); free(uM); free(uP); free(vM); free(vP);
  free(nM); free(nP); free(huM); free(huP); free(hvM); free(hvP); free(F4M); free(G4M); free(F4P); free(G4P); free(spd); free(dF);
  free(dFm); free(F4); free(G4); free(a); free(b); free(sr);
  return 0;
}

/* sw2d::computeTimeStep (main.cpp:253-277); use_abs=1 gives the sw2d-simple variant (abs(Fscale), :108-109).
 * Returns max_i(Fscale_i*spd_i) in *fsc_max as well so ranks can reduce it. */
double oracle_sw2d_timestep(const bdg_dgcontext2d *dg, const bdg_sw2d_params *prm, const double *h,
                            const double *hu, const double *hv, int use_abs, double *fsc_max) {
  const int Np = dg->Np, Nfn = dg->Nfaces * dg->Nfp, K = dg->K, N = dg->N;
  const size_t nF = (size_t)Nfn * K, nT = (size_t)Np * K;
  double *u = dalloc(nT), *v = dalloc(nT), *uV = dalloc(nT), *vV = dalloc(nT), *hV = dalloc(nT), *fs = dalloc(nF);
  double *uM = dalloc(nF), *vM = dalloc(nF), *hM = dalloc(nF);
  for (size_t m = 0; m < nT; ++m) { u[m] = hu[m] / h[m]; v[m] = hv[m] / h[m]; }
  oracle_full_to_vector(u, uV, Np, K);
  oracle_full_to_vector(v, vV, Np, K);
  oracle_full_to_vector(h, hV, Np, K);
  oracle_full_to_vector(dg->Fscale, fs, Nfn, K);
  oracle_apply_index_map(uV, dg->vmapM, uM, (int)nF);
  oracle_apply_index_map(vV, dg->vmapM, vM, (int)nF);
  oracle_apply_index_map(hV, dg->vmapM, hM, (int)nF);
  double mx = -INFINITY;
  for (size_t i = 0; i < nF; ++i) {
    double spd = sqrt(uM[i] * uM[i] + vM[i] * vM[i]) + sqrt(prm->g * hM[i]);
    double val = (use_abs ? fabs(fs[i]) : fs[i]) * spd;
    if (val > mx) mx = val;
  }
  free(u); free(v); free(uV); free(vV); free(hV); free(fs); free(uM); free(vM); free(hM);
  if (fsc_max) *fsc_max = mx;
  return prm->CFL / ((N + 1) * (N + 1) * 0.5 * mx);
}

/* LSERK4 coefficients exactly as written in include/LSERK4.hpp:17-27 */
static const double rk4a[5] = {0.0, -567301805773.0 / 1357537059087.0, -2404267990393.0 / 2016746695238.0,
                               -3550918686646.0 / 2091501179385.0, -1275806237668.0 / 842570457699.0};
static const double rk4b[5] = {1432997174477.0 / 9575080441755.0, 5161836677717.0 / 13612068292357.0,
                               1720146321549.0 / 2090206949498.0, 3134564353537.0 / 4481467310338.0,
                               2277821191437.0 / 14882151754819.0};

void oracle_lserk4_coeffs(double *a, double *b) {
  memcpy(a, rk4a, sizeof(rk4a));
  memcpy(b, rk4b, sizeof(rk4b));
}

/* one stage: res = a*res + dt*RHS ; q += b*res  (src/advec1d/main.cpp:98,101) */
void oracle_lserk4_stage(int stage, double dt, size_t n, const double *rhs, double *res, double *q) {
  const double a = rk4a[stage], b = rk4b[stage];
  PFOR
  for (size_t m = 0; m < n; ++m) res[m] = a * res[m] + dt * rhs[m];
  PFOR
  for (size_t m = 0; m < n; ++m) q[m] = q[m] + b * res[m];
}

/* nsteps LSERK4 steps on one rank; the stage time is the step-start time (LSERK4.hpp has no rk4c). */
int oracle_sw2d_step(const bdg_dgcontext2d *dg, const bdg_sw2d_params *prm, const double *H, const double *Hx,
                     const double *Hy, double *h, double *hu, double *hv, double *res1, double *res2,
                     double *res3, int nsteps, double dt, double *t) {
  const size_t nT = (size_t)dg->Np * dg->K;
  double *R1 = dalloc(nT), *R2 = dalloc(nT), *R3 = dalloc(nT);
  for (int s = 0; s < nsteps; ++s) {
    double dts = dt > 0 ? dt : oracle_sw2d_timestep(dg, prm, h, hu, hv, 0, NULL);
    for (int st = 0; st < 5; ++st) {
      oracle_sw2d_rhs(dg, prm, H, Hx, Hy, h, hu, hv, *t, 0, NULL, R1, R2, R3);
      oracle_lserk4_stage(st, dts, nT, R1, res1, h);
      oracle_lserk4_stage(st, dts, nT, R2, res2, hu);
      oracle_lserk4_stage(st, dts, nT, R3, res3, hv);
    }
    *t += dts;
  }
  free(R1); free(R2); free(R3);
  return 0;
}

/* ---- the rest of the src/sw2d driver around computeRHS ---------------------------- */

/* sw2d::buildSpongeCoeff (src/sw2d/main.cpp:517-553): for every volume node the distance to the closest node of
 * an Out face (x,y of the '-' side through vmapM); inside radInfl the coefficient falls linearly from
 * spongeStrength at the boundary to 0.  x, y, sponge are [Np][K] row-major; out_nodes = BCmap[Out] (face-node ids).
 * Nodes out of reach are left untouched (the driver zero-fills before the call, :185). */
void oracle_build_sponge(const bdg_dgcontext2d *dg, const double *x, const double *y, const int32_t *out_nodes,
                         int n_out, double strength, double radInfl, double *sponge) {
  const int Np = dg->Np, K = dg->K;
  const size_t nT = (size_t)Np * K;
  double *xV = dalloc(nT), *yV = dalloc(nT);
  oracle_full_to_vector(x, xV, Np, K);
  oracle_full_to_vector(y, yV, Np, K);
  for (int k = 0; k < K; ++k)
    for (int n = 0; n < Np; ++n) {
      const double x0 = x[(size_t)n * K + k], y0 = y[(size_t)n * K + k];
      double closest = 1.0e12;
      for (int i = 0; i < n_out; ++i) {
        const int32_t o = dg->vmapM[out_nodes[i]];
        const double dist = hypot(x0 - xV[o], y0 - yV[o]);
        if (dist < radInfl && dist < closest) closest = dist;
      }
      if (closest < 1.0e12) sponge[(size_t)n * K + k] = strength * (1.0 - (closest / radInfl));
    }
  free(xV); free(yV);
}

/* sw2d::readDepthData (src/sw2d/main.cpp:486-515) after the file has been parsed: values arrive element-major
 * (count runs over k then n), are floored at 150 and stored to H(n,k); missing values are 0 -> 150. */
void oracle_depth_to_H(const double *values, size_t nvalues, int Np, int K, double *H) {
  size_t count = 0;
  for (int k = 0; k < K; ++k)
    for (int n = 0; n < Np; ++n) {
      double val = count < nvalues ? values[count] : 0.0;
      if (val < 150.0) val = 150.0;
      H[(size_t)n * K + k] = val;
      ++count;
    }
}

/* bed slopes (src/sw2d/main.cpp:129-133): Hx = rx*(Dr H) + sx*(Ds H), Hy likewise, then both filtered. */
void oracle_bed_slopes(const bdg_dgcontext2d *dg, const double *H, double *Hx, double *Hy) {
  const int Np = dg->Np, K = dg->K;
  const size_t nT = (size_t)Np * K;
  double *dr = dalloc(nT), *ds = dalloc(nT), *tx = dalloc(nT), *ty = dalloc(nT);
  contract(dg->Dr, Np, Np, H, K, dr);
  contract(dg->Ds, Np, Np, H, K, ds);
  for (size_t m = 0; m < nT; ++m) {
    tx[m] = dg->rx[m] * dr[m] + dg->sx[m] * ds[m];
    ty[m] = dg->ry[m] * dr[m] + dg->sy[m] * ds[m];
  }
  if (dg->Filt) {
    contract(dg->Filt, Np, Np, tx, K, Hx);
    contract(dg->Filt, Np, Np, ty, K, Hy);
  } else {
    memcpy(Hx, tx, sizeof(double) * nT);
    memcpy(Hy, ty, sizeof(double) * nT);
  }
  free(dr); free(ds); free(tx); free(ty);
}

/* the driver's time loop (src/sw2d/main.cpp:192-244): dt = computeTimeStep, Heun's SSP-RK2 with both RHS
 * evaluations at the step-start time, sponge relaxation of hu, hv after each stage; sponge may be NULL. */
int oracle_sw2d_step_ssprk2(const bdg_dgcontext2d *dg, const bdg_sw2d_params *prm, const double *H, const double *Hx,
                            const double *Hy, const double *sponge, double *h, double *hu, double *hv, int nsteps,
                            double dt, double *t) {
  const size_t nT = (size_t)dg->Np * dg->K;
  double *R1 = dalloc(nT), *R2 = dalloc(nT), *R3 = dalloc(nT), *h1 = dalloc(nT), *hu1 = dalloc(nT), *hv1 = dalloc(nT);
  for (int s = 0; s < nsteps; ++s) {
    const double dts = dt > 0 ? dt : oracle_sw2d_timestep(dg, prm, h, hu, hv, 0, NULL);
    oracle_sw2d_rhs(dg, prm, H, Hx, Hy, h, hu, hv, *t, 0, NULL, R1, R2, R3);
    for (size_t m = 0; m < nT; ++m) {
      h1[m] = h[m] + dts * R1[m];
      hu1[m] = hu[m] + dts * R2[m];
      hv1[m] = hv[m] + dts * R3[m];
    }
    if (sponge)
      for (size_t m = 0; m < nT; ++m) {
        hu1[m] /= (1.0 + sponge[m] * hu1[m] * hu1[m]);
        hv1[m] /= (1.0 + sponge[m] * hv1[m] * hv1[m]);
      }
    oracle_sw2d_rhs(dg, prm, H, Hx, Hy, h1, hu1, hv1, *t, 0, NULL, R1, R2, R3);
    for (size_t m = 0; m < nT; ++m) {
      h[m] = 0.5 * (h[m] + h1[m] + dts * R1[m]);
      hu[m] = 0.5 * (hu[m] + hu1[m] + dts * R2[m]);
      hv[m] = 0.5 * (hv[m] + hv1[m] + dts * R3[m]);
    }
    if (sponge)
      for (size_t m = 0; m < nT; ++m) {
        hu[m] /= (1.0 + sponge[m] * hu[m] * hu[m]);
        hv[m] /= (1.0 + sponge[m] * hv[m] * hv[m]);
      }
    *t += dts;
  }
  free(R1); free(R2); free(R3); free(h1); free(hu1); free(hv1);
  return 0;
}
