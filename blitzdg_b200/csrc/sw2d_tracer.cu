// Tracer: the 4th conserved field hN of the reference's Python drivers (swhelpers/rhs.py:206-297, swhelpers/flux.py:18-20;
// sw2d.py:30-33,129-140 -- SURVEY.md 8f rank 4):
//     F4 = hN u,  G4 = hN v,   dFlux4 = ((F4- - F4+) nx + (G4- - G4+) ny - lambda (hN- - hN+)) / 2,
//     RHS4 = -(rx Dr F4 + sx Ds F4) - (ry Dr G4 + sy Ds G4) + LIFT (Fscale dFlux4),
// with the per-face Lax-Friedrichs speed of the flow (rhs.py:259-268), reflected momentum and hN+ = hN- on walls.
// One generic launch per stage (any element shape / order, per-node metrics) beside the flow's stage kernel: it reads
// the SAME pre-stage flow state (the flow kernels ping-pong, so q_in stays intact while the stage runs) and keeps its own
// ping-pong pair and residual.  EB elements per CTA, max(Np, Nfn) threads per element.
#include "sw2d_common.cuh"

namespace {

__global__ void k_tracer(Sw2dView v, const double *__restrict__ q, const double *__restrict__ hN_in, double *__restrict__ hN_out,
                         double *__restrict__ resN, double *__restrict__ rhs_out, int mode, double a, double b, int EB, int T) {
  extern __shared__ double sm[];
  const int Np = v.Np, Nfn = v.Nfn, Nfp = v.Nfp;
  double *sF = sm, *sG = sF + EB * Np, *sS = sG + EB * Np, *sL = sS + EB * Nfn;  // F4, G4, Fscale*dFlux4, speeds
  const int e = threadIdx.x / T, j = threadIdx.x - e * T;
  const long long k = (long long)blockIdx.x * EB + e;
  const bool live = e < EB && k < v.K;
  const double *h = q, *hu = q + v.stride, *hv = q + 2 * v.stride;
  if (live && j < Np) {  // volume fluxes (flux.py:6-20)
    const long long m = k * Np + j;
    const double u = hu[m] / h[m], vv = hv[m] / h[m];
    sF[e * Np + j] = hN_in[m] * u;
    sG[e * Np + j] = hN_in[m] * vv;
  }
  double hM = 1, huM = 0, hvM = 0, nM = 0, hP = 1, huP = 0, hvP = 0, nP = 0, nx = 0, ny = 0;
  if (live && j < Nfn) {  // traces (rhs.py:206-238)
    const long long i = k * Nfn + j, mM = k * Np + v.fmask[j], mP = v.vmapP[i];
    hM = h[mM]; huM = hu[mM]; hvM = hv[mM]; nM = hN_in[mM];
    hP = h[mP]; huP = hu[mP]; hvP = hv[mP]; nP = hN_in[mP];
    nx = v.nx[i]; ny = v.ny[i];
    if (v.bctag[i] == 3) {
      const double d = huM * nx + hvM * ny;
      huP = huM - 2 * nx * d;
      hvP = hvM - 2 * ny * d;
    }
    const double uM = huM / hM, vM = hvM / hM, uP = huP / hP, vP = hvP / hP;
    const double sM = sqrt(uM * uM + vM * vM) + sqrt(v.g * hM), sP = sqrt(uP * uP + vP * vP) + sqrt(v.g * hP);
    sL[e * Nfn + j] = sM > sP ? sM : sP;
  }
  __syncthreads();
  if (live && j < Nfn) {
    const int f = j / Nfp;
    double lam = sL[e * Nfn + f * Nfp];
    for (int t = 1; t < Nfp; ++t) lam = fmax(lam, sL[e * Nfn + f * Nfp + t]);  // rhs.py:266-268: max over the face
    const double uM = huM / hM, vM = hvM / hM, uP = huP / hP, vP = hvP / hP;
    const double dF = 0.5 * ((nM * uM - nP * uP) * nx + (nM * vM - nP * vP) * ny - lam * (nM - nP));
    sS[e * Nfn + j] = v.Fscale[k * Nfn + j] * dF;
  }
  __syncthreads();
  if (live && j < Np) {
    const long long m = k * Np + j;
    double drF = 0, dsF = 0, drG = 0, dsG = 0, lf = 0;
    for (int c = 0; c < Np; ++c) {
      const double dr = v.Dr[j * Np + c], ds = v.Ds[j * Np + c];
      drF += dr * sF[e * Np + c]; dsF += ds * sF[e * Np + c];
      drG += dr * sG[e * Np + c]; dsG += ds * sG[e * Np + c];
    }
    for (int c = 0; c < Nfn; ++c) lf += v.Lift[j * Nfn + c] * sS[e * Nfn + c];
    const double rhs = -(v.rx[m] * drF + v.sx[m] * dsF) - (v.ry[m] * drG + v.sy[m] * dsG) + lf;
    if (mode == 0) {
      rhs_out[m] = rhs;
    } else {
      const double dt = v.scal[S_DT];
      double st = dt * rhs;
      if (mode == 1) {
        st = a * resN[m] + st;
        resN[m] = st;
      }
      hN_out[m] = stepper_value(mode, b, st, hN_in[m], hN_out[m], 0, 0.0);
    }
  }
}

}  // namespace

cudaError_t launch_tracer(bdg_sw2d *s, const double *q, const double *hN_in, double *hN_out, double *rhs_out, int mode, double a, double b) {
  const int T = s->v.Np > s->v.Nfn ? s->v.Np : s->v.Nfn;
  int EB = 256 / T;
  if (EB < 1) EB = 1;
  const size_t smem = sizeof(double) * EB * (2 * (size_t)s->v.Np + 2 * (size_t)s->v.Nfn);
  k_tracer<<<(unsigned)((s->v.K + EB - 1) / EB), EB * T, smem, s->stream>>>(s->v, q, hN_in, hN_out, s->resN, rhs_out, mode, a, b, EB, T);
  s->launches++;
  return cudaGetLastError();
}
