// Translation unit of tests/test_adapter_compiles.py: the driver loop of INTEGRATION.md written against
// include/blitzdg_b200_adapter.hpp and the reference's own DGContext2D.hpp (found under $BLITZDG_REFERENCE/include).
#include "blitzdg_b200_adapter.hpp"

double drive(const blitzdg::DGContext2D& dg, double finalTime, int outputInterval) {
  using blitzdg::real_matrix_type;
  const int Np = dg.numLocalPoints(), K = dg.numElements();
  real_matrix_type h(Np, K), hu(Np, K), hv(Np, K), H(Np, K), Hx(Np, K), Hy(Np, K), sponge(Np, K), R1(Np, K), R2(Np, K), R3(Np, K);
  double t = 0.0;
  blitzdg::B200Sw2d gpu(dg, 9.81, 2.5e-3, 1.0070e-4, 0.25, /*fullPhysics=*/true);
  gpu.setBathymetry(H, Hx, Hy);
  gpu.setState(h, hu, hv, t);
  gpu.computeRHS(R1, R2, R3);
  const double dt0 = gpu.computeTimeStep();
  while (t < finalTime) {
    gpu.step(outputInterval, /*dt=*/0.0);
    gpu.checkStability();
    gpu.getState(h, hu, hv, t);
  }
  gpu.setSponge(sponge);
  gpu.stepSSPRK2(outputInterval, dt0);
  return t;
}
