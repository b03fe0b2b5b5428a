// blitzdg_b200_adapter.hpp -- the thin C++ adapter a blitzdg maintainer drops next to src/sw2d/main.cpp: it fills
// bdg_dgcontext2d from the reference's DGContext2D accessors (reference include/DGContext2D.hpp:82-196) and wraps the
// handle.  Header-only; needs the reference's own headers (DGContext2D.hpp, BCtypes.hpp, Types.hpp) and blitz++ on the
// include path, plus libblitzdg_b200.so at link time.  INTEGRATION.md shows the driver loop around it.
//
// Compile-checked in this repository against the reference's headers where they lie (tests/test_adapter_compiles.py;
// blitz++ itself is absent from this image, oracle/blitz_shim stands in for <blitz/array.h>).
#ifndef BLITZDG_B200_ADAPTER_HPP
#define BLITZDG_B200_ADAPTER_HPP

#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "BCtypes.hpp"
#include "DGContext2D.hpp"
#include "blitzdg_b200.h"

namespace blitzdg {

struct B200Sw2d {
  bdg_sw2d* h = nullptr;

  // Arrays are read in the reference's native storage: blitz row-major [Np][K] / [Nfaces*Nfp][K]
  // (include/BlitzHelpers.hpp:166-169), index vectors as built by buildMaps (src/TriangleNodesProvisioner.cpp:895-1020).
  // The library copies everything during create; `dg` may go away afterwards.
  // fullPhysics: the src/sw2d right-hand side (global LF speed, star states, Out BC with tide, sources) instead of the
  // src/sw2d-simple one (per-face LF speed, wall BC).  useFilter: RHS := Filt*RHS (src/sw2d-simple/main.cpp:134-136).
  B200Sw2d(const DGContext2D& dg, double g, double CD, double f, double CFL, bool fullPhysics, bool useFilter = false, int device = 0) {
    bdg_dgcontext2d c{};
    c.N = dg.order();
    c.Np = dg.numLocalPoints();
    c.Nfp = dg.numFacePoints();
    c.Nfaces = dg.numFaces();
    c.K = dg.numElements();
    c.Dr = dg.Dr().data();
    c.Ds = dg.Ds().data();
    c.Lift = dg.lift().data();
    c.Filt = useFilter ? dg.filter().data() : nullptr;
    c.rx = dg.rx().data();
    c.sx = dg.sx().data();
    c.ry = dg.ry().data();
    c.sy = dg.sy().data();
    c.nx = dg.nx().data();
    c.ny = dg.ny().data();
    c.Fscale = dg.fscale().data();
    c.vmapM = dg.vmapM().data();
    c.vmapP = dg.vmapP().data();
    std::vector<std::vector<int32_t>> bc(9);  // keeps the BC lists alive during create only
    for (const auto& kv : dg.bcmap())         // index_hashmap: tag -> face-node ids; an absent tag is an empty list
      if (kv.first >= 1 && kv.first <= 8) bc[kv.first].assign(kv.second.begin(), kv.second.end());
    for (int t = 1; t <= 8; ++t) {
      c.bc_face_nodes[t] = bc[t].data();
      c.bc_count[t] = static_cast<int32_t>(bc[t].size());
    }
    bdg_sw2d_params p{};
    p.g = g;
    p.CD = CD;
    p.f = f;
    p.CFL = CFL;
    p.lf_mode = fullPhysics ? 1 : 0;
    p.sources = fullPhysics ? 1 : 0;
    p.filter = useFilter ? 1 : 0;
    p.tide_amp = 3.0;                 // src/sw2d/main.cpp:283
    p.tide_period = 3600 * 12.42;     // :281
    p.tide_ramp = 0.15 / 3600;        // :352
    if (bdg_sw2d_create(&h, &c, &p, nullptr, device) != BDG_OK) throw std::runtime_error(bdg_last_error(nullptr));
  }
  ~B200Sw2d() { bdg_sw2d_destroy(h); }
  B200Sw2d(const B200Sw2d&) = delete;
  B200Sw2d& operator=(const B200Sw2d&) = delete;

  // throws like the reference does: std::runtime_error, "A numerical instability has occurred!" included (BDG_E_UNSTABLE)
  void check(int rc) const {
    if (rc != BDG_OK) throw std::runtime_error(bdg_last_error(h));
  }
  // blitz matrices [Np][K] in, as sw2d::fields holds them (src/sw2d/SW2d.hpp:32-49)
  void setBathymetry(const real_matrix_type& H, const real_matrix_type& Hx, const real_matrix_type& Hy) {
    check(bdg_sw2d_set_bathymetry(h, H.data(), Hx.data(), Hy.data()));
  }
  void setState(const real_matrix_type& hh, const real_matrix_type& hu, const real_matrix_type& hv, double t) {
    check(bdg_sw2d_set_state(h, hh.data(), hu.data(), hv.data(), t));
  }
  void getState(real_matrix_type& hh, real_matrix_type& hu, real_matrix_type& hv, double& t) {
    check(bdg_sw2d_get_state(h, hh.data(), hu.data(), hv.data(), &t));
  }
  // sw2d::computeRHS (src/sw2d/SW2d.hpp:53): one evaluation on the device state
  void computeRHS(real_matrix_type& RHS1, real_matrix_type& RHS2, real_matrix_type& RHS3) {
    check(bdg_sw2d_rhs(h, RHS1.data(), RHS2.data(), RHS3.data()));
  }
  // sw2d::computeTimeStep (src/sw2d/SW2d.hpp:54)
  double computeTimeStep() {
    double dt = 0.0;
    check(bdg_sw2d_timestep(h, &dt));
    return dt;
  }
  // nsteps LSERK4 steps (include/LSERK4.hpp:16-27), 5 fused RHS+stage launches each; dt <= 0: adaptive on the device
  void step(int nsteps, double dt) { check(bdg_sw2d_step(h, nsteps, dt)); }
  // the src/sw2d driver's own scheme (SSP-RK2 + sponge relaxation, src/sw2d/main.cpp:213-236)
  void setSponge(const real_matrix_type& spongeCoeff) { check(bdg_sw2d_set_sponge(h, spongeCoeff.data())); }
  void stepSSPRK2(int nsteps, double dt) { check(bdg_sw2d_step_ssprk2(h, nsteps, dt)); }
  // the blow-up guard of the driver loop (src/sw2d/main.cpp:238-240)
  double checkStability() {
    double etaMax = 0.0;
    check(bdg_sw2d_check(h, &etaMax));
    return etaMax;
  }
};

}  // namespace blitzdg

#endif  // BLITZDG_B200_ADAPTER_HPP
