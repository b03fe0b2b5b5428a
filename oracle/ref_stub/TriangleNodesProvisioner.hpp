// TEST INFRASTRUCTURE -- stand-in for the reference's include/TriangleNodesProvisioner.hpp, found FIRST on the include
// path when oracle/make_ref.py compiles the reference's own SW2d.hpp / computeRHS translation-unit extracts.  The real
// class needs LAPACK, SuiteSparse, METIS and boost headers (absent here) and is not on the hot path: the
// RHS only READS arrays through its accessors (src/sw2d-simple/main.cpp:188-212, 342-343).  This holder offers exactly
// those accessor names over arrays the caller supplies; the reference's DGContext2D.hpp is the real one.
#pragma once
#include "DGContext2D.hpp"
#include "Types.hpp"

namespace blitzdg {
class TriangleNodesProvisioner {
 public:
  static const index_type NumFaces = 3;  // reference include/TriangleNodesProvisioner.hpp (public constant of the class)
  index_type Np_, Nfp_, K_;
  real_matrix_type Dr_, Ds_, Lift_, Filt_, rx_, sx_, ry_, sy_, nx_, ny_, Fscale_, x_, y_;
  index_vector_type vmapM_, vmapP_;
  index_hashmap bcMap_;
  DGContext2D ctx_;

  const real_matrix_type& get_Dr() const { return Dr_; }
  const real_matrix_type& get_Ds() const { return Ds_; }
  const real_matrix_type& get_Lift() const { return Lift_; }
  const real_matrix_type& get_Filter() const { return Filt_; }
  const real_matrix_type& get_rx() const { return rx_; }
  const real_matrix_type& get_sx() const { return sx_; }
  const real_matrix_type& get_ry() const { return ry_; }
  const real_matrix_type& get_sy() const { return sy_; }
  const real_matrix_type& get_nx() const { return nx_; }
  const real_matrix_type& get_ny() const { return ny_; }
  const real_matrix_type& get_Fscale() const { return Fscale_; }
  const real_matrix_type& get_xGrid() const { return x_; }
  const real_matrix_type& get_yGrid() const { return y_; }
  const index_vector_type& get_vmapM() const { return vmapM_; }
  const index_vector_type& get_vmapP() const { return vmapP_; }
  const index_hashmap& get_bcMap() const { return bcMap_; }
  DGContext2D get_DGContext() const { return ctx_; }
  index_type get_NumFacePoints() const { return Nfp_; }
  index_type get_NumLocalPoints() const { return Np_; }
  index_type get_NumElements() const { return K_; }
};
}  // namespace blitzdg
