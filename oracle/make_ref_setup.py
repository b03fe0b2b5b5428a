#!/usr/bin/env python
"""TEST INFRASTRUCTURE -- builds oracle/_ref/libsetup_ref.so: the reference's OWN builders of every array the hot path is
handed (DGContext2D), compiled verbatim from where they lie under /root/reference:

  TriangleNodesProvisioner   src/TriangleNodesProvisioner.cpp
    computeEquilateralNodes :549, computeWarpFactor :604, xyTors :406, buildNodes :678 (r, s, Fmask)
    buildLift :1060, buildPhysicalGrid :730 (V, Vinv, Dr, Ds, x, y, J, rx..sy, nx, ny, Fscale), buildMaps :895 (vmapM, vmapP,
    mapP, mapB, vmapB), buildBCHash :1022/:1028, buildFilter :515, and the basis helpers they call (:383-513, :642)
  QuadNodesProvisioner       src/QuadNodesProvisioner.cpp
    computeVandermondeMatrix :76, computeGradVandermondeMatrix :94, computeDifferentiationMatrices :114, buildFilter :170,
    buildNodes :204, buildPhysicalGrid :277, buildMaps :450, buildBCHash :577/:583, buildLift :615
  JacobiBuilders             src/JacobiBuilders.cpp: computeJacobiPolynomial :18, computeGaussLobottoPoints :98, computeGradJacobi :117
  VandermondeBuilders        include/VandermondeBuilders.hpp: computeVandermondeMatrix :45, computeGradVandermonde :65

run in the order of the reference's constructors (src/TriangleNodesProvisioner.cpp:72-77, src/QuadNodesProvisioner.cpp:67-72):
buildNodes, buildLift, buildPhysicalGrid, buildMaps.  Nothing is copied into the repository (function bodies are cut out at
build time into a generated translation unit under oracle/_ref/, git-ignored); Types.hpp, BlitzHelpers.hpp (fullToVector,
reshapeMatTo1D, uniquetol), LinAlgHelpers.hpp, Constants.hpp and BCtypes.hpp are the reference's own.

What is NOT the reference's: <blitz/array.h> (oracle/blitz_shim); the class shells around the extracted methods (members
only; MeshManager is a holder of the five tables the builders read, filled by the caller -- the tables themselves are held
to the reference's MeshManager specs in tests/test_setup_golden.py); and the LAPACK-backed bodies, replaced by plain
stand-ins because LAPACK is absent: DenseMatrixInverter::computeInverse (Gauss-Jordan), DirectSolver::solve
(Gaussian elimination), DenseCholeskyFactorizer (unused on this path) and JacobiBuilders::computeJacobiQuadWeights (an
eigenvalue problem: the caller supplies the Gauss-Jacobi points, the host mirror's, themselves held to the reference's
JacobiGQ / JacobiGL golden vectors).

Runs only where /root/reference exists (the build container); the GPU box uses the prebuilt .so.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_ref_output import OUT, REF, function_at, using_block  # noqa: E402

LAPACK_STANDINS = r"""
class DenseMatrixInverter {  // stand-in for LAPACK dgetrf/dgetri (src/DenseMatrixInverter.cpp): Gauss-Jordan with partial pivoting
 public:
  void computeInverse(const real_matrix_type& A, real_matrix_type& Ainv) const {
    const int n = A.rows();
    std::vector<double> M((size_t)n * 2 * n, 0.0);
    for (int i = 0; i < n; ++i) { for (int j = 0; j < n; ++j) M[(size_t)i * 2 * n + j] = A(i, j); M[(size_t)i * 2 * n + n + i] = 1.0; }
    for (int c = 0; c < n; ++c) {
      int piv = c;
      for (int i = c + 1; i < n; ++i) if (std::fabs(M[(size_t)i * 2 * n + c]) > std::fabs(M[(size_t)piv * 2 * n + c])) piv = i;
      if (piv != c) for (int j = 0; j < 2 * n; ++j) std::swap(M[(size_t)c * 2 * n + j], M[(size_t)piv * 2 * n + j]);
      const double d = M[(size_t)c * 2 * n + c];
      for (int j = 0; j < 2 * n; ++j) M[(size_t)c * 2 * n + j] /= d;
      for (int i = 0; i < n; ++i) if (i != c) { const double f = M[(size_t)i * 2 * n + c]; if (f != 0.0) for (int j = 0; j < 2 * n; ++j) M[(size_t)i * 2 * n + j] -= f * M[(size_t)c * 2 * n + j]; }
    }
    for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) Ainv(i, j) = M[(size_t)i * 2 * n + n + j];
  }
};
class DirectSolver {  // stand-in for LAPACK dsgesv (src/DirectSolver.cpp): X = A^-1 B by Gaussian elimination with partial pivoting
 public:
  void solve(const real_matrix_type& A, const real_matrix_type& B, real_matrix_type& X) const {
    const int n = A.rows(), m = B.cols();
    if (A.cols() != n || B.rows() != n || X.rows() != n || X.cols() != m) return;
    std::vector<double> M((size_t)n * (n + m));
    for (int i = 0; i < n; ++i) { for (int j = 0; j < n; ++j) M[(size_t)i * (n + m) + j] = A(i, j); for (int j = 0; j < m; ++j) M[(size_t)i * (n + m) + n + j] = B(i, j); }
    for (int c = 0; c < n; ++c) {
      int piv = c;
      for (int i = c + 1; i < n; ++i) if (std::fabs(M[(size_t)i * (n + m) + c]) > std::fabs(M[(size_t)piv * (n + m) + c])) piv = i;
      if (piv != c) for (int j = 0; j < n + m; ++j) std::swap(M[(size_t)c * (n + m) + j], M[(size_t)piv * (n + m) + j]);
      const double d = M[(size_t)c * (n + m) + c];
      if (d == 0.0) continue;
      for (int j = 0; j < n + m; ++j) M[(size_t)c * (n + m) + j] /= d;
      for (int i = 0; i < n; ++i) if (i != c) { const double f = M[(size_t)i * (n + m) + c]; if (f != 0.0) for (int j = 0; j < n + m; ++j) M[(size_t)i * (n + m) + j] -= f * M[(size_t)c * (n + m) + j]; }
    }
    for (int i = 0; i < n; ++i) for (int j = 0; j < m; ++j) X(i, j) = M[(size_t)i * (n + m) + n + j];
  }
};
class DenseCholeskyFactorizer { public: void computeCholesky(const real_matrix_type&, real_matrix_type&) const {} };  // not on this path
"""

SHELL = r"""// GENERATED by oracle/make_ref_setup.py -- not product source, never committed
#include "Types.hpp"
#include "BlitzHelpers.hpp"
#include "LinAlgHelpers.hpp"
#include "BCtypes.hpp"
#include "Constants.hpp"
#include <blitz/array.h>
#include <cmath>
#include <cstring>
#include <limits>
#include <map>
#include <memory>
#include <utility>
#include <vector>

namespace blitzdg {
class MeshManager {  // shell: the tables the builders read, filled by the caller
 public:
  index_vector_type EToV, EToE, EToF, BCType;
  real_vector_type Vert;
  index_type NumVerts = 0, NumElements = 0;
  const index_vector_type& get_Elements() const { return EToV; }
  const real_vector_type& get_Vertices() const { return Vert; }
  index_type get_NumVerts() const { return NumVerts; }
  index_type get_NumElements() const { return NumElements; }
  const index_vector_type& get_EToE() const { return EToE; }
  const index_vector_type& get_EToF() const { return EToF; }
  const index_vector_type& get_BCType() const { return BCType; }
};
""" + LAPACK_STANDINS + r"""
class JacobiBuilders {
 public:
  // stand-in for the LAPACK eigenproblem of computeJacobiQuadWeights: Gauss-Jacobi points and weights supplied by the caller, keyed by N
  mutable std::map<int, std::pair<std::vector<double>, std::vector<double>>> gq;
  void computeJacobiPolynomial(real_vector_type const & x, real_type alpha, real_type beta, index_type N, real_vector_type & p) const;
  void computeGradJacobi(real_vector_type const & x, real_type alpha, real_type beta, index_type N, real_vector_type& dp) const;
  void computeGaussLobottoPoints(real_type alpha, real_type beta, index_type N, real_vector_type& x) const;
  void computeJacobiQuadWeights(real_type, real_type, index_type N, real_vector_type& x, real_vector_type& w) const {
    const auto& t = gq.at((int)N);
    for (int i = 0; i <= N; ++i) { x(i) = t.first[i]; w(i) = t.second[i]; }
  }
};
class VandermondeBuilders {
  JacobiBuilders Jacobi;
  DenseMatrixInverter Inverter;
 public:
  void computeVandermondeMatrix(const real_vector_type & r, real_matrix_type& V, real_matrix_type& Vinv, bool includeInverse=true) const;
  void computeGradVandermonde(const real_vector_type & r, real_matrix_type& DVr) const;
};
class TriangleNodesProvisioner {  // shell: the members of the reference's include/TriangleNodesProvisioner.hpp the extracted methods use
 public:
  static const index_type NumFaces;
  static const real_type NodeTol;
  index_type NumElements, NOrder, NumLocalPoints, NumFacePoints;
  std::unique_ptr<real_matrix_type> xGrid, yGrid;
  std::unique_ptr<real_vector_type> rGrid, sGrid;
  std::unique_ptr<real_matrix_type> V, Dr, Ds, Drw, Dsw, Lift, J, rx, sx, ry, sy, nx, ny, Vinv, Filter;
  std::unique_ptr<index_matrix_type> Fmask;
  std::unique_ptr<real_matrix_type> Fscale, Fx, Fy;
  std::unique_ptr<index_vector_type> vmapM, vmapP, mapP, vmapB, mapB;
  std::unique_ptr<index_hashmap> BCmap;
  std::unique_ptr<std::vector<index_type>> gather, scatter;
  MeshManager Mesh2D;
  JacobiBuilders Jacobi;
  VandermondeBuilders Vandermonde;
  DirectSolver LinSolver;
  DenseMatrixInverter Inverter;
  DenseCholeskyFactorizer CholeskyFactorizer;
  void buildNodes(); void buildLift(); void buildPhysicalGrid(); void buildMaps(); void buildBCHash(); void buildBCHash(const index_vector_type& bcType);
  void buildFilter(real_type Nc, index_type s);
  void computeEquilateralNodes(real_vector_type & x, real_vector_type & y) const;
  void computeWarpFactor(const real_vector_type & r, real_vector_type & warpFactor) const;
  void xyTors(const real_vector_type & x, const real_vector_type & y, real_vector_type & r, real_vector_type & s) const;
  void evaluateSimplexPolynomial(const real_vector_type & a, const real_vector_type & b, index_type i, index_type j, real_vector_type & p) const;
  void evaluateGradSimplex(const real_vector_type & a, const real_vector_type & b, index_type id, index_type jd, real_vector_type & dpdr, real_vector_type & dpds) const;
  void rsToab(const real_vector_type & r, const real_vector_type & s, real_vector_type & a, real_vector_type & b) const;
  void computeVandermondeMatrix(index_type N, const real_vector_type & r, const real_vector_type & s, real_matrix_type & V) const;
  void computeGradVandermondeMatrix(index_type N,  const real_vector_type & r, const real_vector_type & s, real_matrix_type & V2Dr, real_matrix_type & V2Ds) const;
  void computeDifferentiationMatrices(const real_matrix_type & V2Dr, const real_matrix_type & V2Ds, const real_matrix_type & V, const real_matrix_type & Vc, real_matrix_type & Dr, real_matrix_type & Ds, real_matrix_type& Drw, real_matrix_type& Dsw) const;
};
const index_type TriangleNodesProvisioner::NumFaces = 3;
const real_type TriangleNodesProvisioner::NodeTol = 1.e-5;
class QuadNodesProvisioner {  // shell: the members of the reference's include/QuadNodesProvisioner.hpp the extracted methods use
 public:
  static const index_type NumFaces;
  static const real_type NodeTol;
  index_type NumElements, NOrder, NumLocalPoints, NumFacePoints;
  std::unique_ptr<real_matrix_type> xGrid, yGrid;
  std::unique_ptr<real_vector_type> rGrid, sGrid;
  std::unique_ptr<real_matrix_type> V, Dr, Ds, Drw, Dsw, Lift, J, rx, sx, ry, sy, nx, ny, Vinv, Filter;
  std::unique_ptr<index_matrix_type> Fmask;
  std::unique_ptr<real_matrix_type> Fscale;
  std::unique_ptr<index_vector_type> vmapM, vmapP, mapP, vmapB, mapB;
  std::unique_ptr<index_hashmap> BCmap;
  std::unique_ptr<std::vector<index_type>> gather, scatter;
  MeshManager Mesh2D;
  JacobiBuilders Jacobi;
  VandermondeBuilders Vandermonde;
  DirectSolver LinSolver;
  DenseMatrixInverter Inverter;
  void buildNodes(); void buildLift(); void buildPhysicalGrid(); void buildMaps(); void buildBCHash(); void buildBCHash(const index_vector_type& bcType);
  void buildFilter(real_type Nc, index_type s);
  void computeVandermondeMatrix(index_type N, const real_vector_type & r, const real_vector_type & s, real_matrix_type & V) const;
  void computeGradVandermondeMatrix(index_type N,  const real_vector_type & r, const real_vector_type & s, real_matrix_type & V2Dr, real_matrix_type & V2Ds) const;
  void computeDifferentiationMatrices(const real_matrix_type & V2Dr, const real_matrix_type & V2Ds, const real_matrix_type & V, real_matrix_type & Dr, real_matrix_type & Ds, real_matrix_type& Drw, real_matrix_type& Dsw) const;
};
const index_type QuadNodesProvisioner::NumFaces = 4;
const real_type QuadNodesProvisioner::NodeTol = 1.e-5;
const real_type pi = blitzdg::constants::pi;
}
using std::tgamma;
using std::hypot;
"""

WRAP_COMMON = r"""
namespace {
using namespace blitzdg;
void put(double* dst, const real_matrix_type& A) { for (int i = 0; i < A.rows(); ++i) for (int j = 0; j < A.cols(); ++j) dst[(size_t)i * A.cols() + j] = A(i, j); }
void puti(int* dst, const index_matrix_type& A) { for (int i = 0; i < A.rows(); ++i) for (int j = 0; j < A.cols(); ++j) dst[(size_t)i * A.cols() + j] = A(i, j); }
void putv(double* dst, const real_vector_type& v) { for (int i = 0; i < v.rows(); ++i) dst[i] = v(i); }
void putiv(int* dst, const index_vector_type& v) { for (int i = 0; i < v.rows(); ++i) dst[i] = v(i); }
void fill_mesh(MeshManager& m, int Nf, int K, int Nv, const int* EToV, const double* Vert, const int* EToE, const int* EToF, const int* BCType) {
  m.NumElements = K; m.NumVerts = Nv;
  m.EToV.resize(Nf * K); m.EToE.resize(Nf * K); m.EToF.resize(Nf * K); m.BCType.resize(Nf * K); m.Vert.resize(3 * Nv);
  for (int i = 0; i < Nf * K; ++i) { m.EToV(i) = EToV[i]; m.EToE(i) = EToE[i]; m.EToF(i) = EToF[i]; m.BCType(i) = BCType[i]; }
  for (int i = 0; i < 3 * Nv; ++i) m.Vert(i) = Vert[i];
}
void fill_gq(JacobiBuilders& J, int n, const double* x, const double* w) {  // the n interior Gauss-Jacobi(1,1) points of the order-(n+1) GLL set
  if (n > 0) J.gq[n - 1] = std::make_pair(std::vector<double>(x, x + n), std::vector<double>(w, w + n));
}
// BCmap: for every tag 1..8 the list of face-node ids in the order the reference inserted them; out = [count_1, ids..., count_2, ids..., ...]
int put_bcmap(int* out, const index_hashmap& bc) {
  int pos = 0;
  for (int tag = 1; tag <= 8; ++tag) {
    auto it = bc.find(tag);
    const int n = it == bc.end() ? 0 : (int)it->second.size();
    out[pos++] = n;
    for (int i = 0; i < n; ++i) out[pos++] = it->second[i];
  }
  return pos;
}
}
"""

WRAP_TRI = r"""
struct ref_tri_out {  // every array row-major in the reference's shapes
  double *r, *s;                       // [Np]
  int *Fmask;                          // [Nfp][3]
  double *V, *Vinv, *Dr, *Ds, *Lift;   // [Np][Np] x4, [Np][3 Nfp]
  double *x, *y, *J, *rx, *sx, *ry, *sy;  // [Np][K]
  double *nx, *ny, *Fscale;            // [3 Nfp][K]
  int *vmapM, *vmapP, *mapP;           // [3 Nfp K]
  int *mapB, *vmapB; int nB;           // boundary face nodes (capacity 3 Nfp K)
  int *bcmap;                          // put_bcmap layout (capacity 8 + 3 Nfp K)
  double *Filter;                      // [Np][Np] or NULL
};
extern "C" int ref_provision_triangles(int N, int K, int Nv, const int* EToV, const double* Vert, const int* EToE, const int* EToF, const int* BCType,
                                       int ngq, const double* gqx, const double* gqw, double filtNc, int filtS, ref_tri_out* o) {
  TriangleNodesProvisioner p;
  const int Np = (N + 1) * (N + 2) / 2, Nfp = N + 1, Nfn = 3 * Nfp;
  p.NumElements = K; p.NOrder = N; p.NumLocalPoints = Np; p.NumFacePoints = Nfp;
  fill_mesh(p.Mesh2D, 3, K, Nv, EToV, Vert, EToE, EToF, BCType);
  fill_gq(p.Jacobi, ngq, gqx, gqw);
  // allocations of the reference's constructor (src/TriangleNodesProvisioner.cpp:39-71)
  p.xGrid.reset(new real_matrix_type(Np, K)); p.yGrid.reset(new real_matrix_type(Np, K));
  p.rGrid.reset(new real_vector_type(Np)); p.sGrid.reset(new real_vector_type(Np));
  p.V.reset(new real_matrix_type(Np, Np)); p.Dr.reset(new real_matrix_type(Np, Np)); p.Ds.reset(new real_matrix_type(Np, Np));
  p.Drw.reset(new real_matrix_type(Np, Np)); p.Dsw.reset(new real_matrix_type(Np, Np)); p.Lift.reset(new real_matrix_type(Np, Nfn));
  p.J.reset(new real_matrix_type(Np, K)); p.rx.reset(new real_matrix_type(Np, K)); p.sx.reset(new real_matrix_type(Np, K));
  p.ry.reset(new real_matrix_type(Np, K)); p.sy.reset(new real_matrix_type(Np, K));
  p.nx.reset(new real_matrix_type(Nfn, K)); p.ny.reset(new real_matrix_type(Nfn, K));
  p.Vinv.reset(new real_matrix_type(Np, Np)); p.Filter.reset(new real_matrix_type(Np, Np));
  p.Fmask.reset(new index_matrix_type(Nfp, 3));
  p.Fscale.reset(new real_matrix_type(Nfn, K)); p.Fx.reset(new real_matrix_type(Nfn, K)); p.Fy.reset(new real_matrix_type(Nfn, K));
  p.vmapM.reset(new index_vector_type(Nfn * K)); p.vmapP.reset(new index_vector_type(Nfn * K)); p.mapP.reset(new index_vector_type(Nfn * K));
  p.BCmap.reset(new index_hashmap());
  // ... and its body (:72-77)
  p.buildNodes();
  p.buildLift();
  p.buildPhysicalGrid();
  p.buildMaps();
  putv(o->r, *p.rGrid); putv(o->s, *p.sGrid); puti(o->Fmask, *p.Fmask);
  put(o->V, *p.V); put(o->Vinv, *p.Vinv); put(o->Dr, *p.Dr); put(o->Ds, *p.Ds); put(o->Lift, *p.Lift);
  put(o->x, *p.xGrid); put(o->y, *p.yGrid); put(o->J, *p.J); put(o->rx, *p.rx); put(o->sx, *p.sx); put(o->ry, *p.ry); put(o->sy, *p.sy);
  put(o->nx, *p.nx); put(o->ny, *p.ny); put(o->Fscale, *p.Fscale);
  putiv(o->vmapM, *p.vmapM); putiv(o->vmapP, *p.vmapP); putiv(o->mapP, *p.mapP);
  o->nB = p.mapB->rows(); putiv(o->mapB, *p.mapB); putiv(o->vmapB, *p.vmapB);
  put_bcmap(o->bcmap, *p.BCmap);
  if (o->Filter) { p.buildFilter(filtNc, filtS); put(o->Filter, *p.Filter); }
  return 0;
}
"""

WRAP_QUAD = r"""
extern "C" int ref_provision_quads(int N, int K, int Nv, const int* EToV, const double* Vert, const int* EToE, const int* EToF, const int* BCType,
                                   int ngq, const double* gqx, const double* gqw, double filtNc, int filtS, ref_tri_out* o) {
  QuadNodesProvisioner p;  // same output struct; Fmask [Nfp][4], Lift [Np][4 Nfp], face arrays [4 Nfp][K]
  const int Np = (N + 1) * (N + 1), Nfp = N + 1, Nfn = 4 * Nfp;
  p.NumElements = K; p.NOrder = N; p.NumLocalPoints = Np; p.NumFacePoints = Nfp;
  fill_mesh(p.Mesh2D, 4, K, Nv, EToV, Vert, EToE, EToF, BCType);
  fill_gq(p.Jacobi, ngq, gqx, gqw);
  // allocations of the reference's constructor (src/QuadNodesProvisioner.cpp:36-66)
  p.xGrid.reset(new real_matrix_type(Np, K)); p.yGrid.reset(new real_matrix_type(Np, K));
  p.rGrid.reset(new real_vector_type(Np)); p.sGrid.reset(new real_vector_type(Np));
  p.V.reset(new real_matrix_type(Np, Np)); p.Dr.reset(new real_matrix_type(Np, Np)); p.Ds.reset(new real_matrix_type(Np, Np));
  p.Drw.reset(new real_matrix_type(Np, Np)); p.Dsw.reset(new real_matrix_type(Np, Np)); p.Lift.reset(new real_matrix_type(Np, Nfn));
  p.J.reset(new real_matrix_type(Np, K)); p.rx.reset(new real_matrix_type(Np, K)); p.sx.reset(new real_matrix_type(Np, K));
  p.ry.reset(new real_matrix_type(Np, K)); p.sy.reset(new real_matrix_type(Np, K));
  p.nx.reset(new real_matrix_type(Nfn, K)); p.ny.reset(new real_matrix_type(Nfn, K));
  p.Vinv.reset(new real_matrix_type(Np, Np)); p.Filter.reset(new real_matrix_type(Np, Np));
  p.Fmask.reset(new index_matrix_type(Nfp, 4));
  p.Fscale.reset(new real_matrix_type(Nfn, K));
  p.vmapM.reset(new index_vector_type(Nfn * K)); p.vmapP.reset(new index_vector_type(Nfn * K)); p.mapP.reset(new index_vector_type(Nfn * K));
  p.BCmap.reset(new index_hashmap());
  // ... and its body (:67-72)
  p.buildNodes();
  p.buildLift();
  p.buildPhysicalGrid();
  p.buildMaps();
  putv(o->r, *p.rGrid); putv(o->s, *p.sGrid); puti(o->Fmask, *p.Fmask);
  put(o->V, *p.V); put(o->Vinv, *p.Vinv); put(o->Dr, *p.Dr); put(o->Ds, *p.Ds); put(o->Lift, *p.Lift);
  put(o->x, *p.xGrid); put(o->y, *p.yGrid); put(o->J, *p.J); put(o->rx, *p.rx); put(o->sx, *p.sx); put(o->ry, *p.ry); put(o->sy, *p.sy);
  put(o->nx, *p.nx); put(o->ny, *p.ny); put(o->Fscale, *p.Fscale);
  putiv(o->vmapM, *p.vmapM); putiv(o->vmapP, *p.vmapP); putiv(o->mapP, *p.mapP);
  o->nB = p.mapB->rows(); putiv(o->mapB, *p.mapB); putiv(o->vmapB, *p.vmapB);
  put_bcmap(o->bcmap, *p.BCmap);
  if (o->Filter) { p.buildFilter(filtNc, filtS); put(o->Filter, *p.Filter); }
  return 0;
}
"""


def main():
    tnp = os.path.join(REF, "src", "TriangleNodesProvisioner.cpp")
    qnp = os.path.join(REF, "src", "QuadNodesProvisioner.cpp")
    jac = os.path.join(REF, "src", "JacobiBuilders.cpp")
    vdm = os.path.join(REF, "include", "VandermondeBuilders.hpp")
    if not (os.path.exists(tnp) and os.path.exists(jac)):
        print("make_ref_setup: reference sources not found under", REF, "-- nothing built")
        return 0
    os.makedirs(OUT, exist_ok=True)

    def method(path, line, name, cls):
        """A method defined inside a class body (header): re-emitted as an out-of-class definition."""
        txt = function_at(path, line, name)
        head, body = txt.split("\n", 1)
        sig_end = body.index("{")
        sig = body[:sig_end].replace("bool includeInverse=true", "bool includeInverse").replace(name + "(", cls + "::" + name + "(", 1)
        return head + "\n" + sig + body[sig_end:]

    T, Q = "TriangleNodesProvisioner::", "QuadNodesProvisioner::"
    tu = (SHELL + using_block(tnp, 20, 32) + "namespace blitzdg {\n"
          + function_at(jac, 18, "JacobiBuilders::computeJacobiPolynomial")
          + function_at(jac, 98, "JacobiBuilders::computeGaussLobottoPoints")
          + function_at(jac, 117, "JacobiBuilders::computeGradJacobi")
          + method(vdm, 45, "computeVandermondeMatrix", "VandermondeBuilders")
          + method(vdm, 65, "computeGradVandermonde", "VandermondeBuilders")
          + "".join(function_at(tnp, ln, T + nm) for ln, nm in (
              (383, "evaluateSimplexPolynomial"), (395, "rsToab"), (406, "xyTors"), (418, "computeVandermondeMatrix"),
              (436, "computeGradVandermondeMatrix"), (456, "computeDifferentiationMatrices"), (515, "buildFilter"),
              (549, "computeEquilateralNodes"), (604, "computeWarpFactor"), (642, "evaluateGradSimplex"), (678, "buildNodes"),
              (730, "buildPhysicalGrid"), (895, "buildMaps"), (1022, "buildBCHash()"), (1028, "buildBCHash(const"), (1060, "buildLift")))
          + "".join(function_at(qnp, ln, Q + nm) for ln, nm in (
              (76, "computeVandermondeMatrix"), (94, "computeGradVandermondeMatrix"), (114, "computeDifferentiationMatrices"),
              (170, "buildFilter"), (204, "buildNodes"), (277, "buildPhysicalGrid"), (450, "buildMaps"), (577, "buildBCHash()"),
              (583, "buildBCHash(const"), (615, "buildLift")))
          + "}\n" + WRAP_COMMON + WRAP_TRI + WRAP_QUAD)
    path = os.path.join(OUT, "ref_setup.cpp")
    with open(path, "w") as fh:
        fh.write(tu)
    so = os.path.join(OUT, "libsetup_ref.so")
    subprocess.check_call(["g++", "-std=c++14", "-O2", "-ffp-contract=off", "-fPIC", "-w", "-shared", "-I" + os.path.join(HERE, "blitz_shim"),
                           "-I" + os.path.join(REF, "include"), path, "-o", so])
    print("built", so)
    return 0


if __name__ == "__main__":
    sys.exit(main())
