"""Cubature volume mesh and Gauss face nodes of curved triangles: the two extra contexts the reference's
over-integrated RHS reads (swhelpers/rhs.py:6-176).

Host-side mirrors of
  TriangleNodesProvisioner::buildCubatureVolumeMesh  (reference src/TriangleNodesProvisioner.cpp:81-202)
  TriangleNodesProvisioner::buildGaussFaceNodes      (reference src/TriangleNodesProvisioner.cpp:205-381)
  CubatureContext2D / GaussFaceContext2D             (reference include/CubatureContext2D.hpp, include/GaussFaceContext2D.hpp)
with the property names pyblitzdg exposes (V, Dr, Ds, W, rx, ..., MMChol; Interp, nx, ny, W, BCmap, mapM, mapP).

One deviation, stated here and in DESIGN.md: the reference takes its cubature points from tables
(include/TriangleCubatureRules.hpp, orders 1..28).  Those tables are data of the reference and are not copied; this
mirror uses the Gauss-Legendre x Gauss-Jacobi(1,0) product rule on the collapsed square, exact to the same total
degree with more points (64 instead of 54 at order 15).  sw2dComputeRHS_curved only sees the resulting arrays, so the
device kernel and the oracle are independent of which rule produced them.
A caller who holds the reference's points passes them as build_cubature_volume_mesh(..., rule=(r, s, w)); with them every
array built here agrees with the reference's compiled builders (tests/test_reference_compiled_curved.py).
"""
from dataclasses import dataclass, field
from typing import Dict

import numpy as np

from .jacobi import jacobi_gauss


@dataclass
class CubatureContext2D:
    NCubature: int
    NumCubaturePoints: int
    r: np.ndarray
    s: np.ndarray
    w: np.ndarray
    V: np.ndarray       # [Ncub][Np]   nodal values -> cubature points
    rx: np.ndarray      # [Ncub][K]
    sx: np.ndarray
    ry: np.ndarray
    sy: np.ndarray
    J: np.ndarray
    Dr: np.ndarray      # [Ncub][Np]   d/dr of the nodal basis at the cubature points
    Ds: np.ndarray
    MM: np.ndarray      # [Np][Np][K]  element mass matrices V^T diag(w J) V
    MMChol: np.ndarray  # [Np][Np][K]  their upper Cholesky factors
    x: np.ndarray
    y: np.ndarray
    W: np.ndarray       # [Ncub][K]    w * J


@dataclass
class GaussFaceContext2D:
    NGauss: int
    nx: np.ndarray      # [3*NGauss][K]
    ny: np.ndarray
    sJ: np.ndarray
    Jac: np.ndarray
    rx: np.ndarray
    ry: np.ndarray
    sx: np.ndarray
    sy: np.ndarray
    BCmap: Dict[int, np.ndarray] = field(default_factory=dict)
    x: np.ndarray = None
    y: np.ndarray = None
    W: np.ndarray = None       # [3*NGauss][K]  Gauss weight * surface Jacobian
    Interp: np.ndarray = None  # [3*NGauss][Np] nodal values -> Gauss face nodes
    mapM: np.ndarray = None    # [3*NGauss*K] int32, g + 3*NGauss*k
    mapP: np.ndarray = None

    @property
    def J(self):  # the name pyblitzdg gives it (src/pyblitzdg/pyblitzdg.cpp:130)
        return self.Jac


def triangle_cubature(order):
    """Points (r, s) and weights on the reference triangle {r, s >= -1, r + s <= 0} (area 2), exact for polynomials of
    total degree <= order: n = order//2 + 1 Gauss-Legendre points in a times n Gauss-Jacobi(1,0) points in b on the
    collapsed square, r = (1 + a)(1 - b)/2 - 1, s = b, dr ds = (1 - b)/2 da db."""
    n = int(order) // 2 + 1
    a, wa = jacobi_gauss(0.0, 0.0, n - 1)
    b, wb = jacobi_gauss(1.0, 0.0, n - 1)
    A, B = np.meshgrid(a, b, indexing="ij")
    r = 0.5 * (1.0 + A) * (1.0 - B) - 1.0
    w = 0.5 * np.outer(wa, wb)
    return r.reshape(-1), B.reshape(-1).copy(), w.reshape(-1)


def set_coordinates(nodes, x, y):
    """TriangleNodesProvisioner::setCoordinates: replaces the physical nodes (curved elements); the nodal metrics are
    repaired by build_cubature_volume_mesh as in the reference (:123-147)."""
    nodes.xGrid = np.ascontiguousarray(x, dtype=np.float64)
    nodes.yGrid = np.ascontiguousarray(y, dtype=np.float64)


def build_cubature_volume_mesh(nodes, NCubature, rule=None):
    """rule: (r, s, w) of the cubature points to use instead of this mirror's product rule -- e.g. the reference's own table
    of that order, which a caller holding pyblitzdg (or the tests, through the compiled TriangleCubatureRules.hpp) can supply."""
    Np, K = nodes.NumLocalPoints, nodes.NumElements
    rc, sc, wc = triangle_cubature(NCubature) if rule is None else (np.asarray(a, dtype=np.float64) for a in rule)
    Ncub = len(rc)
    Vcub = nodes.computeInterpMatrix(rc, sc)
    V2Dr, V2Ds = nodes.computeGradVandermondeMatrix(nodes.NOrder, rc, sc)
    Drc, Dsc = np.ascontiguousarray(V2Dr @ nodes.Vinv), np.ascontiguousarray(V2Ds @ nodes.Vinv)  # (:456-488) Dr = V2Dr V^-1
    x, y = nodes.xGrid, nodes.yGrid
    xr, xs, yr, ys = Drc @ x, Dsc @ x, Drc @ y, Dsc @ y
    Jc = -xs * yr + xr * ys
    rxc, sxc, ryc, syc = ys / Jc, -yr / Jc, -xs / Jc, xr / Jc
    # nodal metrics from the (possibly moved) nodes, as the reference repairs them here
    nxr, nyr, nxs, nys = nodes.Dr @ x, nodes.Dr @ y, nodes.Ds @ x, nodes.Ds @ y
    Jn = nxr * nys - nxs * nyr
    nodes.J = Jn
    nodes.rx, nodes.ry, nodes.sx, nodes.sy = nys / Jn, -nxs / Jn, -nyr / Jn, nxr / Jn
    W = wc[:, None] * Jc
    MM = np.empty((Np, Np, K))
    MMChol = np.zeros((Np, Np, K))
    for k in range(K):
        M = Vcub.T @ (W[:, k, None] * Vcub)
        MM[:, :, k] = M
        MMChol[:, :, k] = np.linalg.cholesky(M).T  # upper factor: M = U^T U
    return CubatureContext2D(int(NCubature), Ncub, rc, sc, wc, np.ascontiguousarray(Vcub), rxc, sxc, ryc, syc, Jc, Drc, Dsc,
                             MM, MMChol, Vcub @ x, Vcub @ y, W)


def build_gauss_face_nodes(nodes, NGauss):
    Np, K, Nf = nodes.NumLocalPoints, nodes.NumElements, 3
    mesh = nodes.Mesh2D
    z, w = jacobi_gauss(0.0, 0.0, NGauss - 1)
    one = np.ones_like(z)
    face_rs = [(z, -one), (-z, z), (-one, -z)]
    Interp = np.vstack([nodes.computeVandermondeMatrix(nodes.NOrder, r, s) @ nodes.Vinv for r, s in face_rs])
    Ng3 = Nf * NGauss
    x, y = nodes.xGrid, nodes.yGrid
    gxr, gxs = (Interp @ nodes.Dr) @ x, (Interp @ nodes.Ds) @ x
    gyr, gys = (Interp @ nodes.Dr) @ y, (Interp @ nodes.Ds) @ y
    Jac = gxr * gys - gxs * gyr
    rx, ry, sx, sy = gys / Jac, -gxs / Jac, -gyr / Jac, gxr / Jac
    nx, ny = np.empty_like(Jac), np.empty_like(Jac)
    f1, f2, f3 = slice(0, NGauss), slice(NGauss, 2 * NGauss), slice(2 * NGauss, 3 * NGauss)
    nx[f1], ny[f1] = -sx[f1], -sy[f1]
    nx[f2], ny[f2] = rx[f2] + sx[f2], ry[f2] + sy[f2]
    nx[f3], ny[f3] = -rx[f3], -ry[f3]
    sJ = np.sqrt(nx * nx + ny * ny)
    nx, ny = nx / sJ, ny / sJ
    sJ = sJ * Jac
    # maps: node g of element k has id g + 3*NGauss*k; across a face the neighbour's Gauss nodes run backwards (:335-349)
    mapM = np.arange(Ng3 * K, dtype=np.int64)
    mapP = mapM.copy()
    EToE = np.asarray(mesh.get_EToE()).reshape(K, Nf)
    EToF = np.asarray(mesh.get_EToF()).reshape(K, Nf)
    bc = np.asarray(mesh.get_BCType()).reshape(K, Nf)
    lists = {tag: [] for tag in range(1, 9)}
    ig = np.arange(NGauss)
    for f in range(Nf):
        for k in range(K):
            k2, f2 = int(EToE[k, f]), int(EToF[k, f])
            ids1 = f * NGauss + ig + k * Ng3
            if k2 != k:
                mapP[ids1] = NGauss * (f2 + 1) - ig - 1 + k2 * Ng3
            else:
                tag = int(bc[k, f])
                if tag in lists:
                    lists[tag].extend(ids1.tolist())
    W = np.tile(w, Nf)[:, None] * sJ
    BCmap = {tag: np.asarray(v, dtype=np.int32) for tag, v in lists.items() if v}
    return GaussFaceContext2D(int(NGauss), nx, ny, sJ, Jac, rx, ry, sx, sy, BCmap, Interp @ x, Interp @ y, W,
                              np.ascontiguousarray(Interp), mapM.astype(np.int32), mapP.astype(np.int32))
