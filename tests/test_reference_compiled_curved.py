"""Cubature volume mesh and Gauss face nodes against the REFERENCE'S OWN builders, compiled verbatim
from /root/reference by oracle/make_ref_curved.py (oracle/_ref/libcurved_ref.so):
TriangleNodesProvisioner::buildCubatureVolumeMesh (src/TriangleNodesProvisioner.cpp:81-203), buildGaussFaceNodes (:207-381),
the Vandermonde / gradient / differentiation-matrix helpers they call and the tabulated rules of
include/TriangleCubatureRules.hpp.

Held to them: the host mirrors of blitzdg_b200/curved.py -- every array sw2dComputeRHS_curved and the device kernel read
(V, Dr, Ds, W, rx, sx, ry, sy, MMChol; Interp, nx, ny, sJ, W, mapM, mapP, the wall list) and the nodal metrics the cubature
builder repairs in place.  The mirror is given the reference's own tabulated points (rule=...), so the comparison is array
by array; its own product rule is checked against the tables as a quadrature (same integrals to the rule's order).
The library is prebuilt where the reference tree exists and travels with the snapshot; absent library and absent
reference => skip.  What the library does NOT take from the reference (LAPACK is absent here): the Cholesky factor loop, the
Gauss nodes and weights (the mirror's, themselves held to the reference's JacobiGQ golden vectors in test_setup_golden.py).
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import make_curved_reference_vectors as gen  # noqa: E402  (build_case only: nothing here reads /root/reference)

LIB = os.path.join(ROOT, "oracle", "_ref", "libcurved_ref.so")
_dp, _ip = C.POINTER(C.c_double), C.POINTER(C.c_int)


def _p(a):
    return a.ctypes.data_as(_dp)


def _pi(a):
    return a.ctypes.data_as(_ip)


def _c(a):
    return np.ascontiguousarray(a, dtype=np.float64)


@pytest.fixture(scope="module")
def ref():
    if not os.path.exists(LIB):
        if not os.path.exists("/root/reference/src/TriangleNodesProvisioner.cpp"):
            pytest.skip("compiled reference not built and reference tree not present")
        subprocess.run([sys.executable, os.path.join(ROOT, "oracle", "make_ref_curved.py")], check=True, capture_output=True)
    return C.CDLL(LIB)


def table_rule(ref, order):
    buf = [np.zeros(256) for _ in range(3)]
    n = ref.ref_cubature_rule(order, _p(buf[0]), _p(buf[1]), _p(buf[2]))
    return tuple(b[:n].copy() for b in buf)


def close(a, b, tol):
    return float(np.abs(np.asarray(a) - b).max()) <= tol * max(1.0, float(np.abs(b).max()))


@pytest.mark.parametrize("order", [1, 2, 5, 9, 15, 21, 28])
def test_product_rule_integrates_what_the_reference_tables_integrate(ref, order):
    """Two different point sets, the same integrals: every monomial r^i s^j of total degree <= order."""
    from blitzdg_b200.curved import triangle_cubature

    r, s, w = table_rule(ref, order)
    rr, ss, ww = triangle_cubature(order)
    assert len(rr) == (order // 2 + 1) ** 2 and abs(w.sum() - 2.0) < 1e-13 and abs(ww.sum() - 2.0) < 1e-13
    assert (r + s <= 1e-14).all() and (r >= -1 - 1e-14).all() and (s >= -1 - 1e-14).all()
    for i in range(order + 1):
        for j in range(order + 1 - i):
            assert abs((w * r**i * s**j).sum() - (ww * rr**i * ss**j).sum()) < 5e-13, (order, i, j)


def test_reference_tables_of_order_3_and_4_are_unusable_as_written(ref):
    """include/TriangleCubatureRules.hpp:34-42: the 6-point rule shared by orders 3 and 4 is written with three commas missing
    ('0.633... - 0.816...' and two line ends), so the list initialiser sees 15 values for 18 slots, two of them differences of
    neighbours; the weights then sum to -0.37 instead of the triangle's area 2.  Recorded, not mirrored: the product rule of
    blitzdg_b200/curved.py is the only rule this repo offers at those orders, and it integrates exactly."""
    from blitzdg_b200.curved import triangle_cubature

    for order in (3, 4):
        r, s, w = table_rule(ref, order)
        assert len(w) == 6 and abs(w.sum() - 2.0) > 1.0
        assert abs(triangle_cubature(order)[2].sum() - 2.0) < 1e-14


def test_stored_rule_is_the_reference_table(ref):
    """The rule committed beside the curved fixture is what the compiled header returns today."""
    fx = np.load(os.path.join(ROOT, "tests", "golden", "curved_reference_rhs.npz"))
    for name in gen.TABLE_RULE_CASES:
        r, s, w = table_rule(ref, gen.CASES[name][3])
        assert np.array_equal(fx[name + "_rule"], np.stack([r, s, w]))


@pytest.fixture(scope="module", params=[(5, 4, 2, 9, 6), (6, 5, 4, 15, 10), (3, 3, 1, 5, 3), (3, 2, 7, 21, 12)],
                ids=lambda c: "N%d-cub%d-g%d" % (c[2], c[3], c[4]))
def built(request, ref):
    nx, ny, N, ncub, ng = request.param
    rule = table_rule(ref, ncub)
    nodes, ctx, cub, gauss, curved = gen.build_case(nx, ny, N, ncub, ng, rule=rule)
    return nodes, cub, gauss, ncub, ng


def test_cubature_volume_mesh_matches_the_compiled_reference(ref, built):
    nodes, cub, gauss, ncub, ng = built
    N, Np, K, n = nodes.NOrder, nodes.NumLocalPoints, nodes.NumElements, cub.NumCubaturePoints
    V, Vinv, Dr, Ds, x, y = map(_c, (nodes.V, nodes.Vinv, nodes.Dr, nodes.Ds, nodes.xGrid, nodes.yGrid))
    cV, cDr, cDs = (np.zeros((n, Np)) for _ in range(3))
    cW, crx, csx, cry, csy = (np.zeros((n, K)) for _ in range(5))
    chol = np.zeros((Np, Np, K))
    nJ, nrx, nry, nsx, nsy = (np.zeros((Np, K)) for _ in range(5))
    got = ref.ref_build_cubature(N, K, ncub, _p(V), _p(Vinv), _p(Dr), _p(Ds), _p(x), _p(y), _p(cV), _p(cDr), _p(cDs), _p(cW), _p(crx),
                                 _p(csx), _p(cry), _p(csy), _p(chol), _p(nJ), _p(nrx), _p(nry), _p(nsx), _p(nsy))
    assert got == n
    # tolerances: rounding of sums of Np terms; the metric quotients carry the conditioning of 1/J
    for nm, mine, want, tol in (("V", cub.V, cV, 1e-14), ("Dr", cub.Dr, cDr, 1e-13), ("Ds", cub.Ds, cDs, 1e-13), ("W", cub.W, cW, 1e-14),
                                ("rx", cub.rx, crx, 1e-12), ("sx", cub.sx, csx, 1e-12), ("ry", cub.ry, cry, 1e-12), ("sy", cub.sy, csy, 1e-12),
                                ("MMChol", cub.MMChol, chol, 1e-13),
                                ("J", nodes.J, nJ, 1e-14), ("nodal rx", nodes.rx, nrx, 1e-12), ("nodal ry", nodes.ry, nry, 1e-12),
                                ("nodal sx", nodes.sx, nsx, 1e-12), ("nodal sy", nodes.sy, nsy, 1e-12)):
        assert close(mine, want, tol), (nm, float(np.abs(mine - want).max()), float(np.abs(want).max()))
    assert np.abs(np.tril(chol[:, :, 0], -1)).max() == 0.0  # upper factors, as sw2dComputeRHS_curved's two triangular solves assume


def test_gauss_face_nodes_match_the_compiled_reference(ref, built):
    from blitzdg_b200.jacobi import jacobi_gauss

    nodes, cub, gauss, ncub, ng = built
    N, Np, K = nodes.NOrder, nodes.NumLocalPoints, nodes.NumElements
    V, Vinv, Dr, Ds, x, y = map(_c, (nodes.V, nodes.Vinv, nodes.Dr, nodes.Ds, nodes.xGrid, nodes.yGrid))
    gz, gw = (_c(a) for a in jacobi_gauss(0.0, 0.0, ng - 1))
    mesh = nodes.Mesh2D
    EToE, EToF, BC = (np.ascontiguousarray(a, dtype=np.int32).reshape(-1) for a in (mesh.get_EToE(), mesh.get_EToF(), mesh.get_BCType()))
    Interp = np.zeros((3 * ng, Np))
    gnx, gny, gsJ, gW = (np.zeros((3 * ng, K)) for _ in range(4))
    mapM, mapP, wall = (np.zeros(3 * ng * K, dtype=np.int32) for _ in range(3))
    nw = ref.ref_build_gauss(N, K, ng, _p(V), _p(Vinv), _p(Dr), _p(Ds), _p(x), _p(y), _pi(EToE), _pi(EToF), _pi(BC), _p(gz), _p(gw),
                             _p(Interp), _p(gnx), _p(gny), _p(gsJ), _p(gW), _pi(mapM), _pi(mapP), _pi(wall))
    for nm, mine, want, tol in (("Interp", gauss.Interp, Interp, 1e-14), ("nx", gauss.nx, gnx, 1e-13), ("ny", gauss.ny, gny, 1e-13),
                                ("sJ", gauss.sJ, gsJ, 1e-13), ("W", gauss.W, gW, 1e-13)):
        assert close(mine, want, tol), (nm, float(np.abs(mine - want).max()), float(np.abs(want).max()))
    assert np.array_equal(gauss.mapM, mapM) and np.array_equal(gauss.mapP, mapP)
    assert sorted(gauss.BCmap) == [3] and np.array_equal(gauss.BCmap[3], wall[:nw])  # same ids in the same order
