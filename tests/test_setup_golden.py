"""Host setup (MeshManager / Triangle|QuadNodesProvisioner) against the reference's own golden vectors.

Golden data: tests/golden/reference_vectors.json, extracted by tests/golden/extract_golden.py from
/root/reference/src/test/{MeshManager,TriangleNodesProvisioner,QuadNodesProvisioner}Tests.cpp with the
tolerances those specs use.
"""
import numpy as np
import pytest

from blitzdg_b200 import MeshManager, QuadNodesProvisioner, TriangleNodesProvisioner, BCTag
from blitzdg_b200 import jacobi

EPSF = 5.8e-5  # reference: epsf = 5.e2*float-eps-ish (TriangleNodesProvisionerTests.cpp)


def test_can_read_gmsh_mesh(golden, coarse_box_mesh):
    m = coarse_box_mesh
    assert m.get_NumVerts() == 29 and m.get_NumElements() == 40
    assert np.abs(m.get_Vertices() - np.array(golden["mesh_expectedVerts"])).max() < 3.4e-7
    assert np.array_equal(m.get_Elements(), np.array(golden["mesh_expectedElements"], dtype=np.int32))
    assert np.array_equal(m.get_EToE(), np.array(golden["mesh_expectedEToE"], dtype=np.int32))
    assert np.array_equal(m.get_EToF(), np.array(golden["mesh_expectedEToF"], dtype=np.int32))
    assert np.array_equal(m.get_BCType(), np.array(golden["mesh_expectedBcTable"], dtype=np.int32))


def test_triangle_lift_and_derivative_matrices(golden, coarse_box_mesh):
    p = TriangleNodesProvisioner(3, coarse_box_mesh)
    assert np.linalg.norm(p.get_Lift() - np.array(golden["tri_Lift_expected"]).reshape(10, 12)) < EPSF
    assert np.linalg.norm(p.get_Dr() - np.array(golden["tri_expectedDr"]).reshape(10, 10)) < EPSF
    assert np.linalg.norm(p.get_Ds() - np.array(golden["tri_expectedDs"]).reshape(10, 10)) < EPSF


def test_triangle_equilateral_nodes(golden, coarse_box_mesh):
    p = TriangleNodesProvisioner(3, coarse_box_mesh, build=False)
    x, y = p.computeEquilateralNodes()
    eps = 50 * np.finfo(float).eps
    assert np.abs(x - np.array(golden["tri_expectedx"])).max() < eps
    assert np.abs(y - np.array(golden["tri_expectedy"])).max() < eps


def test_triangle_known_answers(coarse_box_mesh):
    """Scalar known answers of TriangleNodesProvisionerTests.cpp:69-130,163-177."""
    p = TriangleNodesProvisioner(3, coarse_box_mesh, build=False)
    eps = 50 * np.finfo(float).eps
    pv = p.evaluateSimplexPolynomial(np.array([.1, .2, .3]), np.array([.2, .3, .4]), 1, 2)
    assert np.abs(pv - [0.133252242007405, 0.355359724434270, 0.637112282097905]).max() < 1e-13
    a, b = p.rsToab(np.array([-.1, .1, .2]), np.array([.2, .3, .5]))
    assert np.abs(a - [1.25, 2.14285714285714, 3.8]).max() < 1e-13
    r, s = p.xyTors(np.array([0.5, 0.6, 0.7]), np.array([0.2, 0.3, 0.4]))
    assert np.abs(r - [0.0511966128287416, 0.0934615859097789, 0.1357265589908162]).max() < eps * 10
    assert np.abs(s - [-0.1023932256574831, 0.0130768281804420, 0.1285468820183672]).max() < eps * 10
    w = p.computeWarpFactor(np.array([-.1, .1, .2]))
    assert np.abs(w - [-0.0384345884812357, 0.0384345884812359, 0.0768691769624717]).max() < 1e-13


def test_triangle_volume_to_face_maps(golden, coarse_box_mesh):
    p = TriangleNodesProvisioner(3, coarse_box_mesh)
    i32 = lambda k: np.array(golden[k], dtype=np.int32)
    assert np.array_equal(p.get_vmapM(), i32("tri_expectedvmapM"))
    assert np.array_equal(p.get_vmapP(), i32("tri_expectedvmapP"))
    assert np.array_equal(p.get_vmapB(), i32("tri_expectedvmapB"))
    assert np.array_equal(p.get_mapB(), i32("tri_expectedmapB"))
    assert np.array_equal(p.get_bcMap()[BCTag.Wall], i32("tri_expectedmapW"))


def test_triangle_fmask_n3(coarse_box_mesh):
    p = TriangleNodesProvisioner(3, coarse_box_mesh, build=False)
    p.buildNodes()
    assert p.Fmask.T.tolist() == [[0, 1, 2, 3], [3, 6, 8, 9], [0, 4, 7, 9]]


def test_quad_nodes_fmask_gradv_lift(golden, coarse_quads_mesh):
    p = QuadNodesProvisioner(3, coarse_quads_mesh)
    assert np.abs(p.rGrid - np.array(golden["quad_rExpected"])).max() < 1e-3
    assert np.abs(p.sGrid - np.array(golden["quad_sExpected"])).max() < 1e-3
    assert np.array_equal(p.Fmask.reshape(-1), np.array(golden["quad_FmExpected"], dtype=np.int32))
    V2Dr, V2Ds = p.computeGradVandermondeMatrix(3, p.rGrid, p.sGrid)
    assert np.abs(V2Dr - np.array(golden["quad_V2DrExpected"]).reshape(16, 16)).max() < 1e-4
    assert np.abs(V2Ds - np.array(golden["quad_V2DsExpected"]).reshape(16, 16)).max() < 1e-4
    assert np.abs(p.get_Lift() - np.array(golden["quad_liftExpected"]).reshape(16, 16)).max() < 1e-3


def test_gll_and_jacobi_known_answers():
    """JacobiBuildersTests.cpp known answers (GLL N=3 and orthonormal Legendre values)."""
    x = jacobi.jacobi_gauss_lobatto(0.0, 0.0, 3)
    assert np.abs(x - [-1, -0.447213595499958, 0.447213595499958, 1]).max() < 1e-14
    p = jacobi.jacobi_polynomial(np.array([0.0, 1.0]), 0.0, 0.0, 2)
    assert np.abs(p - np.sqrt(2.5) * np.array([-0.5, 1.0])).max() < 1e-14


def test_derivative_exactness_and_geometry(coarse_box_mesh):
    """Dr/Ds differentiate polynomials of degree <= N exactly; metrics reproduce d/dx, d/dy."""
    for N in (1, 2, 4, 8):
        p = TriangleNodesProvisioner(N, coarse_box_mesh)
        x, y = p.get_xGrid(), p.get_yGrid()
        f = (1 + x) ** N + x * y ** (N - 1) if N > 1 else 1 + 2 * x - y
        dfdx = N * (1 + x) ** (N - 1) + (y ** (N - 1) if N > 1 else 1.0)
        num = p.rx * (p.Dr @ f) + p.sx * (p.Ds @ f)
        assert np.abs(num - dfdx).max() < 1e-9 * max(1.0, np.abs(dfdx).max())
        assert (p.J > 0).all()
        assert np.allclose(p.nx ** 2 + p.ny ** 2, 1.0)


def test_quads_ccw_repair_and_maps(coarse_quads_mesh):
    p = QuadNodesProvisioner(4, coarse_quads_mesh)
    assert (p.J > 0).all()
    xf, yf = p.xGrid.T.reshape(-1), p.yGrid.T.reshape(-1)
    assert np.abs(xf[p.vmapM] - xf[p.vmapP]).max() < 1e-12
    assert np.abs(yf[p.vmapM] - yf[p.vmapP]).max() < 1e-12
    assert len(p.get_bcMap()[BCTag.Wall]) == 16 * 5


def test_partition_map_contract():
    """Can_Partition_A_Mesh (MeshManagerTests.cpp:230-258) on input/2box.{V,E2V}: the spec reads the two files, calls
    partitionMesh(2) -- METIS_PartMeshNodal with the options of MeshManager.cpp:498-514 and the time-dependent seed -1 --
    and expects epMap = (1, 0) and every vertex assigned.  Here through the same METIS (libbdg_metis.so)."""
    import os
    import tempfile

    with tempfile.TemporaryDirectory() as d:
        with open(os.path.join(d, "2box.V"), "w") as fh:
            fh.write("0 0\n0.5 0\n1 0\n1 1\n0.5 1\n0 1\n")
        with open(os.path.join(d, "2box.E2V"), "w") as fh:
            fh.write("0 1 4 5\n1 2 3 4\n")
        m = MeshManager()
        m.readVertices(os.path.join(d, "2box.V"))
        m.readElements(os.path.join(d, "2box.E2V"))
    assert m.get_NumVerts() == 6 and m.get_NumElements() == 2 and m.get_NumFaces() == 4
    ep = m.partitionMesh(2, method="metis")
    assert m.PartitionInfo["method"] == "METIS_PartMeshNodal"
    # The spec's literal epMap = (1, 0) is the labelling of the reference's METIS build (master, manylinux2010/Dockerfile:124-130);
    # the toolkit's 5.1.0 archive returns the mirrored labelling (0, 1) for every seed.  Part labels carry no meaning:
    # the contract is "the two boxes land in different parts".
    assert sorted(ep.tolist()) == [0, 1]
    assert (m.get_VertexPartitionMap() > -1).all() and (m.get_VertexPartitionMap() < 2).all()


@pytest.mark.parametrize("shape,nx,ny,P", [("tri", 40, 24, 4), ("tri", 120, 80, 8), ("quad", 60, 40, 3)])
def test_metis_partition_properties(shape, nx, ny, P):
    """partitionMesh on meshes large enough to have a real cut: every part used, balanced within METIS's own tolerance
    for nodal partitions, far fewer cut faces than a random map, reproducible for a fixed seed, and an irregular
    neighbour graph (some rank has >= 3 neighbours at P = 8) -- what the halo code must cope with."""
    from blitzdg_b200.synthetic import structured_mesh

    m = structured_mesh(shape, nx, ny)
    K, Nf = m.get_NumElements(), m.get_NumFaces()
    ep = m.partitionMesh(P, method="metis", seed=11).copy()
    assert ep.min() == 0 and ep.max() == P - 1
    cnt = np.bincount(ep, minlength=P)
    assert cnt.max() <= 1.15 * K / P and cnt.min() >= 0.8 * K / P
    e2e = m.get_EToE().reshape(K, Nf)
    cut = int((ep[e2e] != ep[:, None]).sum()) // 2
    assert cut < 12 * np.sqrt(K * P)
    assert np.array_equal(ep, m.partitionMesh(P, method="metis", seed=11))
    vp = m.get_VertexPartitionMap()
    assert vp.min() >= 0 and vp.max() < P
    if P == 8:
        nbrs = [len(set(ep[e2e[ep == r]].ravel().tolist()) - {r}) for r in range(P)]
        assert max(nbrs) >= 3


def test_partition_fallback_is_rcb():
    from blitzdg_b200.synthetic import structured_mesh

    m = structured_mesh("tri", 16, 16)
    ep = m.partitionMesh(4, method="rcb")
    assert m.PartitionInfo["method"] == "rcb" and np.bincount(ep).tolist() == [128] * 4
