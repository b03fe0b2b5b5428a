"""Every array the hot path is handed (DGContext2D: SURVEY.md 8(a) a18, 8(b)) against the REFERENCE'S OWN provisioners,
compiled verbatim from /root/reference by oracle/make_ref_setup.py (oracle/_ref/libsetup_ref.so): the bodies of the
TriangleNodesProvisioner / QuadNodesProvisioner constructors -- buildNodes, buildLift, buildPhysicalGrid, buildMaps (+
buildBCHash) -- and buildFilter, with the Jacobi / Vandermonde helpers they call.

The reference's specs pin these arrays at N = 3 on coarse_box to 5-6 digits (Lift, Dr, Ds to 5.8e-5);
here the host mirrors (blitzdg_b200/triangle_nodes.py, quad_nodes.py) are held to the reference's arithmetic at the orders
the BASELINE configs run (N = 4, 6, 8) on the reference's gmsh meshes and on jittered synthetic ones: real arrays to 2e-12
(measured 1e-15 .. 9e-13, growing with N as V^-1 does), integer tables -- Fmask, vmapM, vmapP, mapP, mapB, vmapB, BCmap
with its insertion order -- identical.  Finally the whole chain: the reference's provisioner feeding the reference's
computeRHS against the mirror feeding the same function.
The library is prebuilt where the reference tree exists and travels with the snapshot; absent library and absent
reference => skip.  Not the reference's (LAPACK is absent): matrix inverse / dense solve loops and the Gauss-Jacobi points
behind the GLL nodes (the mirror's, held to the reference's JacobiGL golden vectors in tests/test_setup_golden.py).
"""
import os

import numpy as np
import pytest

from oracle import ref_setup_binding as rsb

REAL = ("r", "s", "V", "Vinv", "Dr", "Ds", "Lift", "x", "y", "J", "rx", "sx", "ry", "sy", "nx", "ny", "Fscale")
MINE = dict(r="rGrid", s="sGrid", x="xGrid", y="yGrid")
INTS = ("Fmask", "vmapM", "vmapP", "mapP", "mapB", "vmapB")
TOL = 2e-12


@pytest.fixture(scope="module")
def ref():
    if not rsb.available():
        pytest.skip("compiled reference not built and reference tree not present")
    return rsb


def _mesh(kind, coarse_box_mesh, coarse_quads_mesh):
    from blitzdg_b200.synthetic import structured_mesh

    return {"coarse_box": coarse_box_mesh, "coarse_quads": coarse_quads_mesh,
            "tri_jitter": structured_mesh("tri", 7, 5, jitter=0.2), "quad_jitter": structured_mesh("quad", 6, 4, jitter=0.2),
            "tri_shuffled": structured_mesh("tri", 6, 6, shuffle=5)}[kind]


def _check(want, t, filt):
    worst = 0.0
    for k in REAL + (("Filter",) if filt else ()):
        mine, w = np.asarray(getattr(t, MINE.get(k, k))), want[k]
        assert mine.shape == w.shape, k
        err = float(np.abs(mine - w).max() / max(1.0, np.abs(w).max()))
        assert err <= TOL, (k, err)
        worst = max(worst, err)
    for k in INTS:
        assert np.array_equal(np.asarray(getattr(t, k)), want[k]), k
    assert sorted(t.BCmap) == sorted(want["BCmap"])
    for tag, ids in want["BCmap"].items():
        assert np.array_equal(np.asarray(t.BCmap[tag]), ids), tag  # same face-node ids in the reference's insertion order
    return worst


@pytest.mark.parametrize("kind", ["coarse_box", "tri_jitter", "tri_shuffled"])
@pytest.mark.parametrize("N", [1, 2, 3, 4, 6, 8])
def test_triangle_provisioner_matches_the_compiled_reference(ref, kind, N, coarse_box_mesh, coarse_quads_mesh):
    from blitzdg_b200 import TriangleNodesProvisioner

    mesh = _mesh(kind, coarse_box_mesh, coarse_quads_mesh)
    want = ref.provision(mesh, N, quad=False, filt=(0.9 * N, 4))
    t = TriangleNodesProvisioner(N, mesh)
    t.buildFilter(0.9 * N, 4)
    _check(want, t, True)
    # what get_DGContext hands to the C ABI is those arrays
    ctx = t.get_DGContext()
    for k in ("Dr", "Ds", "Lift", "rx", "sx", "ry", "sy", "nx", "ny", "Fscale"):
        assert np.abs(np.asarray(getattr(ctx, k)) - want[k]).max() <= TOL * max(1.0, np.abs(want[k]).max()), k
    assert np.array_equal(ctx.vmapM, want["vmapM"]) and np.array_equal(ctx.vmapP, want["vmapP"])


@pytest.mark.parametrize("kind", ["coarse_quads", "quad_jitter"])
@pytest.mark.parametrize("N", [1, 2, 3, 6])
def test_quad_provisioner_matches_the_compiled_reference(ref, kind, N, coarse_box_mesh, coarse_quads_mesh):
    """Includes the quad buildFilter as the reference executes it (only (N+1)(N+2)/2 of the (N+1)^2 diagonal entries are
    filled, src/QuadNodesProvisioner.cpp:185-196): the mirror reproduces it, the drivers do not enable it."""
    from blitzdg_b200 import QuadNodesProvisioner

    mesh = _mesh(kind, coarse_box_mesh, coarse_quads_mesh)
    want = ref.provision(mesh, N, quad=True, filt=(0.9 * N, 4))
    t = QuadNodesProvisioner(N, mesh)
    t.buildFilter(0.9 * N, 4)
    _check(want, t, True)


@pytest.mark.parametrize("quad,N", [(False, 4), (False, 7)])  # the reference's C++ RHS is a triangle code (src/sw2d-simple/main.cpp:34)
def test_reference_setup_into_reference_rhs_against_mirror_setup(ref, quad, N, coarse_box_mesh, coarse_quads_mesh):
    """The chain as the reference runs it -- its provisioner's arrays into its computeRHS (oracle/_ref/libsw2d_ref.so) --
    against the mirror's arrays into the same function: the setup difference seen through the RHS."""
    import dataclasses

    from blitzdg_b200 import QuadNodesProvisioner, TriangleNodesProvisioner
    from blitzdg_b200.initial_conditions import gaussian_hump
    from oracle import ref_binding

    if not ref_binding.available():
        pytest.skip("compiled reference RHS not built")
    mesh = coarse_quads_mesh if quad else coarse_box_mesh
    want = ref.provision(mesh, N, quad=quad)
    ctx = (QuadNodesProvisioner if quad else TriangleNodesProvisioner)(N, mesh).get_DGContext()
    rctx = dataclasses.replace(ctx, **{k: np.ascontiguousarray(want[k]) for k in ("Dr", "Ds", "Lift", "rx", "sx", "ry", "sy", "nx", "ny", "Fscale", "vmapM", "vmapP", "J", "x", "y")},
                               BCmap={k: v.astype(np.int32) for k, v in want["BCmap"].items()})
    H, h, hu, hv = gaussian_hump(ctx.x, ctx.y)
    hu = 0.3 * h * np.sin(2 * ctx.x + ctx.y)
    hv = -0.2 * h * np.cos(ctx.x - 3 * ctx.y)
    Ra = ref_binding.CompiledReference(rctx).rhs_simple(h, hu, hv)
    Rb = ref_binding.CompiledReference(ctx).rhs_simple(h, hu, hv)
    for a, b in zip(Ra, Rb):
        assert np.abs(a - b).max() <= 1e-12 * np.abs(a).max()


# ---------------------------------------------------------------------------------------------- mesh tables (MeshManager)
def _raw_mesh(shape, nx, ny, seed, clockwise_fraction=0.0):
    from blitzdg_b200.synthetic import shuffle_numbering, structured_arrays

    etov, vert = structured_arrays(shape, nx, ny, jitter=0.2)
    if seed is not None:
        etov, vert = shuffle_numbering(etov, vert, seed)
    etov = np.asarray(etov).copy()
    if clockwise_fraction:
        flip = np.random.default_rng(1).random(len(etov)) < clockwise_fraction
        etov[flip] = etov[flip][:, ::-1]  # reversed orientation
    if vert.shape[1] == 2:
        vert = np.concatenate([vert, np.zeros((len(vert), 1))], axis=1)
    return etov, vert


@pytest.fixture(scope="module")
def refmesh():
    from oracle import ref_mesh_binding as rmb

    if not rmb.available():
        pytest.skip("compiled reference not built and reference tree not present")
    return rmb


@pytest.mark.parametrize("shape,nx,ny,seed,cw", [("tri", 9, 7, None, 0.0), ("tri", 30, 20, 3, 0.3), ("tri", 64, 48, 9, 0.5),
                                                 ("quad", 8, 5, None, 0.0), ("quad", 25, 18, 11, 0.0)])
def test_mesh_tables_match_the_compiled_reference(refmesh, shape, nx, ny, seed, cw):
    """EToV after the counter-clockwise repair, EToE / EToF from the reference's sparse face-to-face product
    (src/MeshManager.cpp:383-489), BCType from buildBCTable(Wall): the mirror's sort-based tables are the same integers, on
    structured, randomly renumbered and partly clockwise input."""
    from blitzdg_b200 import MeshManager

    etov, vert = _raw_mesh(shape, nx, ny, seed, cw)
    m = MeshManager()
    m.buildMesh(etov, vert)
    K, Nf = etov.shape
    for mine, want in zip((m.get_Elements(), m.get_EToE(), m.get_EToF(), m.get_BCType()), refmesh.mesh_tables(etov, vert, tag=3)):
        assert np.array_equal(np.asarray(mine).reshape(K, Nf), want)
    if cw:
        assert not np.array_equal(np.asarray(m.get_Elements()).reshape(K, Nf), etov)  # the repair did something


def test_gmsh_files_of_the_reference_give_the_reference_tables(refmesh, coarse_box_mesh, coarse_quads_mesh):
    """The parser is the mirror's (the reference's needs boost); everything after it is the reference's code."""
    for mesh in (coarse_box_mesh, coarse_quads_mesh):
        K, Nf = mesh.get_NumElements(), mesh.get_NumFaces()
        etov = np.asarray(mesh.get_Elements()).reshape(K, Nf)
        vert = np.asarray(mesh.get_Vertices()).reshape(-1, 3)
        want = refmesh.mesh_tables(etov, vert, tag=3)
        for mine, w in zip((mesh.get_Elements(), mesh.get_EToE(), mesh.get_EToF(), mesh.get_BCType()), want):
            assert np.array_equal(np.asarray(mine).reshape(K, Nf), w)


def test_clockwise_quads_are_the_one_documented_deviation(refmesh):
    """the reference swaps local vertices 1 and 2 of a clockwise element whatever its type
    (src/MeshManager.cpp:292-306), which makes a bow-tie of a quad; the mirror reverses the orientation (1 <-> 3)."""
    from blitzdg_b200 import MeshManager

    etov, vert = _raw_mesh("quad", 4, 3, None, 1.0)
    m = MeshManager()
    m.buildMesh(etov, vert)
    mine = np.asarray(m.get_Elements()).reshape(etov.shape)
    want = refmesh.mesh_tables(etov, vert, tag=3)[0]
    assert np.array_equal(want, etov[:, [0, 2, 1, 3]]) and np.array_equal(mine, etov[:, [0, 3, 2, 1]])

    def signed_area(e):
        x, y = vert[e, 0], vert[e, 1]
        return 0.5 * (x * np.roll(y, -1, axis=1) - np.roll(x, -1, axis=1) * y).sum(axis=1)

    cell = signed_area(mine)
    assert (cell > 0).all()                                    # the mirror's elements are valid counter-clockwise quads
    assert (np.abs(signed_area(want)) < 0.2 * cell).all()      # the reference's are self-intersecting: the two lobes cancel
