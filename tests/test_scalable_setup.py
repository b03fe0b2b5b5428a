"""Host setup at scale: the sort-based connectivity and '+' maps of the host mirror give
what the reference's sparse-product / all-pairs algorithms give, on meshes whose element and vertex numbering carries
no structure, and provision 10^5-10^6 elements in seconds."""
import time

import numpy as np
import pytest

from blitzdg_b200.synthetic import provision, structured_mesh


def shuffled_mesh(shape, nx, ny, jitter, seed):
    """Structured cells with jittered interior vertices and a random numbering (synthetic.shuffle_numbering): nothing
    of the generator's regularity is left."""
    return structured_mesh(shape, nx, ny, jitter, shuffle=seed)


@pytest.mark.parametrize("shape,N,nx,ny", [("tri", 3, 9, 7), ("quad", 2, 8, 5)])
def test_sorted_setup_equals_reference_algorithms(shape, N, nx, ny):
    from oracle.setup_numpy import build_connectivity, build_vmapP

    m = shuffled_mesh(shape, nx, ny, 0.25, seed=11)
    K, Nf, Nv = m.get_NumElements(), m.get_NumFaces(), m.get_NumVerts()
    E2E, E2F = build_connectivity(m.get_Elements(), K, Nf, Nv)
    assert np.array_equal(m.get_EToE(), E2E) and np.array_equal(m.get_EToF(), E2F)
    P = provision(m, N)
    vM, vP = build_vmapP(P.get_xGrid(), P.get_yGrid(), P.Fmask, E2E, E2F, P.NumLocalPoints, P.NumFacePoints, Nf, K)
    assert np.array_equal(P.get_vmapM(), vM) and np.array_equal(P.get_vmapP(), vP)
    nb = (E2E.reshape(K, Nf) == np.arange(K)[:, None]).sum()
    assert nb == (2 * (nx + ny))  # boundary faces self-reference
    assert len(P.get_bcMap()[3]) == nb * P.NumFacePoints


def test_unstructured_numbering_at_scale():
    """200 k triangles with a random numbering: seconds, and the maps are a consistent involution on matching nodes."""
    t0 = time.perf_counter()
    m = shuffled_mesh("tri", 400, 250, 0.2, seed=5)
    P = provision(m, 4)
    ctx = P.get_DGContext()
    el = time.perf_counter() - t0
    assert el < 60.0, el
    vM, vP = np.asarray(ctx.vmapM), np.asarray(ctx.vmapP)
    xv, yv = ctx.x.T.reshape(-1), ctx.y.T.reshape(-1)
    assert np.abs(xv[vM] - xv[vP]).max() < 1e-9 and np.abs(yv[vM] - yv[vP]).max() < 1e-9
    # '+' of '+' is the node itself (a corner node sits on two faces, so match on (node, partner) pairs)
    interior = vM != vP
    pairs = set(zip(vM[interior].tolist(), vP[interior].tolist()))
    assert all((b, a) in pairs for a, b in pairs)
    assert (~interior).sum() == 2 * (400 + 250) * ctx.Nfp
    assert np.allclose(ctx.nx ** 2 + ctx.ny ** 2, 1.0)
