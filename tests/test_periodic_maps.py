"""Re-mapped '+' traces: vmapP edited on the host so that two opposite boundaries become one periodic seam.

The reference's Python drivers do this with swhelpers/maps.py:3-43 (makeMapsPeriodic: the '+' node of a boundary face
node becomes the node at the same y on the opposite boundary) before handing vmapM/vmapP to the RHS
(sw2d_curved.py:144).  For the hot path this means vmapP may point anywhere in the mesh, not only into a face
neighbour: the device element order, the fused kernel's tiles and the global-speed launch must cope.
"""
import numpy as np
import pytest
from conftest import kernel_variants, mean_geometry

from blitzdg_b200.synthetic import provision, structured_mesh

RHS_TOL = 1e-12
STEP_TOL = 1e-10


def relerr(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


def periodic_in_x(ctx):
    """Pair the faces on x = -1 with the faces on x = +1 spanning the same y interval, node by node; returns the number
    of paired face nodes.  Edits ctx.vmapP and ctx.BCmap[3] in place (same result as makeMapsPeriodic for this geometry,
    built face-wise so the corner nodes shared by two faces pair with the right one)."""
    Nfp, Nf, K, Np = ctx.Nfp, ctx.Nfaces, ctx.K, ctx.Np
    xf, yf = ctx.x.flatten("F")[ctx.vmapM], ctx.y.flatten("F")[ctx.vmapM]
    wall = set(int(i) for i in ctx.BCmap[3])
    faces = {}
    for fid in range(Nf * K):
        ids = np.arange(fid * Nfp, (fid + 1) * Nfp)
        if int(ids[0]) not in wall:
            continue
        for side, xs in (("L", -1.0), ("R", 1.0)):
            if np.all(np.abs(xf[ids] - xs) < 1e-12):
                key = (side, round(float(yf[ids].min()), 9), round(float(yf[ids].max()), 9))
                faces[key] = ids
    vmapP = ctx.vmapP.copy()
    paired = 0
    for (side, ya, yb), ids in faces.items():
        if side != "L":
            continue
        other = faces[("R", ya, yb)]
        for i in ids:
            j = other[np.argmin(np.abs(yf[other] - yf[i]))]
            assert abs(yf[j] - yf[i]) < 1e-12
            vmapP[i], vmapP[j] = ctx.vmapM[j], ctx.vmapM[i]
            wall.discard(int(i))
            wall.discard(int(j))
            paired += 2
    ctx.vmapP = vmapP
    ctx.BCmap = dict(ctx.BCmap)
    ctx.BCmap[3] = np.array(sorted(wall), dtype=np.int32)
    return paired


def _case(shape, N, nx=6, ny=4):
    ctx = provision(structured_mesh(shape, nx, ny), N).get_DGContext()
    n = periodic_in_x(ctx)
    assert n == 2 * ny * ctx.Nfp
    x, y = ctx.x, ctx.y
    H = np.full_like(x, 10.0)
    # periodic in x, with flow through the seam (a modest mean flow: a large uniform momentum makes RHS1 a small
    # difference of large flux derivatives and the 1e-12 bar, relative to max|RHS1|, a statement about conditioning)
    h = H + 0.5 * np.exp(-8 * y * y) * (1 + 0.3 * np.cos(np.pi * x))
    hu = 0.1 * h + 0.2 * np.sin(np.pi * x)
    hv = 0.1 * np.sin(np.pi * x) * np.cos(0.5 * np.pi * y)
    return ctx, H, h, hu, hv


def _oracle(*a, **k):
    from oracle.binding import Oracle

    return Oracle(*a, **k)


@pytest.mark.parametrize("shape,N", [("tri", 4), ("quad", 3)])
def test_oracle_periodic_seam(shape, N):
    """Second derivation agrees; a seam-normal uniform flow crosses the seam without a residual."""
    from oracle import sw2d_numpy

    ctx, H, h, hu, hv = _case(shape, N)
    R = _oracle(ctx).rhs(h, hu, hv)
    Rn = sw2d_numpy.compute_rhs(ctx, h, hu, hv)
    for a, b in zip(R, Rn):
        assert relerr(a, b) < 1e-12  # BLAS dot products vs sequential sums
    # uniform flow in x: with walls on x = +-1 it would reflect; through the seam the RHS vanishes
    one = np.full_like(h, 10.0)
    for a in _oracle(ctx).rhs(one, 3.0 * one, 0.0 * one):
        assert np.abs(a).max() < 1e-10


@pytest.mark.gpu
@pytest.mark.parametrize("shape,N", [("tri", 4), ("tri", 7), ("quad", 3)])
@pytest.mark.parametrize("mode", ["simple", "full"])
def test_cuda_periodic_seam(shape, N, mode):
    from blitzdg_b200.sw2d import SW2dSolver

    ctx, H, h, hu, hv = _case(shape, N)
    full = mode == "full"
    orc = _oracle(ctx, lf_mode=int(full), sources=int(full))
    z = np.zeros_like(H)
    orc.set_bathymetry(H, z, z)
    sol = SW2dSolver(ctx, mode=mode)
    sol.set_bathymetry(H, z, z)
    variants = kernel_variants(sol)
    Ro = orc.rhs(h, hu, hv)
    ho, huo, hvo = h.copy(), hu.copy(), hv.copy()
    res = [np.zeros_like(h) for _ in range(3)]
    dt = 1e-3
    orc.step(ho, huo, hvo, res, 50, dt)
    # The fused kernel reads 13 geometry doubles per straight-sided triangle -- the element / face means of the per-node copies
    # the reference stores.  Those copies scatter by the rounding noise of the reference's own setup; where a field's RHS is a
    # small remainder of large terms (RHS1 here: 0.65 against flux derivatives of 50-160) that scatter alone moves the
    # reference-ordered result by more than 1e-12 at N = 7 (5e-12, measured with the extended-precision arbiter on both
    # geometries -- no kernel involved).  Variant 2 is therefore held to the oracle on the SAME averaged geometry at 1e-12,
    # and to the plain oracle within that measured effect.
    Rom = None
    if 2 in variants and ctx.Nfaces == 3:
        om = _oracle(mean_geometry(ctx), lf_mode=int(full), sources=int(full))
        om.set_bathymetry(H, z, z)
        Rom = om.rhs(h, hu, hv)
    for var in variants:
        sol.set_variant(var)
        sol.set_state(h, hu, hv)
        R = sol.computeRHS()
        for c in range(3):
            if var == 2 and Rom is not None:
                assert relerr(R[c], Rom[c]) < RHS_TOL, (var, c)
                assert relerr(R[c], Ro[c]) < RHS_TOL + 1.1 * relerr(Rom[c], Ro[c]), (var, c)
            else:
                assert relerr(R[c], Ro[c]) < RHS_TOL, (var, c)
        sol.step(50, dt)
        out = sol.get_state()
        for a, b in zip(out[:3], (ho, huo, hvo)):
            assert relerr(a, b) < STEP_TOL, var
    sol.close()
