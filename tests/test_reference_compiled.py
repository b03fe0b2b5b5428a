"""The oracle against the reference's OWN code: blitzdg::sw2d::computeRHS of both drivers, computeTimeStep,
buildSpongeCoeff and the SSP-RK2 loop body, compiled VERBATIM from /root/reference/src/sw2d/main.cpp:211-236, 251-484,
517-553 and src/sw2d-simple/main.cpp:179-356 by oracle/make_ref.py (with the reference's own BlitzHelpers.hpp /
DGContext2D.hpp / Types.hpp; only <blitz/array.h> is a stand-in, oracle/blitz_shim).  This pins the lines that only the
C++ driver has and that no Python reference covers: global Lax-Friedrichs maximum, star states on a DISCONTINUOUS bed,
Out faces with tidal forcing, drag / Coriolis / bed-slope sources in the C++ sign convention, SSP-RK2 + sponge.
The C restatement (oracle/sw2d_oracle.c) follows the reference's operation order, so the agreement is BITWISE."""
import copy

import numpy as np
import pytest

from blitzdg_b200 import TriangleNodesProvisioner
from blitzdg_b200.initial_conditions import gaussian_hump
from oracle.binding import Oracle
from oracle.ref_binding import CompiledReference, available

pytestmark = pytest.mark.skipif(not available(), reason="oracle/_ref/libsw2d_ref.so not built (needs /root/reference at build time)")


def _ctx(mesh, N, split_out=True):
    ctx = TriangleNodesProvisioner(N, mesh).get_DGContext()
    if split_out:
        ctx = copy.copy(ctx)
        wall = ctx.BCmap[3]
        nf = len(wall) // ctx.Nfp
        ctx.BCmap = {3: wall[: (nf // 2) * ctx.Nfp], 2: wall[(nf // 2) * ctx.Nfp:]}
    return ctx


def _state(ctx, seed, bed="slope"):
    _, h, _, _ = gaussian_hump(ctx.x, ctx.y)
    rng = np.random.default_rng(seed)
    hu, hv = 0.4 * rng.standard_normal(h.shape), 0.4 * rng.standard_normal(h.shape)
    H = 10 + 0.5 * ctx.x + 0.25 * ctx.y
    if bed == "steps":  # a different bed level in every element: HM != HP on every interior face, the star states bite
        H = H + 0.3 * rng.standard_normal(ctx.K)[None, :]
    Hx, Hy = np.full_like(H, 0.5), np.full_like(H, 0.25)
    return H, Hx, Hy, np.ascontiguousarray(H + (h - 10.0)), hu, hv


@pytest.mark.parametrize("N", [1, 2, 4, 7])
def test_simple_rhs_bitwise(coarse_box_mesh, N):
    ctx = _ctx(coarse_box_mesh, N, split_out=False)
    _, _, _, h, hu, hv = _state(ctx, N)
    Rr = CompiledReference(ctx).rhs_simple(h, hu, hv)
    Ro = Oracle(ctx).rhs(h, hu, hv)
    for a, b in zip(Ro, Rr):
        assert np.array_equal(a, b)


@pytest.mark.parametrize("N,bed", [(1, "slope"), (4, "slope"), (4, "steps"), (6, "steps")])
def test_full_rhs_bitwise(coarse_box_mesh, N, bed):
    ctx = _ctx(coarse_box_mesh, N)
    H, Hx, Hy, h, hu, hv = _state(ctx, 10 + N, bed)
    ref = CompiledReference(ctx)
    ref.set_bathymetry(H, Hx, Hy)
    orc = Oracle(ctx, lf_mode=1, sources=1)
    orc.set_bathymetry(H, Hx, Hy)
    for t in (0.0, 4321.0, 3600 * 12.42 + 900.0):
        for a, b in zip(orc.rhs(h, hu, hv, t=t), ref.rhs_full(h, hu, hv, t)):
            assert np.array_equal(a, b), (N, bed, t)
    assert orc.timestep(h, hu, hv)[0] == ref.timestep(h, hu, hv, CFL=0.25)


def test_full_rhs_without_out_faces(coarse_box_mesh):
    """All-wall mesh: the reference would throw at bcHash.at(Out); with the key present and empty it runs."""
    ctx = _ctx(coarse_box_mesh, 3, split_out=False)
    H, Hx, Hy, h, hu, hv = _state(ctx, 3)
    ref = CompiledReference(ctx)
    ref.set_bathymetry(H, Hx, Hy)
    orc = Oracle(ctx, lf_mode=1, sources=1)
    orc.set_bathymetry(H, Hx, Hy)
    for a, b in zip(orc.rhs(h, hu, hv), ref.rhs_full(h, hu, hv)):
        assert np.array_equal(a, b)


def test_sponge_and_ssprk2_loop_bitwise(coarse_box_mesh):
    ctx = _ctx(coarse_box_mesh, 4)
    H, Hx, Hy, h, hu, hv = _state(ctx, 5, "steps")
    ref = CompiledReference(ctx)
    ref.set_bathymetry(H, Hx, Hy)
    orc = Oracle(ctx, lf_mode=1, sources=1)
    orc.set_bathymetry(H, Hx, Hy)
    out_nodes = np.asarray(ctx.BCmap[2], dtype=np.int32)
    sp_o = orc.build_sponge(ctx.x, ctx.y, out_nodes, 40.0, 0.6)
    sp_r = ref.build_sponge(ctx.x, ctx.y, 40.0, 0.6)
    assert np.array_equal(sp_o, sp_r) and sp_r.max() > 0
    for dt in (2e-4, 0.0):  # fixed and computeTimeStep-driven
        a = [h.copy(), hu.copy(), hv.copy()]
        b = [h.copy(), hu.copy(), hv.copy()]
        ta = orc.step_ssprk2(a[0], a[1], a[2], 12, dt, t=100.0, sponge=sp_o)
        tb = ref.step_ssprk2(b[0], b[1], b[2], 12, dt, t=100.0, sponge=sp_r)
        assert ta == tb
        for x, y in zip(a, b):
            assert np.array_equal(x, y)
