"""CPU checks of the extended-precision arbiter (oracle/sw2d_arbiter.c) and of the strip-wise full-mesh oracle driver
used by tests/test_fullsize_parity.py."""
import numpy as np
import pytest

from blitzdg_b200 import QuadNodesProvisioner, TriangleNodesProvisioner
from blitzdg_b200.initial_conditions import gaussian_hump
from blitzdg_b200.synthetic import provision, structured_mesh
from oracle.binding import Arbiter, Oracle


def relerr(a, b):
    return np.abs(a - b).max() / np.abs(b).max()


def _state(ctx, seed):
    H, h, hu, hv = gaussian_hump(ctx.x, ctx.y)
    rng = np.random.default_rng(seed)
    return H, h, 0.3 * rng.standard_normal(h.shape), 0.3 * rng.standard_normal(h.shape)


@pytest.mark.parametrize("quad,N", [(False, 4), (False, 7), (True, 3), (True, 6)])
def test_arbiter_agrees_with_oracle_on_unit_size_meshes(coarse_box_mesh, coarse_quads_mesh, quad, N):
    """A third derivation of the RHS (element-wise, long double) against the reference-ordered fp64 oracle: on the
    coarse box the fp64 rounding noise is ~1e-14, far below the 1e-12 bar."""
    P = QuadNodesProvisioner if quad else TriangleNodesProvisioner
    prov = P(N, coarse_quads_mesh if quad else coarse_box_mesh)
    prov.buildFilter(0.9 * N, 4)
    ctx = prov.get_DGContext()
    H, h, hu, hv = _state(ctx, N)
    for kw in (dict(), dict(filter=0 if quad else 1)):
        Ro, Ra = Oracle(ctx, **kw).rhs(h, hu, hv), Arbiter(ctx, **kw).rhs(h, hu, hv, threads=3)
        assert max(relerr(a, b) for a, b in zip(Ro, Ra)) < 1e-13


def test_arbiter_full_mode_open_boundary(coarse_box_mesh):
    """Global LF, star states on a sloping bed, Out faces with tide, sources: the C++-only lines of src/sw2d/main.cpp
    through the independent element-wise derivation."""
    ctx = TriangleNodesProvisioner(4, coarse_box_mesh).get_DGContext()
    _, h, hu, hv = _state(ctx, 7)
    H = 10 + 0.5 * ctx.x + 0.25 * ctx.y
    Hx, Hy = np.full_like(H, 0.5), np.full_like(H, 0.25)
    h = H + (h - 10.0)
    wall = ctx.BCmap[3]
    nf = len(wall) // ctx.Nfp
    ctx.BCmap = {3: wall[: (nf // 2) * ctx.Nfp], 2: wall[(nf // 2) * ctx.Nfp:]}
    kw = dict(lf_mode=1, sources=1)
    o, a = Oracle(ctx, **kw), Arbiter(ctx, **kw)
    o.set_bathymetry(H, Hx, Hy)
    a.set_bathymetry(H, Hx, Hy)
    for t in (0.0, 4321.0):
        assert max(relerr(x, y) for x, y in zip(o.rhs(h, hu, hv, t=t), a.rhs(h, hu, hv, t=t))) < 1e-13


def test_long_double_arbiter_against_float128_build(coarse_box_mesh):
    ctx = TriangleNodesProvisioner(5, coarse_box_mesh).get_DGContext()
    H, h, hu, hv = _state(ctx, 1)
    try:
        Rq = Arbiter(ctx, quad=True).rhs(h, hu, hv)
    except Exception as exc:  # no libquadmath on this host
        pytest.skip("float128 build unavailable: %s" % exc)
    Ra = Arbiter(ctx).rhs(h, hu, hv)
    assert max(relerr(a, b) for a, b in zip(Ra, Rq)) < 1e-15  # the double rounding of the output dominates


def test_fp64_noise_floor_grows_with_resolution():
    """The reason 1e-12 per RHS cannot hold against ANY fp64 evaluation on the BASELINE-size meshes: the reference
    arithmetic's own distance from the exact operator grows like 1/h_cell (measured: 4e-13 at 64x32, 2e-12 at 250x125,
    8e-12 at 1000x500 cells)."""
    errs = []
    for nx, ny in ((32, 16), (128, 64)):
        ctx = provision(structured_mesh("tri", nx, ny), 4).get_DGContext()
        H, h, hu, hv = gaussian_hump(ctx.x, ctx.y)
        Ro, Ra = Oracle(ctx).rhs(h, hu, hv), Arbiter(ctx).rhs(h, hu, hv)
        errs.append(max(relerr(a, b) for a, b in zip(Ro[1:], Ra[1:])))
    assert errs[1] > 2.0 * errs[0] and errs[1] < 2e-11


def test_strip_wise_oracle_equals_whole_mesh_oracle():
    from test_fullsize_parity import oracle_rhs_strips

    mesh = structured_mesh("tri", 24, 20)
    ctx = provision(mesh, 3).get_DGContext()
    H, h, hu, hv = _state(ctx, 3)
    Ro = Oracle(ctx).rhs(h, hu, hv)
    Rs, mask = oracle_rhs_strips(mesh, 3, (h, hu, hv), 5)
    assert mask.all()
    assert max(relerr(a, b) for a, b in zip(Rs, Ro)) < 1e-14
    Rs, mask = oracle_rhs_strips(mesh, 3, (h, hu, hv), 5, strips=(1, 3))
    assert mask.sum() == 2 * ctx.K // 5
    assert max(relerr(a[:, mask], b[:, mask]) for a, b in zip(Rs, Ro)) < 1e-14
