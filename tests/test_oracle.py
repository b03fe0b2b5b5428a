"""CPU tests of the oracle: the reference's BlitzHelpers known answers, the LSERK4 table, and the
self-checks that stand in for the missing golden RHS vectors (SURVEY.md §8c i-iv)."""
import ctypes as C

import numpy as np
import pytest

from blitzdg_b200 import LSERK4, QuadNodesProvisioner, TriangleNodesProvisioner
from blitzdg_b200.initial_conditions import gaussian_hump
from oracle import binding, sw2d_numpy
from oracle.binding import Oracle

_dp = C.POINTER(C.c_double)


def test_full_to_vector_known_answer():
    """/root/reference/src/test/BlitzHelpersTests.cpp:99-144 (column-wise ordering)."""
    mat = np.array([[2, 3, 0, 0, 0], [3, 0, 4, 0, 6], [0, -1, -3, 2, 0], [0, 0, 1, 0, 0], [0, 4, 2, 0, 1]], dtype=float)
    expected = np.array([2, 3, 0, 0, 0, 3, 0, -1, 0, 4, 0, 4, -3, 1, 2, 0, 0, 2, 0, 0, 0, 6, 0, 0, 1], dtype=float)
    vec = np.zeros(25)
    binding.lib().oracle_full_to_vector(mat.ctypes.data_as(_dp), vec.ctypes.data_as(_dp), 5, 5)
    assert np.array_equal(vec, expected)
    back = np.zeros((5, 5))
    binding.lib().oracle_vector_to_full(vec.ctypes.data_as(_dp), back.ctypes.data_as(_dp), 5, 5)
    assert np.array_equal(back, mat)  # BlitzHelpersTests.cpp:146-192


def test_apply_index_map_known_answer():
    """/root/reference/src/test/BlitzHelpersTests.cpp:194-210."""
    vec = np.array([1.1, 1.2, 1.3, 1.4, 1.5])
    mp = np.array([0, 2, 4], dtype=np.int32)
    out = np.zeros(3)
    binding.lib().oracle_apply_index_map(vec.ctypes.data_as(_dp), mp.ctypes.data_as(C.POINTER(C.c_int32)), out.ctypes.data_as(_dp), 3)
    assert np.array_equal(out, [1.1, 1.3, 1.5])


def test_lserk4_table_matches_header():
    a, b = np.zeros(5), np.zeros(5)
    binding.lib().oracle_lserk4_coeffs(a.ctypes.data_as(_dp), b.ctypes.data_as(_dp))
    assert np.array_equal(a, LSERK4.rk4a) and np.array_equal(b, LSERK4.rk4b)
    assert a[0] == 0.0 and abs(sum(b) - 1.0) > 0  # not a convex table; just pinned values


def _ctx(mesh, N, quad=False):
    P = QuadNodesProvisioner if quad else TriangleNodesProvisioner
    p = P(N, mesh)
    p.buildFilter(0.9 * N, N)
    return p.get_DGContext()


@pytest.mark.parametrize("quad,N", [(False, 1), (False, 4), (False, 8), (True, 3), (True, 6)])
def test_c_oracle_agrees_with_numpy_transcription_simple(coarse_box_mesh, coarse_quads_mesh, quad, N):
    ctx = _ctx(coarse_quads_mesh if quad else coarse_box_mesh, N, quad)
    H, h, hu, hv = gaussian_hump(ctx.x, ctx.y)
    rng = np.random.default_rng(1)
    hu = 0.3 * rng.standard_normal(h.shape)
    hv = 0.3 * rng.standard_normal(h.shape)
    R = Oracle(ctx, lf_mode=0, sources=0).rhs(h, hu, hv)
    Rn = sw2d_numpy.compute_rhs(ctx, h, hu, hv)
    for a, b in zip(R, Rn):
        assert np.abs(a - b).max() <= 1e-13 * max(1.0, np.abs(b).max())


def test_c_oracle_agrees_with_numpy_transcription_full(coarse_box_mesh):
    ctx = _ctx(coarse_box_mesh, 4)
    H = 10 + 0.5 * ctx.x + 0.25 * ctx.y
    Hx, Hy = np.full_like(H, 0.5), np.full_like(H, 0.25)
    h = H + np.exp(-10 * ctx.x ** 2 - 10 * ctx.y ** 2)
    rng = np.random.default_rng(2)
    hu, hv = 0.5 * rng.standard_normal(h.shape), 0.5 * rng.standard_normal(h.shape)
    # tag a few boundary faces Out to exercise the tidal branch
    wall = ctx.BCmap[3]
    ctx.BCmap = {3: wall[: len(wall) // 2], 2: wall[len(wall) // 2:]}
    o = Oracle(ctx, lf_mode=1, sources=1)
    o.set_bathymetry(H, Hx, Hy)
    R = o.rhs(h, hu, hv, t=1234.5)
    Rn = sw2d_numpy.compute_rhs(ctx, h, hu, hv, lf_mode=1, sources=True, H=H, Hx=Hx, Hy=Hy, t=1234.5)
    for a, b in zip(R, Rn):
        assert np.abs(a - b).max() <= 1e-13 * np.abs(b).max()
    # filtered variant
    o2 = Oracle(ctx, lf_mode=0, sources=0, filter=1)
    R2 = o2.rhs(h, hu, hv)
    R2n = sw2d_numpy.compute_rhs(ctx, h, hu, hv, use_filter=True)
    for a, b in zip(R2, R2n):
        assert np.abs(a - b).max() <= 1e-13 * np.abs(b).max()


def test_still_water_is_steady(coarse_box_mesh, coarse_quads_mesh):
    """(i) h=H, u=v=0 on a flat bottom => RHS == 0 to round-off, both modes."""
    for mesh, quad, N in ((coarse_box_mesh, False, 4), (coarse_quads_mesh, True, 4)):
        ctx = _ctx(mesh, N, quad)
        h = np.full_like(ctx.x, 10.0)
        z = np.zeros_like(h)
        for mode in (dict(lf_mode=0, sources=0), dict(lf_mode=1, sources=1)):
            o = Oracle(ctx, **mode)
            o.set_bathymetry(h, z, z)
            for R in o.rhs(h, z, z):
                assert np.abs(R).max() < 1e-9


def test_mass_conservation_all_wall(coarse_box_mesh):
    """(iii) sum_k J_k w^T RHS1 = 0 with all-wall BC; w = column sums of the mass matrix (V V^T)^-1."""
    ctx = _ctx(coarse_box_mesh, 4)
    H, h, hu, hv = gaussian_hump(ctx.x, ctx.y)
    rng = np.random.default_rng(3)
    hu, hv = 0.2 * rng.standard_normal(h.shape), 0.2 * rng.standard_normal(h.shape)
    # make momentum single-valued at shared nodes is not needed: DG conserves mass for any state
    R1 = Oracle(ctx).rhs(h, hu, hv)[0]
    M = np.linalg.inv(ctx.V @ ctx.V.T)
    w = M.sum(axis=0)
    total = (w[:, None] * ctx.J * R1).sum()
    scale = (w[:, None] * ctx.J * np.abs(R1)).sum()
    assert abs(total) < 1e-12 * scale


def test_lserk4_fourth_order(coarse_box_mesh):
    """(iv) temporal convergence under dt-halving (error vs a dt/8 run)."""
    ctx = _ctx(coarse_box_mesh, 2)
    o = Oracle(ctx)
    H, h0, hu0, hv0 = gaussian_hump(ctx.x, ctx.y)
    dt0, _ = o.timestep(h0, hu0, hv0)
    T = 8 * dt0

    def run(nsub):
        h, hu, hv = h0.copy(), hu0.copy(), hv0.copy()
        res = [np.zeros_like(h) for _ in range(3)]
        o.step(h, hu, hv, res, 8 * nsub, T / (8 * nsub))
        return h

    ref = run(8)
    e1 = np.abs(run(1) - ref).max()
    e2 = np.abs(run(2) - ref).max()
    assert e1 / e2 > 12.0  # 2^4 = 16 asymptotically


def test_timestep_formula(coarse_box_mesh):
    ctx = _ctx(coarse_box_mesh, 4)
    H, h, hu, hv = gaussian_hump(ctx.x, ctx.y)
    dt, mx = Oracle(ctx, CFL=0.25).timestep(h, hu, hv)
    hM = h.flatten("F")[ctx.vmapM]
    fs = ctx.Fscale.flatten("F")
    assert mx == pytest.approx((fs * np.sqrt(9.81 * hM)).max(), rel=1e-15)
    assert dt == pytest.approx(0.25 / (25 * 0.5 * mx), rel=1e-15)
