"""Parity against outputs of the reference's OWN Python RHS code (tests/golden/python_reference_rhs.npz).

The fixture was produced in the build container by tests/golden/make_python_reference_vectors.py, which executes
/root/reference/sw2dquads.py:10-123, /root/reference/swhelpers/rhs.py:178-311 and /root/reference/swhelpers/limiters.py:3-21
unmodified on this repo's DGContext2D arrays.  The CPU tests pin the oracle to it; the gpu tests compare the CUDA path
with it through the C ABI.  Tolerance: BASELINE.json's relative 1e-12 per RHS evaluation (max-norm per field) -- the
Python drivers contract with BLAS dot products and write hu*u where the C++ writes (hu*hu)/h, so they agree with the
C++ evaluation order to round-off, not bitwise.
"""
import os

import numpy as np
import pytest
from conftest import kernel_variants

from blitzdg_b200 import MeshManager, QuadNodesProvisioner, TriangleNodesProvisioner

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
RHS_TOL = 1e-12
STEP_TOL = 1e-10  # BASELINE.json: after 1000 LSERK4 steps
STEP_CASES = [("tri4", "coarse_box.msh", False, 4), ("quad3", "coarse_box_quads.msh", True, 3)]
G = 9.81
F_COR = 1.0070e-4
CASES = [("tri4", "coarse_box.msh", False, 4), ("tri7", "coarse_box.msh", False, 7),
         ("quad3", "coarse_box_quads.msh", True, 3), ("quad6", "coarse_box_quads.msh", True, 6)]


@pytest.fixture(scope="module")
def ref():
    return np.load(os.path.join(GOLDEN, "python_reference_rhs.npz"))


def relerr(a, b):
    return np.abs(a - b).max() / np.abs(b).max()


def _ctx(msh, quad, N):
    mm = MeshManager()
    mm.readMesh(os.path.join(GOLDEN, msh))
    return (QuadNodesProvisioner if quad else TriangleNodesProvisioner)(N, mm).get_DGContext()


def _oracle(*a, **k):
    from oracle.binding import Oracle

    return Oracle(*a, **k)


def _solver(*a, **k):
    from blitzdg_b200.sw2d import SW2dSolver

    return SW2dSolver(*a, **k)


class _phys:
    g = G
    f = F_COR

    def __init__(self, CD):
        self.CD = CD


def _drag_sign_fix(ref, CD):
    """swhelpers/rhs.py:306 reads `RHS3 -= f*hu - CD_norm_u*v`, i.e. RHS3 += -f*hu + CD*|u|*v: the drag enters the
    v-momentum equation with the opposite sign to the u-momentum line above it (:305) and to src/sw2d/main.cpp:479
    (`RHS3 += -CD*v*|u|`).  The hot path follows the C++ driver; the fixture is compared after removing the difference."""
    h, hu, hv = ref["swhelpers_h"], ref["swhelpers_hu"], ref["swhelpers_hv"]
    u, v = hu / h, hv / h
    return -2.0 * CD * np.hypot(u, v) * v


# ---------------------------------------------------------------------------------------------- CPU: the oracle
@pytest.mark.parametrize("tag,msh,quad,N", CASES)
def test_oracle_matches_sw2dquads_rhs(ref, tag, msh, quad, N):
    ctx = _ctx(msh, quad, N)
    h, hu, hv = (ref["sw2dquads_%s_%s" % (tag, n)] for n in ("h", "hu", "hv"))
    assert h.shape == (ctx.Np, ctx.K)
    R = _oracle(ctx, lf_mode=0, sources=0).rhs(h, hu, hv)
    for c in range(3):
        assert relerr(R[c], ref["sw2dquads_%s_RHS%d" % (tag, c + 1)]) < RHS_TOL


@pytest.mark.parametrize("tag,CD", [("cd0", 0.0), ("cd", 2.5e-3)])
def test_oracle_matches_swhelpers_rhs_with_sources(ref, tag, CD):
    """Per-face LF + Coriolis + drag + bed slope (zx = -Hx); the star states of src/sw2d/main.cpp:355-368 are the
    identity up to round-off on a continuous bed."""
    ctx = _ctx("coarse_box.msh", False, 4)
    h, hu, hv, H = (ref["swhelpers_" + n] for n in ("h", "hu", "hv", "H"))
    o = _oracle(ctx, lf_mode=0, sources=1, CD=CD, f=F_COR)
    o.set_bathymetry(H, np.full_like(H, 0.5), np.full_like(H, 0.25))
    R = o.rhs(h, hu, hv, t=0.0)
    want = [ref["swhelpers_%s_RHS%d" % (tag, c + 1)] for c in range(3)]
    want[2] = want[2] + _drag_sign_fix(ref, CD)
    for c in range(3):
        assert relerr(R[c], want[c]) < RHS_TOL, c


@pytest.mark.parametrize("tag,msh,quad,N", STEP_CASES)
def test_oracle_matches_1000_lserk4_steps_of_reference_rhs(ref, tag, msh, quad, N):
    """The fixture's loop is advec1d.py:80-88's stage idiom around the reference's sw2dquads RHS function."""
    ctx = _ctx(msh, quad, N)
    h, hu, hv = (ref["sw2dquads_%s_%s" % (tag, n)].copy() for n in ("h", "hu", "hv"))
    res = [np.zeros_like(h) for _ in range(3)]
    _oracle(ctx, lf_mode=0, sources=0).step(h, hu, hv, res, 1000, float(ref["lserk4_1000_%s_dt" % tag]))
    for a, n in zip((h, hu, hv), ("h", "hu", "hv")):
        assert relerr(a, ref["lserk4_1000_%s_%s" % (tag, n)]) < STEP_TOL, n


def test_numpy_transcription_matches_tracer_equation(ref):
    """4th (tracer) equation of swhelpers/rhs.py -- SURVEY.md §8f rank 4; no kernel yet, the second derivation is held to
    the reference's output so the kernel can be built against both."""
    from oracle import sw2d_numpy

    ctx = _ctx("coarse_box.msh", False, 4)
    h, hu, hv, H, hN = (ref["swhelpers_" + n] for n in ("h", "hu", "hv", "H", "hN"))
    R = sw2d_numpy.compute_rhs(ctx, h, hu, hv, g=G, lf_mode=0, sources=True, H=H, Hx=np.full_like(H, 0.5),
                               Hy=np.full_like(H, 0.25), CD=0.0, f=F_COR, hN=hN)
    assert len(R) == 4
    for c in range(4):
        assert relerr(R[c], ref["swhelpers_cd0_RHS%d" % (c + 1)]) < RHS_TOL, c


def test_oracle_tracer_matches_reference_rhs4(ref):
    """oracle_sw2d_rhs_tracer (C restatement of swhelpers/rhs.py:206-297) against RHS4 of the reference's own function;
    RHS4 does not see the sources, so both fixture variants (with / without drag) must give it."""
    ctx = _ctx("coarse_box.msh", False, 4)
    h, hu, hv, hN = (ref["swhelpers_" + n] for n in ("h", "hu", "hv", "hN"))
    R4 = _oracle(ctx).rhs_tracer(h, hu, hv, hN)
    for tag in ("cd0", "cd"):
        assert relerr(R4, ref["swhelpers_%s_RHS4" % tag]) < RHS_TOL, tag


def test_limiter_restatement_matches_reference_function(ref):
    from oracle.output_numpy import positivity_limiter

    out = positivity_limiter(ref["limiter_h_in"], ref["swhelpers_hu"], ref["swhelpers_hv"])
    for a, n in zip(out, ("h", "hu", "hv")):
        assert np.abs(a - ref["limiter_%s_out" % n]).max() < 1e-13


# ---------------------------------------------------------------------------------------------- GPU: the CUDA path
@pytest.mark.gpu
@pytest.mark.parametrize("tag,msh,quad,N", CASES)
def test_cuda_matches_sw2dquads_rhs(ref, tag, msh, quad, N):
    ctx = _ctx(msh, quad, N)
    h, hu, hv = (ref["sw2dquads_%s_%s" % (tag, n)] for n in ("h", "hu", "hv"))
    sol = _solver(ctx)
    sol.set_state(h, hu, hv)
    variants = kernel_variants(sol)
    for var in variants:
        sol.set_variant(var)
        R = sol.computeRHS()
        for c in range(3):
            assert relerr(R[c], ref["sw2dquads_%s_RHS%d" % (tag, c + 1)]) < RHS_TOL, (var, c)
    sol.close()


@pytest.mark.gpu
@pytest.mark.parametrize("tag,msh,quad,N", STEP_CASES)
def test_cuda_matches_1000_lserk4_steps_of_reference_rhs(ref, tag, msh, quad, N):
    ctx = _ctx(msh, quad, N)
    h, hu, hv = (ref["sw2dquads_%s_%s" % (tag, n)] for n in ("h", "hu", "hv"))
    sol = _solver(ctx)
    variants = kernel_variants(sol)
    for var in variants:
        sol.set_variant(var)
        sol.set_state(h, hu, hv)
        sol.step(1000, float(ref["lserk4_1000_%s_dt" % tag]))
        out = sol.get_state()
        for a, n in zip(out[:3], ("h", "hu", "hv")):
            assert relerr(a, ref["lserk4_1000_%s_%s" % (tag, n)]) < STEP_TOL, (var, n)
    sol.close()


@pytest.mark.gpu
@pytest.mark.parametrize("tag,CD", [("cd0", 0.0), ("cd", 2.5e-3)])
def test_cuda_matches_swhelpers_rhs_with_sources(ref, tag, CD):
    ctx = _ctx("coarse_box.msh", False, 4)
    h, hu, hv, H = (ref["swhelpers_" + n] for n in ("h", "hu", "hv", "H"))
    sol = _solver(ctx, phys=_phys(CD), mode="full", lf_mode=0)
    sol.set_bathymetry(H, np.full_like(H, 0.5), np.full_like(H, 0.25))
    sol.set_state(h, hu, hv, 0.0)
    R = sol.computeRHS()
    want = [ref["swhelpers_%s_RHS%d" % (tag, c + 1)] for c in range(3)]
    want[2] = want[2] + _drag_sign_fix(ref, CD)
    for c in range(3):
        assert relerr(R[c], want[c]) < RHS_TOL, c
    sol.close()


@pytest.mark.gpu
def test_cuda_tracer_matches_reference_rhs4_and_oracle_steps(ref):
    """The tracer kernel (sw2d_tracer.cu) through the C ABI: RHS4 against the reference function's output, then LSERK4 and
    SSP-RK2 steps of flow + tracer against the oracle, on both flow kernel variants (the tracer launch reads the pre-stage
    flow state beside either)."""
    from conftest import kernel_variants

    ctx = _ctx("coarse_box.msh", False, 4)
    h, hu, hv, hN = (ref["swhelpers_" + n] for n in ("h", "hu", "hv", "hN"))
    sol = _solver(ctx)
    orc = _oracle(ctx, CFL=0.65)
    dt = 0.5 * orc.timestep(h, hu, hv, use_abs=True)[0]
    q = [np.array(a) for a in (h, hu, hv, hN)]
    res = [np.zeros_like(h) for _ in range(4)]
    orc.step_with_tracer(q[0], q[1], q[2], q[3], res, 40, dt)
    for var in kernel_variants(sol):
        sol.set_variant(var)
        sol.set_state(h, hu, hv)
        sol.set_tracer(hN)
        assert relerr(sol.computeRHS4(), ref["swhelpers_cd0_RHS4"]) < RHS_TOL, var
        sol.step(40, dt)
        got = list(sol.get_state()[:3]) + [sol.get_tracer()]
        for a, b, n in zip(got, q, ("h", "hu", "hv", "hN")):
            assert relerr(a, b) < STEP_TOL, (var, n)
        # tracer mass is conserved with all-wall BC (its flux vanishes with the normal velocity)
        w = np.linalg.inv(ctx.V @ ctx.V.T).sum(axis=0)
        m0, m1 = (w[:, None] * ctx.J * hN).sum(), (w[:, None] * ctx.J * got[3]).sum()
        assert abs(m1 - m0) < 1e-11 * abs(m0)
        # the src/sw2d loop form: hN1 = hN + dt RHS4(q); hN = (hN + hN1 + dt RHS4(q1, hN1)) / 2
        sol.set_state(h, hu, hv)
        sol.set_tracer(hN)
        sol.step_ssprk2(10, dt)
        a = [np.array(x) for x in (h, hu, hv, hN)]
        for _ in range(10):
            R, R4 = orc.rhs(a[0], a[1], a[2]), orc.rhs_tracer(a[0], a[1], a[2], a[3])
            b = [x + dt * r for x, r in zip(a, R + [R4])]
            R, R4 = orc.rhs(b[0], b[1], b[2]), orc.rhs_tracer(b[0], b[1], b[2], b[3])
            a = [0.5 * ((x + y) + dt * r) for x, y, r in zip(a, b, R + [R4])]
        got = list(sol.get_state()[:3]) + [sol.get_tracer()]
        for x, y, n in zip(got, a, ("h", "hu", "hv", "hN")):
            assert relerr(x, y) < STEP_TOL, (var, n)
    sol.set_tracer(None)
    with pytest.raises(RuntimeError):
        sol.get_tracer()
    sol.close()
    full = _solver(ctx, mode="full")  # global LF: not the reference's tracer form
    with pytest.raises(RuntimeError):
        full.set_tracer(hN)
    full.close()


@pytest.mark.gpu
def test_cuda_limiter_matches_reference_function(ref):
    ctx = _ctx("coarse_box.msh", False, 4)
    sol = _solver(ctx)
    sol.set_state(ref["limiter_h_in"], ref["swhelpers_hu"], ref["swhelpers_hv"])
    sol.positivityPreservingLimiter2D()
    out = sol.get_state()
    for a, n in zip(out[:3], ("h", "hu", "hv")):
        assert np.abs(a - ref["limiter_%s_out" % n]).max() < 1e-13
    sol.close()
