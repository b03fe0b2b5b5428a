"""The rest of the src/sw2d driver around computeRHS: open boundary + sponge + bathymetry,
advanced by the driver's own SSP-RK2 loop (reference src/sw2d/main.cpp:192-244).

CPU tests pin the oracle's restatement of that loop and the host-side setup helpers against it; the GPU tests
compare bdg_sw2d_step_ssprk2 / bdg_sw2d_set_sponge with the oracle on the same inputs (1e-10 after 100+ steps).
"""
import os
import subprocess
import sys

import numpy as np
import pytest

from blitzdg_b200 import QuadNodesProvisioner, TriangleNodesProvisioner
from blitzdg_b200.initial_conditions import gaussian_hump
from blitzdg_b200.sw2d import bedSlopes, buildSpongeCoeff, readDepthData
from blitzdg_b200.synthetic import provision, structured_mesh

STEP_TOL = 1e-10
TIDE_T = 3600 * 12.42


def relerr(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


def _oracle(*a, **k):
    from oracle.binding import Oracle

    return Oracle(*a, **k)


def _ctx(mesh, N, quad=False, filt=True):
    p = (QuadNodesProvisioner if quad else TriangleNodesProvisioner)(N, mesh)
    if filt and not quad:  # the quad buildFilter of the reference is defective
        p.buildFilter(0.9 * N, 4)
    return p.get_DGContext()


def coastal_case(ctx, seed=3):
    """A small basin: the second half of the boundary faces is open (Out), depth 150-200 m with a slope, a sponge
    of radius 0.5 behind the open boundary, a perturbed state so that every term of the loop is exercised."""
    wall = np.asarray(ctx.BCmap[3])
    nface = len(wall) // ctx.Nfp
    ctx.BCmap = {3: wall[: (nface // 2) * ctx.Nfp], 2: wall[(nface // 2) * ctx.Nfp:]}
    H = 175.0 + 20.0 * ctx.x + 5.0 * ctx.y * ctx.y
    Hx, Hy = bedSlopes(ctx, H)
    sponge = buildSpongeCoeff(ctx, 10000.0, 0.5)
    rng = np.random.default_rng(seed)
    _, h0, _, _ = gaussian_hump(ctx.x, ctx.y)
    h = H + (h0 - 10.0)
    hu = 2.0 * rng.standard_normal(h.shape)
    hv = 2.0 * rng.standard_normal(h.shape)
    return H, Hx, Hy, sponge, h, hu, hv


# ------------------------------------------------------------------------------------------------- CPU
def test_oracle_ssprk2_is_heun_with_sponge(coarse_box_mesh):
    """One step of the restated loop = the formulas of main.cpp:213-236 written out with the oracle's RHS."""
    ctx = _ctx(coarse_box_mesh, 3)
    H, Hx, Hy, sponge, h, hu, hv = coastal_case(ctx)
    orc = _oracle(ctx, lf_mode=1, sources=1, CFL=0.25)
    orc.set_bathymetry(H, Hx, Hy)
    t0 = TIDE_T
    dt, _ = orc.timestep(h, hu, hv)
    R = orc.rhs(h, hu, hv, t=t0)
    q1 = [h + dt * R[0], hu + dt * R[1], hv + dt * R[2]]
    q1[1] = q1[1] / (1.0 + sponge * q1[1] * q1[1])
    q1[2] = q1[2] / (1.0 + sponge * q1[2] * q1[2])
    R1 = orc.rhs(*q1, t=t0)
    q2 = [0.5 * (a + b + dt * c) for a, b, c in zip((h, hu, hv), q1, R1)]
    q2[1] = q2[1] / (1.0 + sponge * q2[1] * q2[1])
    q2[2] = q2[2] / (1.0 + sponge * q2[2] * q2[2])
    ho, huo, hvo = h.copy(), hu.copy(), hv.copy()
    t1 = orc.step_ssprk2(ho, huo, hvo, 1, 0.0, t=t0, sponge=sponge)
    assert t1 == t0 + dt
    for a, b in zip((ho, huo, hvo), q2):
        assert np.array_equal(a, b)
    assert np.abs(sponge).max() > 0 and relerr(huo, hu) > 1e-3  # the sponge and the forcing did act


def test_oracle_ssprk2_second_order(coarse_box_mesh):
    """Heun's scheme converges with order 2 in dt (sw2d-simple physics, no sponge)."""
    ctx = _ctx(coarse_box_mesh, 3, filt=False)
    _, h, hu, hv = gaussian_hump(ctx.x, ctx.y)
    orc = _oracle(ctx, CFL=0.25)
    dt0, _ = orc.timestep(h, hu, hv)

    def run(nsub):
        a, b, c = h.copy(), hu.copy(), hv.copy()
        orc.step_ssprk2(a, b, c, 16 * nsub, dt0 / nsub)
        return a

    ref = run(16)
    e1, e2, e4 = (np.abs(run(n) - ref).max() for n in (1, 2, 4))
    assert 3.3 < e1 / e2 < 4.8 and 3.3 < e2 / e4 < 4.8, (e1, e2, e4)


def test_sponge_builder_matches_reference_scan(coarse_box_mesh, coarse_quads_mesh):
    for mesh, quad in ((coarse_box_mesh, False), (coarse_quads_mesh, True)):
        ctx = _ctx(mesh, 4, quad)
        coastal_case(ctx)
        orc = _oracle(ctx)
        for rad in (0.3, 0.75, 5.0):
            got = buildSpongeCoeff(ctx, 10000.0, rad)
            want = orc.build_sponge(ctx.x, ctx.y, ctx.BCmap[2], 10000.0, rad)
            assert np.abs(got - want).max() <= 1e-9 * 10000.0
            assert got.max() == 10000.0 and got.min() >= 0.0  # nodes on the open boundary carry the full strength
        if rad == 5.0:
            assert (got > 0).all()
    ctx.BCmap = {3: ctx.BCmap[3]}
    assert not buildSpongeCoeff(ctx, 10000.0, 0.5).any()  # no Out faces: no sponge


def test_depth_reader(tmp_path, coarse_box_mesh):
    ctx = _ctx(coarse_box_mesh, 2)
    Np, K = ctx.Np, ctx.K
    rng = np.random.default_rng(0)
    vals = rng.uniform(20.0, 400.0, Np * K - 7)  # shallow values get floored, the last 7 rows are missing
    f = tmp_path / "H0.oct"
    f.write_text("".join("%.17g\n" % v for v in vals))
    H = readDepthData(str(f), Np, K)
    from oracle.binding import Oracle

    want = Oracle.depth_to_H(vals, Np, K)
    assert np.array_equal(H, want)
    assert H.min() == 150.0 and H[Np - 1, K - 1] == 150.0
    assert H[1, 0] == max(vals[1], 150.0) and H[0, 1] == max(vals[Np], 150.0)  # element-major file order
    f.write_text("12.5\nnot-a-number\n")
    with pytest.raises(RuntimeError):
        readDepthData(str(f), Np, K)


def test_bed_slopes(coarse_box_mesh):
    ctx = _ctx(coarse_box_mesh, 4)
    H = 175.0 + 20.0 * ctx.x - 7.0 * ctx.y
    Hx, Hy = bedSlopes(ctx, H)
    assert np.abs(Hx - 20.0).max() < 1e-9 and np.abs(Hy + 7.0).max() < 1e-9  # a linear bed is differentiated exactly
    H = 175.0 + 20.0 * np.sin(ctx.x) * ctx.y
    got = bedSlopes(ctx, H)
    want = _oracle(ctx).bed_slopes(H)
    assert relerr(got[0], want[0]) < 1e-12 and relerr(got[1], want[1]) < 1e-12


# ------------------------------------------------------------------------------------------------- GPU
def _solver(*a, **k):
    from blitzdg_b200.sw2d import SW2dSolver

    return SW2dSolver(*a, **k)


def _run_coastal(ctx, nsteps, dt):
    H, Hx, Hy, sponge, h, hu, hv = coastal_case(ctx)
    orc = _oracle(ctx, lf_mode=1, sources=1, CFL=0.25)
    orc.set_bathymetry(H, Hx, Hy)
    ho, huo, hvo = h.copy(), hu.copy(), hv.copy()
    to = orc.step_ssprk2(ho, huo, hvo, nsteps, dt, t=TIDE_T, sponge=sponge)
    from conftest import kernel_variants

    sol = _solver(ctx, mode="full")
    sol.set_bathymetry(H, Hx, Hy)
    sol.set_sponge(sponge)
    launches = {}
    for var in kernel_variants(sol):  # [1, 2] on affine triangles N <= 5: the fused single-launch kernel must exist there
        sol.set_variant(var)
        sol.set_state(h, hu, hv, t=TIDE_T)
        n0 = sol.launch_count
        sol.step_ssprk2(nsteps, dt)
        launches[var] = sol.launch_count - n0
        hg, hug, hvg, t = sol.get_state()
        sol.check()
        err = max(relerr(hg, ho), relerr(hug, huo), relerr(hvg, hvo))
        assert abs(t - to) <= 1e-12 * to, (var, t, to)
        assert err < STEP_TOL, (var, err)
    sol.close()
    assert relerr(huo, hu) > 1e-3
    return launches


@pytest.mark.gpu
@pytest.mark.parametrize("quad,N,dt", [(False, 4, 0.0), (False, 4, 2e-5), (False, 8, 0.0), (True, 3, 0.0), (True, 6, 0.0)])
def test_ssprk2_coastal_workflow(coarse_box_mesh, coarse_quads_mesh, quad, N, dt):
    """Out faces with the tidal forcing + star states + bathymetry/drag/Coriolis + sponge, 120 steps of the driver loop."""
    ctx = _ctx(coarse_quads_mesh if quad else coarse_box_mesh, N, quad)
    launches = _run_coastal(ctx, 120, dt)
    assert launches[1] >= 120 * 4  # two-launch path: (speed +) face + volume per stage: the loop ran on the device
    if 2 in launches:  # fused full physics: ONE launch per stage (+ time advance, + the dt reduction in adaptive mode), no k_speed
        assert 120 * 2 <= launches[2] < launches[1], launches  # (the 40-element box adds one edge-kernel launch for its ragged tile)


@pytest.mark.gpu
@pytest.mark.parametrize("shape,N,nx,ny,jit", [("tri", 4, 40, 20, 0.0), ("tri", 4, 24, 12, 0.2), ("quad", 6, 16, 8, 0.2), ("tri", 8, 12, 6, 0.0)])
def test_ssprk2_simple_physics_then_lserk4(shape, N, nx, ny, jit):
    """sw2d-simple physics under the SSP-RK2 loop (no sponge), followed by LSERK4 steps on the same handle: the two
    steppers share the state buffers and the residual planes."""
    ctx = provision(structured_mesh(shape, nx, ny, jit), N).get_DGContext()
    _, h, hu, hv = gaussian_hump(ctx.x, ctx.y)
    orc = _oracle(ctx, CFL=0.25)
    dt, _ = orc.timestep(h, hu, hv)
    ho, huo, hvo = h.copy(), hu.copy(), hv.copy()
    res = [np.zeros_like(h) for _ in range(3)]
    t = orc.step_ssprk2(ho, huo, hvo, 25, dt)
    t = orc.step(ho, huo, hvo, res, 10, dt, t=t)
    t = orc.step_ssprk2(ho, huo, hvo, 5, dt, t=t)
    sol = _solver(ctx)
    sol.set_state(h, hu, hv)
    sol.step_ssprk2(25, dt)
    sol.step(10, dt)
    sol.step_ssprk2(5, dt)
    hg, hug, hvg, tg = sol.get_state()
    sol.close()
    assert abs(tg - t) <= 1e-12 * t
    assert max(relerr(hg, ho), relerr(hug, huo), relerr(hvg, hvo)) < STEP_TOL


@pytest.mark.gpu
def test_ssprk2_thread_per_node_kernels():
    """The study switches select the scalar kernels of the two-launch path (read once per process): same loop, same answer."""
    env = dict(os.environ, BDG_VOLUME_SCALAR="1", BDG_FACE_SCALAR="1")
    code = ("import sys\n"
            "sys.path[:0] = ['.', 'tests']\n"
            "import test_coastal_workflow as T\n"
            "from blitzdg_b200 import MeshManager\n"
            "m = MeshManager()\n"
            "m.readMesh('tests/golden/coarse_box.msh')\n"
            "print('launches', T._run_coastal(T._ctx(m, 4), 60, 0.0))\n")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], cwd=root, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "launches" in r.stdout, r.stdout + r.stderr


# ---- positivity limiter of the Python drivers (swhelpers/limiters.py:3-21) --------------------
def test_oracle_limiter_properties(coarse_box_mesh):
    """What the reference function does as written: the floor `hmin[hmin < 1e-3] = 1e-3` makes theta >= 1 (clipped to 1,
    i.e. no limiting) whenever the element mean exceeds ~9e-4; only nearly dry elements are contracted.  Element means
    are always kept.  The device kernel reproduces exactly this (test_device_limiter_matches_oracle)."""
    from oracle.output_numpy import positivity_limiter

    ctx = _ctx(coarse_box_mesh, 4, filt=False)
    rng = np.random.default_rng(0)
    h = 0.05 + 0.06 * rng.standard_normal(ctx.x.shape)
    h[:, 1::2] = 5e-4 + 2e-4 * rng.standard_normal((ctx.Np, h[:, 1::2].shape[1]))  # nearly dry elements
    hu, hv = rng.standard_normal(h.shape), rng.standard_normal(h.shape)
    a, b, c = positivity_limiter(h, hu, hv)
    wet = h.mean(axis=0) > 1e-3
    assert wet.any() and (~wet).any()
    assert np.abs(a[:, wet] - h[:, wet]).max() < 1e-15 and np.abs(b[:, wet] - hu[:, wet]).max() < 1e-14
    assert np.abs(a[:, ~wet] - h[:, ~wet]).max() > 1e-5  # theta != 1 there
    for x, y in ((a, h), (b, hu), (c, hv)):
        assert np.abs(x.mean(axis=0) - y.mean(axis=0)).max() < 1e-13


@pytest.mark.gpu
@pytest.mark.parametrize("quad,N", [(False, 4), (False, 8), (True, 3), (True, 6)])
def test_device_limiter_matches_oracle(coarse_box_mesh, coarse_quads_mesh, quad, N):
    from oracle.output_numpy import positivity_limiter

    ctx = _ctx(coarse_quads_mesh if quad else coarse_box_mesh, N, quad, filt=False)
    rng = np.random.default_rng(N)
    h = 0.05 + 0.06 * rng.standard_normal(ctx.x.shape)
    h[:, ::3] += 1.0
    h[:, 1::3] = 5e-4 + 2e-4 * rng.standard_normal((ctx.Np, h[:, 1::3].shape[1]))  # nearly dry: theta != 1
    hu, hv = rng.standard_normal(h.shape), rng.standard_normal(h.shape)
    want = positivity_limiter(h, hu, hv)
    sol = _solver(ctx)
    sol.set_state(h, hu, hv)
    sol.positivityPreservingLimiter2D()
    got = sol.get_state()
    sol.close()
    assert np.abs(want[0] - h).max() > 1e-5
    for g, w in zip(got[:3], want):
        assert np.abs(g - w).max() <= 1e-12 * max(np.abs(w).max(), 1.0)
