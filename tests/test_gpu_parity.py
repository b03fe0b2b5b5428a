"""GPU parity tests: the CUDA path through the C ABI against the CPU oracle on the same seeded inputs.

Tolerances are BASELINE.json's: relative 1e-12 (max-norm, per field) per RHS evaluation, 1e-10 after
1000 LSERK4 steps.  Run on the B200 box: python -m pytest tests -m gpu
"""
import numpy as np
import pytest

from blitzdg_b200 import QuadNodesProvisioner, TriangleNodesProvisioner
from blitzdg_b200.initial_conditions import gaussian_hump
from blitzdg_b200.partition import build_local
from blitzdg_b200.synthetic import provision, structured_mesh

pytestmark = pytest.mark.gpu

RHS_TOL = 1e-12
STEP_TOL = 1e-10


def relerr(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


def _solver(*a, **k):
    from blitzdg_b200.sw2d import SW2dSolver

    return SW2dSolver(*a, **k)


def _oracle(*a, **k):
    from oracle.binding import Oracle

    return Oracle(*a, **k)


def _ctx(mesh, N, quad=False, filt=False):
    p = (QuadNodesProvisioner if quad else TriangleNodesProvisioner)(N, mesh)
    if filt:
        p.buildFilter(0.9 * N, 4)
    return p.get_DGContext()


def _state(ctx, seed=0, amp=0.3):
    H, h, hu, hv = gaussian_hump(ctx.x, ctx.y)
    rng = np.random.default_rng(seed)
    return H, h, amp * rng.standard_normal(h.shape), amp * rng.standard_normal(h.shape)


from conftest import kernel_variants as _variants  # noqa: E402


@pytest.mark.parametrize("quad,N", [(False, 1), (False, 2), (False, 3), (False, 4), (False, 5), (False, 8), (True, 3), (True, 6)])
def test_rhs_simple_mode(coarse_box_mesh, coarse_quads_mesh, quad, N):
    ctx = _ctx(coarse_quads_mesh if quad else coarse_box_mesh, N, quad)
    H, h, hu, hv = _state(ctx, seed=N)
    Ro = _oracle(ctx).rhs(h, hu, hv)
    sol = _solver(ctx)
    sol.set_state(h, hu, hv)
    for var in _variants(sol):
        sol.set_variant(var)
        R = sol.computeRHS()
        for c in range(3):
            assert relerr(R[c], Ro[c]) < RHS_TOL, (var, c, relerr(R[c], Ro[c]))
    sol.close()


@pytest.mark.parametrize("quad,N", [(False, 4), (True, 3)])
def test_rhs_full_mode_with_open_boundary(coarse_box_mesh, coarse_quads_mesh, quad, N):
    ctx = _ctx(coarse_quads_mesh if quad else coarse_box_mesh, N, quad)
    H = 10 + 0.5 * ctx.x + 0.25 * ctx.y
    Hx, Hy = np.full_like(H, 0.5), np.full_like(H, 0.25)
    _, h, hu, hv = _state(ctx, seed=7, amp=0.5)
    h = H + (h - 10.0)
    wall = ctx.BCmap[3]
    nfp = ctx.Nfp
    nface = len(wall) // nfp
    ctx.BCmap = {3: wall[: (nface // 2) * nfp], 2: wall[(nface // 2) * nfp:]}
    orc = _oracle(ctx, lf_mode=1, sources=1)
    orc.set_bathymetry(H, Hx, Hy)
    t = 4321.0
    Ro = orc.rhs(h, hu, hv, t=t)
    sol = _solver(ctx, mode="full")
    sol.set_bathymetry(H, Hx, Hy)
    sol.set_state(h, hu, hv, t=t)
    for var in _variants(sol):
        sol.set_variant(var)
        R = sol.computeRHS()
        for c in range(3):
            assert relerr(R[c], Ro[c]) < RHS_TOL, (var, c, relerr(R[c], Ro[c]))
    sol.close()


def test_rhs_with_filter(coarse_box_mesh):
    ctx = _ctx(coarse_box_mesh, 4, filt=True)
    H, h, hu, hv = _state(ctx, seed=3)
    Ro = _oracle(ctx, filter=1).rhs(h, hu, hv)
    sol = _solver(ctx, use_filter=True)
    sol.set_state(h, hu, hv)
    R = sol.computeRHS()
    for c in range(3):
        assert relerr(R[c], Ro[c]) < RHS_TOL
    sol.close()


@pytest.mark.parametrize("variant", [1, 2])
def test_1000_lserk4_steps_coarse_box_n4(coarse_box_mesh, variant):
    """BASELINE config 1: coarse_box.msh, N=4, LSERK4, fixed dt from the reference formula."""
    ctx = _ctx(coarse_box_mesh, 4)
    H, h, hu, hv = gaussian_hump(ctx.x, ctx.y)
    orc = _oracle(ctx, CFL=0.65)
    dt, _ = orc.timestep(h, hu, hv, use_abs=True)
    dt *= 0.5
    sol = _solver(ctx)
    sol.set_variant(variant)  # N=4 affine triangles: both variants must exist
    sol.set_bathymetry(H)
    sol.set_state(h, hu, hv)
    sol.step(1000, dt)
    ho, huo, hvo = h.copy(), hu.copy(), hv.copy()
    res = [np.zeros_like(h) for _ in range(3)]
    to = orc.step(ho, huo, hvo, res, 1000, dt)
    hg, hug, hvg, t = sol.get_state()
    assert abs(t - to) <= 1e-12 * to
    errs = [relerr(hg, ho), relerr(hug, huo), relerr(hvg, hvo)]
    assert max(errs) < STEP_TOL, errs
    assert sol.check() < 1.5
    sol.close()


def test_timestep_and_adaptive_stepping(coarse_box_mesh):
    ctx = _ctx(coarse_box_mesh, 4)
    H, h, hu, hv = _state(ctx, seed=5, amp=0.1)
    orc = _oracle(ctx, CFL=0.25)
    dto, _ = orc.timestep(h, hu, hv)
    sol = _solver(ctx)
    sol.set_state(h, hu, hv)
    assert abs(sol.computeTimeStep() - dto) <= 1e-14 * dto
    sol.step(20, 0.0)  # adaptive: dt recomputed on device every step
    ho, huo, hvo = h.copy(), hu.copy(), hv.copy()
    res = [np.zeros_like(h) for _ in range(3)]
    to = orc.step(ho, huo, hvo, res, 20, 0.0)
    hg, hug, hvg, t = sol.get_state()
    assert abs(t - to) <= 1e-12 * to
    assert max(relerr(hg, ho), relerr(hug, huo), relerr(hvg, hvo)) < STEP_TOL
    sol.close()


def test_blow_up_guard(coarse_box_mesh):
    from blitzdg_b200.sw2d import NumericalInstability

    ctx = _ctx(coarse_box_mesh, 2)
    H, h, hu, hv = gaussian_hump(ctx.x, ctx.y)
    sol = _solver(ctx)
    sol.set_bathymetry(H)
    sol.set_state(h, hu, hv)
    assert sol.check() == pytest.approx(np.abs(h - H).max(), rel=1e-15)
    h2 = h.copy()
    h2[0, 0] = np.nan
    sol.set_state(h2, hu, hv)
    with pytest.raises(NumericalInstability):
        sol.check()
    h2[0, 0] = 1e9
    sol.set_state(h2, hu, hv)
    with pytest.raises(NumericalInstability):
        sol.check()
    sol.close()


@pytest.mark.parametrize("shape,N,nx,ny,jit", [("tri", 4, 64, 32, 0.0), ("tri", 4, 48, 24, 0.2), ("tri", 8, 24, 12, 0.0),
                                              ("quad", 6, 32, 16, 0.2),
                                              # ragged element counts (K % 8 != 0, K % 32 != 0): DMMA column padding,
                                              # partial tiles, every row-block count of the generic DMMA kernels
                                              ("tri", 6, 9, 7, 0.0), ("tri", 3, 11, 5, 0.15), ("tri", 5, 7, 5, 0.1),
                                              ("quad", 4, 9, 7, 0.2), ("quad", 7, 5, 5, 0.1), ("quad", 2, 7, 3, 0.2)])
def test_synthetic_mesh_rhs_and_steps(shape, N, nx, ny, jit):
    mesh = structured_mesh(shape, nx, ny, jitter=jit)
    ctx = provision(mesh, N).get_DGContext()
    H, h, hu, hv = _state(ctx, seed=11, amp=0.2)
    orc = _oracle(ctx, CFL=0.65)
    dt = 0.5 * orc.timestep(h, hu, hv, use_abs=True)[0]
    Ro = orc.rhs(h, hu, hv)
    sol = _solver(ctx)
    sol.set_state(h, hu, hv)
    for var in _variants(sol):
        sol.set_variant(var)
        sol.set_state(h, hu, hv)
        R = sol.computeRHS()
        for c in range(3):
            assert relerr(R[c], Ro[c]) < RHS_TOL, (var, c, relerr(R[c], Ro[c]))
        sol.step(20, dt)
        ho, huo, hvo = h.copy(), hu.copy(), hv.copy()
        res = [np.zeros_like(h) for _ in range(3)]
        orc.step(ho, huo, hvo, res, 20, dt)
        hg, hug, hvg, _ = sol.get_state()
        assert max(relerr(hg, ho), relerr(hug, huo), relerr(hvg, hvo)) < STEP_TOL, var
    sol.close()


@pytest.mark.parametrize("N,nx,ny,bed", [(4, 64, 32, "slope"), (4, 40, 24, "steps"), (3, 33, 17, "steps"), (5, 20, 12, "slope"), (2, 50, 30, "steps")])
def test_full_physics_fused_on_structured_meshes(N, nx, ny, bed):
    """src/sw2d physics (global LF, star states, Out faces with tide, sources) through the fused single-launch kernel on
    meshes with many tiles (persistent + edge kernels): RHS, LSERK4 steps with the global speed carried from stage to stage,
    then the driver's SSP-RK2 loop with a sponge, all against the oracle (which agrees bitwise with the reference's own
    compiled computeRHS, tests/test_reference_compiled.py).  bed = steps: a different bed level per element, so that the star
    states differ from the plain traces on every interior face."""
    mesh = structured_mesh("tri", nx, ny)
    ctx = provision(mesh, N).get_DGContext()
    _, h, hu, hv = _state(ctx, seed=N, amp=0.3)
    rng = np.random.default_rng(40 + N)
    H = 12 + 0.5 * ctx.x + 0.25 * ctx.y
    if bed == "steps":
        H = H + 0.2 * rng.standard_normal(ctx.K)[None, :]
    Hx, Hy = np.full_like(H, 0.5), np.full_like(H, 0.25)
    h = np.ascontiguousarray(H + (h - 10.0))
    wall = np.asarray(ctx.BCmap[3]).reshape(-1, ctx.Nfp)
    yv = ctx.y.T.reshape(-1)[np.asarray(ctx.vmapM)[wall]]
    top = (yv > 1.0 - 1e-9).all(axis=1)
    ctx.BCmap = {3: wall[~top].reshape(-1), 2: wall[top].reshape(-1)}  # open top edge
    orc = _oracle(ctx, lf_mode=1, sources=1, CFL=0.25)
    orc.set_bathymetry(H, Hx, Hy)
    t0 = 3600 * 12.42 + 500.0
    Ro = orc.rhs(h, hu, hv, t=t0)
    dt = 0.5 * orc.timestep(h, hu, hv)[0]
    ho, huo, hvo = h.copy(), hu.copy(), hv.copy()
    res = [np.zeros_like(h) for _ in range(3)]
    t1 = orc.step(ho, huo, hvo, res, 12, dt, t=t0)
    sponge = 30.0 * rng.random(h.shape)
    h2, hu2, hv2 = ho.copy(), huo.copy(), hvo.copy()
    t2 = orc.step_ssprk2(h2, hu2, hv2, 10, dt, t=t1, sponge=sponge)
    sol = _solver(ctx, mode="full")
    sol.set_bathymetry(H, Hx, Hy)
    sol.set_sponge(sponge)
    assert _variants(sol) == [1, 2]
    for var in (1, 2):
        sol.set_variant(var)
        sol.set_state(h, hu, hv, t=t0)
        R = sol.computeRHS()
        for c in range(3):
            assert relerr(R[c], Ro[c]) < RHS_TOL, (var, c, relerr(R[c], Ro[c]))
        sol.step(12, dt)
        hg, hug, hvg, tg = sol.get_state()
        assert abs(tg - t1) <= 1e-12 * t1
        assert max(relerr(hg, ho), relerr(hug, huo), relerr(hvg, hvo)) < STEP_TOL, var
        R = sol.computeRHS()  # an RHS between steps must not disturb the carried speed
        sol.step_ssprk2(10, dt)
        hg, hug, hvg, tg = sol.get_state()
        assert abs(tg - t2) <= 1e-12 * t2
        assert max(relerr(hg, h2), relerr(hug, hu2), relerr(hvg, hv2)) < STEP_TOL, var
        sol.check()
    sol.close()


@pytest.mark.parametrize("nx,ny,P,how", [(16, 12, 2, "strips"), (80, 64, 2, "metis"), (96, 80, 2, "metis"), (60, 44, 3, "metis")])
def test_partitioned_halo_emulated_on_one_gpu(nx, ny, P, how):
    """Element-partitioned RHS with every rank's ghost traces filled through the set_ghosts hook equals the single-rank
    oracle (each element's arithmetic is independent of the partition).  The METIS cases give local element counts that are
    NOT multiples of the 32-element tile with the cut elements in the ragged last tile -- there the halo slots Np*K + g lie
    inside the tile's padded node range and must still be treated as external (a round-2 bug found by the 2-GPU tests)."""
    from blitzdg_b200.partition import make_partition

    mesh = structured_mesh("tri", nx, ny)
    N = 4
    ctx = provision(mesh, N).get_DGContext()
    H, h, hu, hv = _state(ctx, seed=13)
    Ro = _oracle(ctx).rhs(h, hu, hv)
    K = ctx.K
    epart, _ = make_partition(mesh, P, how, seed=5)
    parts = [build_local(mesh, epart, r, N) for r in range(P)]
    if how == "metis":
        assert any(p[0].K % 32 for p in parts)
    for c, halo, info in parts:
        s = _solver(c, halo=halo)
        el = info["elements"]
        s.set_state(h[:, el], hu[:, el], hv[:, el])
        # ghosts in the peers' send-buffer layout (per neighbour segment: 3 planes of cnt values), read from the global state
        gk = info["ghost_keys"]
        gn, ge = gk % c.Np, (gk // c.Np) % K
        ro = halo["recv_offset"]
        buf = np.zeros(3 * len(gk))
        for i in range(len(halo["nbr_rank"])):
            a, b = int(ro[i]), int(ro[i + 1])
            for f, arr in enumerate((h, hu, hv)):
                buf[3 * a + f * (b - a): 3 * a + (f + 1) * (b - a)] = arr[gn[a:b], ge[a:b]]
        s.halo_set_ghosts(buf)
        if P == 2 and how == "strips":  # the pack hook delivers the same values the peer would send
            other = parts[1 - halo["rank"]]
            so = _solver(other[0], halo=other[1])
            so.set_state(h[:, other[2]["elements"]], hu[:, other[2]["elements"]], hv[:, other[2]["elements"]])
            assert np.array_equal(so.halo_pack_to_host(len(other[1]["send_nodes"])), buf)
            so.close()
        for var in _variants(s):
            s.set_variant(var)
            R = s.computeRHS()
            for cc in range(3):
                assert relerr(R[cc], Ro[cc][:, el]) < RHS_TOL, (halo["rank"], var, cc)
        s.close()


# BASELINE's full-size configs (2-5): tests/test_fullsize_parity.py (whole-mesh oracle + extended-precision arbiter)


def test_streaming_state_calls_equal_serial_calls():
    """stage_state / commit_state / get_state_async / get_state_wait: a batch of independent states pushed through the
    overlapped calls gives bitwise what set_state -> step -> get_state gives, in order; misuse is an error."""
    import torch

    ctx = provision(structured_mesh("tri", 48, 20), 4).get_DGContext()
    _, h, hu, hv = _state(ctx, seed=1, amp=0.1)
    sol = _solver(ctx)
    sol.set_state(h, hu, hv)
    dt = 0.5 * sol.computeTimeStep()
    batch = []
    for i in range(4):
        rng = np.random.default_rng(100 + i)
        batch.append([np.ascontiguousarray(a) for a in (h + 0.01 * i, 0.1 * rng.standard_normal(h.shape), 0.1 * rng.standard_normal(h.shape))])
    want = []
    for b in batch:
        sol.set_state(*b, t=1.5)
        sol.step(7, dt)
        want.append(sol.get_state())
    pin = lambda: [torch.empty(h.shape, dtype=torch.float64).pin_memory().numpy() for _ in range(3)]
    src, out = [pin() for _ in batch], [pin() for _ in batch]
    for s_, b in zip(src, batch):
        for a, c in zip(s_, b):
            a[...] = c
    with pytest.raises(RuntimeError):
        sol.commit_state()
    sol.stage_state(*src[0], t=1.5)
    with pytest.raises(RuntimeError):
        sol.stage_state(*src[1], t=1.5)
    for i in range(len(batch)):
        sol.commit_state()
        if i + 1 < len(batch):
            sol.stage_state(*src[i + 1], t=1.5)
        sol.step(7, dt, sync=False)
        sol.get_state_async(out[i])
    sol.get_state_wait()
    for i in range(len(batch)):
        for c in range(3):
            assert np.array_equal(out[i][c], want[i][c]), (i, c)
    assert sol.get_state()[3] == want[-1][3]
    sol.close()
