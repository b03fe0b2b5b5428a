"""Parity at BASELINE.json's FULL sizes: the CUDA path through the C ABI against the CPU oracle over the WHOLE mesh,
with the extended-precision arbiter (oracle/sw2d_arbiter.c) as the yardstick.

What is measured, per config x kernel variant x field, and written to gpurun_out/parity_fullsize.json:
    cuda_vs_oracle    max|R_cuda   - R_oracle | / max|R_oracle|      (BASELINE's parity metric)
    cuda_vs_exact     max|R_cuda   - R_arbiter| / max|R_arbiter|
    oracle_vs_exact   max|R_oracle - R_arbiter| / max|R_arbiter|     (the reference arithmetic's own rounding error)
The volume term (src/sw2d/main.cpp:434-442) differentiates g h^2/2 ~ 490 with metric factors ~ 2/h_cell: its fp64
rounding error grows like 1/h_cell and passes 1e-12 of the result near 250x125 cells for ANY summation order -- the
reference's own included (oracle_vs_exact = 7.8e-12 on the 1 M-triangle mesh).  A kernel cannot be closer to the
reference's fp64 result than the reference is to the exact one unless it imitates its rounding sequence, so at these
sizes the asserted bound is relative to that measured noise floor:
    cuda_vs_exact  <= 1.5 * oracle_vs_exact + 2e-13      (the kernels are at least as accurate as the reference)
    cuda_vs_oracle <= 2.5 * oracle_vs_exact + 2e-13      (triangle inequality of the two)
while the unit-size meshes of tests/test_gpu_parity.py hold the plain 1e-12.

A second, independent effect concerns the fused kernel (variant 2): it reads 13 geometry doubles per straight-sided
triangle (B_alg's 104 B) where the reference stores rx, sx, ry, sy per NODE and nx, ny, Fscale per FACE NODE.  On an
affine element those per-node copies are mathematically equal, but the reference's setup computes each from nodal
coordinates (Dr x with |x| ~ 1 and differences ~ h_cell), so they scatter by eps |Dr| |x| / h_cell (`geometry_noise`,
measured from the arrays: 7e-12 on config 2, 6e-11 on config 3).  The kernel uses their mean.  The one place where that
scatter is amplified is a wall face -- a normal that is off by 1e-13 turns tangential momentum into a reflection jump that
Fscale*LIFT ~ 1e4/h_cell multiplies (5e-10 of max|RHS| on config 2, 4e-8 on config 3 with mean normals, measured with the
arbiter) -- so on wall face nodes the kernel reads the reference's own per-node normal.  What remains is measured without
any kernel: `geometry_effect` = arbiter(per-node geometry) vs arbiter(the geometry the kernel reads).  Variant 2 is held
    to the arbiter on the SAME geometry at the fp64 floor:            v2_cuda_vs_exact_meangeom <= 1.5 * floor + 2e-13
    to the reference-ordered oracle within floor + that effect:       v2_cuda_vs_oracle <= 2.5 * floor + 1.1 * geometry_effect + 2e-13
"""
import json
import os

import numpy as np
import pytest

from blitzdg_b200.initial_conditions import gaussian_hump
from blitzdg_b200.synthetic import provision, structured_mesh

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def relerr(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))


def oracle_rhs_strips(mesh, N, state, nstrips, strips=None, threads=None):
    """Oracle RHS of the global state evaluated strip by strip (oracle/threaded.py).  Returns (R [3][Np][K] with NaN
    outside the evaluated strips, element mask).  strips: which strip indices to evaluate (default all)."""
    from oracle.threaded import StripOracle

    threads = threads or max(1, min((os.cpu_count() or 2), 8))  # ~55 temporaries per strip and thread: bounded memory
    so = StripOracle(mesh, N, nparts=nstrips, threads=threads, parts=strips)
    so.set_state(*state)
    R = so.rhs()
    so.close()
    return R, ~np.isnan(R[0][0])


def geometry_noise(ctx):
    """Largest relative scatter of the per-node copies of the (mathematically constant) affine geometry factors: rx, sx,
    ry, sy within an element, nx, ny, Fscale within a face -- the rounding noise of the reference's own setup."""
    worst = 0.0
    vol = [ctx.rx, ctx.sx, ctx.ry, ctx.sy]
    scale = max(float(np.abs(a).max()) for a in vol)
    for a in vol:
        worst = max(worst, float((a.max(axis=0) - a.min(axis=0)).max()) / scale)
    Nfp = ctx.Nfp
    for a in (ctx.nx, ctx.ny, ctx.Fscale):
        b = a.reshape(ctx.Nfaces, Nfp, -1)
        worst = max(worst, float((b.max(axis=1) - b.min(axis=1)).max()) / float(np.abs(a).max()))
    return worst


from conftest import mean_geometry  # noqa: E402


def record(name, row):
    path = os.path.join(ROOT, "gpurun_out", "parity_fullsize.json")
    try:
        os.makedirs(os.path.dirname(path), exist_ok=True)
        data = json.load(open(path)) if os.path.exists(path) else {}
        data[name] = row
        json.dump(data, open(path, "w"), indent=1, sort_keys=True)
    except OSError:
        pass


CASES = [  # name, shape, N, nx, ny, oracle strips (total, evaluated), fused variant expected
    ("config 2: 1M tri N=4", "tri", 4, 1000, 500, (8, None), True),
    ("config 3: 4M tri N=8", "tri", 8, 2000, 1000, (40, None), True),
    ("config 4: 2M quad N=6", "quad", 6, 2000, 1000, (40, None), False),
    ("config 5: 16M tri N=4", "tri", 4, 4000, 2000, (16, (0, 7, 8)), True),  # oracle on 3 M elements: boundary strip + the two around the hump
]


@pytest.mark.parametrize("name,shape,N,nx,ny,strips,fused", CASES, ids=[c[0].split(":")[0].replace(" ", "") for c in CASES])
def test_full_mesh_rhs_against_oracle_and_arbiter(name, shape, N, nx, ny, strips, fused):
    from blitzdg_b200.sw2d import SW2dSolver
    from oracle.binding import Arbiter

    if name.startswith("config 5") and os.environ.get("BDG_SKIP_16M"):
        pytest.skip("BDG_SKIP_16M set")
    mesh = structured_mesh(shape, nx, ny)
    ctx = provision(mesh, N).get_DGContext()
    H, h, hu, hv = gaussian_hump(ctx.x, ctx.y)
    # a moving state that respects the wall BC (no normal flow at x, y = +-1), so that max|RHS| is set by the interior and
    # not by an O(1/h_cell) reflection jump at the walls: every flux component carries weight
    hu = 0.5 * h * np.sin(np.pi * ctx.x) * np.cos(2.0 * ctx.y)
    hv = 0.4 * h * np.cos(2.5 * ctx.x) * np.sin(np.pi * ctx.y)
    hu, hv = np.ascontiguousarray(hu), np.ascontiguousarray(hv)
    sol = SW2dSolver(ctx)
    sol.set_state(h, hu, hv)
    variants = [1]
    if fused:  # affine triangles N <= 8: the fused kernel is the product, its absence is a failure
        sol.set_variant(2)
        variants.append(2)
    Rc = {}
    for var in variants:
        sol.set_variant(var)
        Rc[var] = sol.computeRHS()
    sol.close()
    Ra = Arbiter(ctx).rhs(h, hu, hv)
    Ro, mask = oracle_rhs_strips(mesh, N, (h, hu, hv), strips[0], strips[1])
    assert mask.sum() >= min(ctx.K, 2_000_000), mask.sum()
    row = {"K": int(ctx.K), "N": N, "shape": shape, "oracle_elements": int(mask.sum()), "fields": ["h", "hu", "hv"]}
    row["oracle_vs_exact"] = [relerr(Ro[c][:, mask], Ra[c][:, mask]) for c in range(3)]
    floor = max(row["oracle_vs_exact"])
    for var in variants:
        row["v%d_cuda_vs_exact" % var] = [relerr(Rc[var][c], Ra[c]) for c in range(3)]
        row["v%d_cuda_vs_oracle" % var] = [relerr(Rc[var][c][:, mask], Ro[c][:, mask]) for c in range(3)]
    if 2 in variants:
        row["geometry_noise"] = geometry_noise(ctx)
        Rm = Arbiter(mean_geometry(ctx)).rhs(h, hu, hv)
        row["geometry_effect"] = [relerr(Rm[c], Ra[c]) for c in range(3)]
        row["v2_cuda_vs_exact_meangeom"] = [relerr(Rc[2][c], Rm[c]) for c in range(3)]
    record(name, row)
    print(name, json.dumps(row))
    assert max(row["v1_cuda_vs_exact"]) <= 1.5 * floor + 2e-13, row
    assert max(row["v1_cuda_vs_oracle"]) <= 2.5 * floor + 2e-13, row
    if 2 in variants:
        assert max(row["v2_cuda_vs_exact_meangeom"]) <= 1.5 * floor + 2e-13, row
        assert max(row["v2_cuda_vs_oracle"]) <= 2.5 * floor + 1.1 * max(row["geometry_effect"]) + 2e-13, row
        assert row["geometry_noise"] < 1e-10
    # the absolute level, so that a regression of either side shows: conditioning ~ eps * |Dr|_inf * (2/h_cell) * g h^2/2 / |RHS|
    assert floor < 1e-9


@pytest.mark.parametrize("name,shape,N,nx,ny", [(c[0], c[1], c[2], c[3], c[4]) for c in CASES],
                         ids=[c[0].split(":")[0].replace(" ", "") for c in CASES])
def test_full_size_properties(name, shape, N, nx, ny):
    """Size-independent properties at full size: still water is steady, mass is conserved over LSERK4 steps with all-wall
    BC, the state stays finite (config 5: all 16 M triangles on ONE GPU, byte offsets beyond 2^31)."""
    from blitzdg_b200.sw2d import SW2dSolver

    if name.startswith("config 5") and os.environ.get("BDG_SKIP_16M"):
        pytest.skip("BDG_SKIP_16M set")
    ctx = provision(structured_mesh(shape, nx, ny), N).get_DGContext()
    sol = SW2dSolver(ctx)
    flat = np.full((ctx.Np, ctx.K), 10.0)
    zero = np.zeros_like(flat)
    sol.set_state(flat, zero, zero)
    for R in sol.computeRHS():
        assert np.abs(R).max() < 1e-6 * N
    del flat, zero
    H, h, hu, hv = gaussian_hump(ctx.x, ctx.y)
    w = np.linalg.inv(ctx.V @ ctx.V.T).sum(axis=0)
    mass0 = (w[:, None] * ctx.J * h).sum()
    sol.set_bathymetry(H)  # eta = h - H in the blow-up guard (reference src/sw2d/main.cpp:193,238)
    sol.set_state(h, hu, hv)
    dt = 0.5 * sol.computeTimeStep() * (0.65 / 0.25)
    assert dt > 0
    sol.step(5, dt)
    hg, hug, hvg, _ = sol.get_state()
    mass1 = (w[:, None] * ctx.J * hg).sum()
    assert abs(mass1 - mass0) < 1e-11 * abs(mass0)
    assert np.isfinite(hg).all() and np.isfinite(hug).all() and sol.check() < 1.01
    sol.close()
