"""Device element order (blitzdg_b200/csrc/sw2d_order.cu): host-side properties on CPU, result invariance on the GPU."""
import os

import numpy as np
import pytest

from blitzdg_b200 import capi
from blitzdg_b200.synthetic import provision, structured_mesh


def _order(ctx, n_interior=None):
    lib = capi.load()
    K = ctx.K
    vmapP = np.ascontiguousarray(ctx.vmapP, dtype=np.int32)
    perm = np.empty(K, dtype=np.int32)
    rc = lib.bdg_element_order(capi._i(vmapP), ctx.Np, ctx.Nfp, ctx.Nfaces, K, K if n_interior is None else n_interior, capi._i(perm))
    assert rc == capi.BDG_OK
    return perm


def _ghosts_per_tile(ctx, perm):
    """distinct '+' nodes outside each 32-element tile, for the element order perm (new -> old)."""
    K, Np, Nfn = ctx.K, ctx.Np, ctx.Nfaces * ctx.Nfp
    inv = np.empty(K, dtype=np.int64)
    inv[perm] = np.arange(K)
    vp = np.asarray(ctx.vmapP, dtype=np.int64).reshape(K, Nfn)[perm]
    tile_of_nbr = inv[vp // Np] // 32
    out = []
    for T in range(K // 32):
        rows = slice(32 * T, 32 * T + 32)
        ext = vp[rows][tile_of_nbr[rows] != T]
        out.append(len(np.unique(ext)))
    return np.array(out)


@pytest.mark.parametrize("shape,nx,ny,N", [("tri", 64, 48, 4), ("quad", 40, 40, 3)])
def test_order_is_a_permutation_and_makes_compact_tiles(shape, nx, ny, N):
    ctx = provision(structured_mesh(shape, nx, ny), N).get_DGContext()
    perm = _order(ctx)
    assert np.array_equal(np.sort(perm), np.arange(ctx.K))
    g_new = _ghosts_per_tile(ctx, perm).mean()
    g_old = _ghosts_per_tile(ctx, np.arange(ctx.K, dtype=np.int32)).mean()
    assert g_new < 0.75 * g_old, (g_new, g_old)
    # inside a tile the caller's relative order is kept
    for T in range(ctx.K // 32):
        assert np.all(np.diff(perm[32 * T:32 * T + 32]) > 0)


def test_order_keeps_the_interior_cut_split_and_small_meshes():
    ctx = provision(structured_mesh("tri", 32, 20), 2).get_DGContext()
    n_int = 1000
    perm = _order(ctx, n_interior=n_int)
    assert np.array_equal(np.sort(perm[:n_int]), np.arange(n_int))
    assert np.array_equal(np.sort(perm[n_int:]), np.arange(n_int, ctx.K))
    small = provision(structured_mesh("tri", 4, 4), 2).get_DGContext()  # K = 32 <= tile: identity
    assert np.array_equal(_order(small), np.arange(small.K))


@pytest.mark.gpu
def test_results_do_not_depend_on_the_device_order():
    """Per-face LF: an element's arithmetic does not depend on where it sits (only the face-slot arrangement, i.e.
    the summation order of the LIFT columns, may differ between two creates) -> equal to round-off."""
    from blitzdg_b200.initial_conditions import gaussian_hump
    from blitzdg_b200.sw2d import SW2dSolver

    ctx = provision(structured_mesh("tri", 48, 40), 4).get_DGContext()
    H, h, hu, hv = gaussian_hump(ctx.x, ctx.y)
    rng = np.random.default_rng(3)
    hu = 0.1 * rng.standard_normal(h.shape)
    hv = 0.1 * rng.standard_normal(h.shape)
    outs = []
    for flag in ("1", "0"):
        os.environ["BDG_RENUMBER"] = flag
        try:
            sol = SW2dSolver(ctx, device=0)
            sol.set_state(h, hu, hv)
            dt = 0.25 * sol.computeTimeStep()
            R = sol.computeRHS()
            sol.step(7, dt)
            outs.append((R, sol.get_state()[:3]))
            sol.close()
        finally:
            os.environ.pop("BDG_RENUMBER", None)
    for a, b in zip(list(outs[0][0]) + list(outs[0][1]), list(outs[1][0]) + list(outs[1][1])):
        assert np.abs(a - b).max() <= 1e-13 * np.abs(b).max()
