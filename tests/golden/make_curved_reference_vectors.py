"""Generate curved_reference_rhs.npz: outputs of the REFERENCE'S OWN sw2dComputeRHS_curved (swhelpers/rhs.py:6-176).

Run once in the build container (reads /root/reference, which does not exist on the GPU box):
    python tests/golden/make_curved_reference_vectors.py

swhelpers/rhs.py and swhelpers/flux.py are imported as they stand.  Only compatibility shim: `np.float`, a NumPy 1.x
alias that NumPy 2.x removed (rhs.py:101), is mapped to float64.  No arithmetic is touched.

The function reads three contexts.  pyblitzdg's are replaced by this repo's host mirrors: DGContext2D (checked against the
reference's C++ specs in tests/test_setup_golden.py) and the cubature / Gauss-face contexts of blitzdg_b200/curved.py
(the cubature points come from a product rule, not from the reference's tables -- the function only sees the arrays).
Inputs are stored beside the outputs, and so is everything needed to rebuild the contexts: mesh size, order, the
deformation that curves the elements, the cubature and Gauss orders, the list of elements sent through the Cholesky path.
"""
import os
import sys
import types

import numpy as np

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, REF)

NSTEPS = 4
CASES = {  # name -> (nx, ny, N, cubature order, Gauss points per face)
    "curved_N4": (6, 5, 4, 15, 10),
    "curved_N2": (5, 4, 2, 9, 6),
    "curved_N4_reftable": (6, 5, 4, 15, 10),
}
# cases whose cubature points are the REFERENCE'S OWN tabulated rule of that order (include/TriangleCubatureRules.hpp; 54 points
# at order 15), read at generation time through the compiled header (oracle/make_ref_curved.py -> ref_cubature_rule) and stored
# beside the vectors as <case>_rule = [r; s; w], so the tests rebuild the same contexts without the reference tree
TABLE_RULE_CASES = ("curved_N4_reftable",)


def deformation(x, y):
    """Smooth map of [-1,1]^2 onto itself that keeps the boundary: every element becomes a genuinely curved triangle."""
    return x + 0.04 * np.sin(np.pi * y) * (1 - x * x), y + 0.03 * np.sin(np.pi * x) * (1 - y * y)


def reference_table_rule(order):
    """(r, s, w) of the reference's tabulated rule, from oracle/_ref/libcurved_ref.so (generation time only)."""
    import ctypes as C
    import subprocess

    lib = os.path.join(ROOT, "oracle", "_ref", "libcurved_ref.so")
    if not os.path.exists(lib):
        subprocess.check_call([sys.executable, os.path.join(ROOT, "oracle", "make_ref_curved.py")])
    ref = C.CDLL(lib)
    buf = [np.zeros(256) for _ in range(3)]
    n = ref.ref_cubature_rule(int(order), *[b.ctypes.data_as(C.POINTER(C.c_double)) for b in buf])
    return np.stack([b[:n] for b in buf])


def build_case(nx, ny, N, ncub, ngauss, rule=None):
    from blitzdg_b200.curved import build_cubature_volume_mesh, build_gauss_face_nodes, set_coordinates
    from blitzdg_b200.synthetic import provision, structured_mesh

    nodes = provision(structured_mesh("tri", nx, ny), N)
    nodes.buildFilter(0.9 * N, 4)  # sw2d_curved.py:49-53
    xd, yd = deformation(nodes.xGrid, nodes.yGrid)
    set_coordinates(nodes, xd, yd)
    cub = build_cubature_volume_mesh(nodes, ncub, rule=rule)  # also repairs the nodal metrics, as the reference does
    gauss = build_gauss_face_nodes(nodes, ngauss)
    ctx = nodes.get_DGContext()
    K = ctx.K
    curved = np.arange(0, K, 3, dtype=np.int32)  # every third element through the Cholesky path, the rest through V V^T / J
    return nodes, ctx, cub, gauss, curved


def state(ctx):
    x, y = ctx.x, ctx.y
    H = 7.5 + 0.5 * x - 0.25 * y * y
    h = H + 0.8 * np.exp(-6 * x * x - 6 * y * y) + 0.05 * np.sin(7 * x - 3 * y)
    hu = 0.5 * np.sin(3 * x + y) + 0.1 * x * y
    hv = 0.3 * np.cos(2 * y - x) - 0.2 * x
    hN = h * np.exp(-4 * ((x - 0.2) ** 2 + (y + 0.1) ** 2))
    z = -H
    zx = ctx.rx * (ctx.Dr @ z) + ctx.sx * (ctx.Ds @ z)
    zy = ctx.ry * (ctx.Dr @ z) + ctx.sy * (ctx.Ds @ z)
    CD = 2.5e-3 * (0.5 + 0.5 * np.tanh(3 * x))  # nodal drag coefficient, as sw2d_curved.py:166-177 builds one
    return H, h, hu, hv, hN, zx, zy, CD


def py_view(ctx):
    return types.SimpleNamespace(numElements=ctx.K, numLocalPoints=ctx.Np, V=ctx.V)


def main():
    if not hasattr(np, "float"):
        np.float = np.float64  # swhelpers/rhs.py:101 (NumPy 1.x alias)
    from swhelpers.rhs import sw2dComputeRHS_curved

    out = {}
    g, f = 9.81 * 1e-3, 7.8825e-5  # reduced gravity and Coriolis parameter of sw2d_curved.py:20-26,158
    for name, (nx, ny, N, ncub, ngauss) in CASES.items():
        rule = reference_table_rule(ncub) if name in TABLE_RULE_CASES else None
        if rule is not None:
            out[name + "_rule"] = rule
        nodes, ctx, cub, gauss, curved = build_case(nx, ny, N, ncub, ngauss, rule=rule)
        H, h, hu, hv, hN, zx, zy, CD = state(ctx)
        R = sw2dComputeRHS_curved(h.copy(), hu.copy(), hv.copy(), hN.copy(), zx, zy, g, H, f, CD, py_view(ctx), cub, gauss,
                                  curved.tolist(), ctx.J, gauss.mapM, gauss.mapP)
        for i in range(4):
            out["%s_RHS%d" % (name, i + 1)] = np.ascontiguousarray(R[i])
        for nm, a in (("H", H), ("h", h), ("hu", hu), ("hv", hv), ("hN", hN), ("zx", zx), ("zy", zy), ("CD", CD), ("curved", curved)):
            out["%s_%s" % (name, nm)] = a
        out["%s_params" % name] = np.array([nx, ny, N, ncub, ngauss, g, f])
        # the driver's time loop (sw2d_curved.py:240-262): filtered RHS, midpoint predictor, corrector
        c0 = np.sqrt(g * H.mean())
        u, v = hu / h, hv / h
        dt = 0.5 * 0.75 / np.max(((N + 1) ** 2) * 0.5 * np.abs(ctx.Fscale.flatten("F")) * (c0 + np.hypot(u, v).flatten("F")[ctx.vmapM]))
        q = [a.copy() for a in (h, hu, hv, hN)]
        Filt = ctx.Filt
        for _ in range(NSTEPS):
            R = sw2dComputeRHS_curved(q[0], q[1], q[2], q[3], zx, zy, g, H, f, CD, py_view(ctx), cub, gauss, curved.tolist(), ctx.J, gauss.mapM,
                                      gauss.mapP)
            q1 = [q[i] + 0.5 * dt * np.dot(Filt, R[i]) for i in range(4)]
            R = sw2dComputeRHS_curved(q1[0], q1[1], q1[2], q1[3], zx, zy, g, H, f, CD, py_view(ctx), cub, gauss, curved.tolist(), ctx.J,
                                      gauss.mapM, gauss.mapP)
            for i in range(4):
                q[i] += dt * np.dot(Filt, R[i])
        for i, nm in enumerate(("h", "hu", "hv", "hN")):
            out["%s_step_%s" % (name, nm)] = q[i]
        out["%s_step_params" % name] = np.array([NSTEPS, dt])
        print(name, "K", ctx.K, "Ncub", cub.NumCubaturePoints, "max|RHS|", [float(np.abs(r).max()) for r in R])
    np.savez_compressed(os.path.join(HERE, "curved_reference_rhs.npz"), **out)


if __name__ == "__main__":
    main()
