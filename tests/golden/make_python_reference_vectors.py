"""Generate python_reference_rhs.npz: outputs of the REFERENCE'S OWN Python code for the sw2d RHS.

Run once in the build container (reads /root/reference, which does not exist on the GPU box):
    python tests/golden/make_python_reference_vectors.py

What is executed, unmodified:
  * /root/reference/sw2dquads.py:10-123  sw2dComputeFluxes / sw2dComputeRHS  (per-face LF, wall BC, no sources).
    The file is a script that imports pyblitzdg (not built in this image) at the top, so the two function
    definitions are cut out of its source with `ast` and exec'd as they stand.
  * /root/reference/swhelpers/rhs.py:178-311  sw2dComputeRHS  (adds Coriolis, drag, bed slope and a 4th tracer
    equation), imported as a module together with swhelpers/flux.py.
  * /root/reference/swhelpers/limiters.py:3-21  positivityPreservingLimiter2D, imported.
Only compatibility shim: both files use NumPy 1.x aliases that NumPy 2.x removed (`np.float`,
`np.dtype('Float64')`); they are mapped to float64 below.  No arithmetic is touched.

The functions take a DGContext2D; pyblitzdg's is replaced by this repo's host mirror, whose arrays are checked
against the reference's C++ specs in tests/test_setup_golden.py.  Inputs (smooth, fixed formulas of the node
coordinates) are stored beside the outputs, so the tests do not depend on regenerating them bit for bit.
"""
import ast
import os
import sys
import types

import numpy as np

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, REF)


class _NumpyCompat(types.ModuleType):
    """numpy with the two removed 1.x aliases the reference scripts use."""

    def __getattr__(self, name):
        if name == "float":
            return np.float64
        return getattr(np, name)

    @staticmethod
    def dtype(x, *a, **k):
        return np.dtype("float64" if x == "Float64" else x, *a, **k)


def load_script_functions(path, names):
    src = open(path).read()
    tree = ast.parse(src)
    ns = {"np": _NumpyCompat("numpy_compat")}
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in names:
            exec(compile(ast.Module(body=[node], type_ignores=[]), path, "exec"), ns)
    return [ns[n] for n in names]


def py_view(ctx):
    """The property set pyblitzdg exposes for a DGContext2D (src/pyblitzdg/pyblitzdg.cpp:160-187) over the host mirror,
    whose accessors follow the C++ class (methods)."""
    v = types.SimpleNamespace(numFacePoints=ctx.Nfp, numFaces=ctx.Nfaces, numElements=ctx.K, numLocalPoints=ctx.Np,
                              order=ctx.N)
    for nm in ("vmapM", "vmapP", "BCmap", "nx", "ny", "rx", "sx", "ry", "sy", "Dr", "Ds", "Lift", "Fscale", "x", "y", "J"):
        setattr(v, nm, getattr(ctx, nm))
    return v


def state(ctx, H):
    x, y = ctx.x, ctx.y
    h = H + 0.8 * np.exp(-10 * x * x - 10 * y * y) + 0.05 * np.sin(7 * x - 3 * y)
    hu = 0.5 * np.sin(3 * x + y) + 0.1 * x * y
    hv = 0.3 * np.cos(2 * y - x) - 0.2 * x
    return h, hu, hv


def main():
    from blitzdg_b200 import MeshManager, QuadNodesProvisioner, TriangleNodesProvisioner

    if not hasattr(np, "float"):
        np.float = np.float64  # swhelpers/rhs.py:266 (module-level numpy alias)
    (quads_rhs,) = load_script_functions(os.path.join(REF, "sw2dquads.py"), ["sw2dComputeFluxes", "sw2dComputeRHS"])[1:]
    from swhelpers.limiters import positivityPreservingLimiter2D
    from swhelpers.rhs import sw2dComputeRHS as helpers_rhs

    out = {}
    g = 9.81
    for tag, msh, quad, N in (("tri4", "coarse_box.msh", False, 4), ("tri7", "coarse_box.msh", False, 7),
                              ("quad3", "coarse_box_quads.msh", True, 3), ("quad6", "coarse_box_quads.msh", True, 6)):
        mm = MeshManager()
        mm.readMesh(os.path.join(HERE, msh))
        ctx = (QuadNodesProvisioner if quad else TriangleNodesProvisioner)(N, mm).get_DGContext()
        H = np.full_like(ctx.x, 10.0)
        h, hu, hv = state(ctx, H)
        R = quads_rhs(h, hu, hv, g, H, py_view(ctx))
        for nm, a in zip(("h", "hu", "hv", "RHS1", "RHS2", "RHS3"), (h, hu, hv) + tuple(R)):
            out["sw2dquads_%s_%s" % (tag, nm)] = np.ascontiguousarray(a)

    # 1000 LSERK4 steps driven by the reference's RHS function: stage idiom of /root/reference/advec1d.py:80-88
    # (`resRK = rk4a[stg]*resRK + dt*RHS; u += rk4b[stg]*resRK`), coefficients of include/LSERK4.hpp:17-27 (host mirror,
    # pinned in tests/test_oracle.py), each of h, hu, hv updated like advec1d's u; fixed dt = 2e-3 (CFL dt here ~3.7e-3).
    from blitzdg_b200 import LSERK4

    for tag, msh, quad, N in (("tri4", "coarse_box.msh", False, 4), ("quad3", "coarse_box_quads.msh", True, 3)):
        mm = MeshManager()
        mm.readMesh(os.path.join(HERE, msh))
        ctx = (QuadNodesProvisioner if quad else TriangleNodesProvisioner)(N, mm).get_DGContext()
        view = py_view(ctx)
        H = np.full_like(ctx.x, 10.0)
        q = [a.copy() for a in state(ctx, H)]
        res = [np.zeros_like(a) for a in q]
        dt = 2e-3
        for _ in range(1000):
            for stg in range(5):
                R = quads_rhs(q[0], q[1], q[2], g, H, view)
                for c in range(3):
                    res[c] = LSERK4.rk4a[stg] * res[c] + dt * R[c]
                    q[c] += LSERK4.rk4b[stg] * res[c]
        out["lserk4_1000_%s_dt" % tag] = np.array(dt)
        for nm, a in zip(("h", "hu", "hv"), q):
            assert np.isfinite(a).all()
            out["lserk4_1000_%s_%s" % (tag, nm)] = np.ascontiguousarray(a)

    # swhelpers/rhs.py on triangles with a sloping bed; z = -H so zx = -Hx.
    mm = MeshManager()
    mm.readMesh(os.path.join(HERE, "coarse_box.msh"))
    ctx = TriangleNodesProvisioner(4, mm).get_DGContext()
    H = 10 + 0.5 * ctx.x + 0.25 * ctx.y
    Hx, Hy = np.full_like(H, 0.5), np.full_like(H, 0.25)
    h, hu, hv = state(ctx, H)
    f = 1.0070e-4
    hN = h * (0.5 + 0.1 * np.sin(2 * ctx.x + ctx.y))  # tracer of the 4th equation (RHS1-3 do not see it)
    out["swhelpers_hN"] = np.ascontiguousarray(hN)
    for tag, CD in (("cd0", 0.0), ("cd", 2.5e-3)):
        R = helpers_rhs(h, hu, hv, hN, -Hx, -Hy, g, H, f, CD, py_view(ctx), ctx.vmapM, ctx.vmapP)
        for nm, a in zip(("RHS1", "RHS2", "RHS3", "RHS4"), R):
            out["swhelpers_%s_%s" % (tag, nm)] = np.ascontiguousarray(a)
    for nm, a in zip(("h", "hu", "hv", "H"), (h, hu, hv, H)):
        out["swhelpers_" + nm] = np.ascontiguousarray(a)

    # limiter: a state with thin water in some elements
    hl = np.maximum(h - 10.2 - 0.4 * ctx.x, -0.05) + 0.02
    lim = positivityPreservingLimiter2D(hl.copy(), hu.copy(), hv.copy())
    out["limiter_h_in"] = hl
    for nm, a in zip(("h", "hu", "hv"), lim):
        out["limiter_%s_out" % nm] = np.ascontiguousarray(a)

    path = os.path.join(HERE, "python_reference_rhs.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()
