"""`import blitzdg_b200.pyblitzdg as dg`: the reference's OWN Python drivers run against it.

The reference's 2-D shallow-water drivers are scripts written against its boost.python module (`import pyblitzdg as dg`,
src/pyblitzdg/pyblitzdg.cpp:97-193).  Here they are executed from where they lie under /root/reference with that module
name bound to blitzdg_b200.pyblitzdg -- every set-up line, the driver's own numpy RHS and its time loop included:

  * sw2dquads.py (quads, N = 4, input/coarse_box_quads_fine.msh, filtered midpoint loop): 20 steps, the end state against the
    same loop around the C oracle, the .vtu files it wrote;
  * sw2d.py (triangles, N = 6, input/R_8km_circle.msh, Coriolis + tracer): 200 steps (the fewest its output arithmetic
    allows), conservation of mass and tracer, the .vtu files;
  * sw2d_curved.py (curved-element workflow: correctBCTable + setBCType, boundary splines, meshhelpers/curved.py, setCoordinates,
    buildGaussFaceNodes, makeMapsPeriodic, buildCubatureVolumeMesh, filtered midpoint loop around sw2dComputeRHS_curved) on a
    channel mesh written here in gmsh 2.2 format (the reference does not ship input/headlands_highres.msh): three steps,
    periodic maps, conservation, outputs, and its state through the C-ABI marshalling;
  * sw2dComputeRHS_curved (swhelpers/rhs.py:6-176) with the contexts of nodes.dgContext() / buildCubatureVolumeMesh /
    buildGaussFaceNodes handed over as they are -- the stored outputs of tests/golden/curved_reference_rhs.npz were generated
    through an adapter object; here there is none.

What is changed in a script: the literal assigned to `finalTime` (40 s / 3 days / 1 day of model time), and `np` is numpy with the two
NumPy-1.x aliases the scripts use (`np.float`, `np.dtype('Float64')`); matplotlib (absent, imported and never used on these
paths) is an empty module.  Skipped where /root/reference is absent (the GPU box).  The last test needs no reference: a
pyblitzdg-style context goes into the C-ABI structure unchanged.
"""
import ast
import os
import sys
import types

import numpy as np
import pytest

REF = "/root/reference"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
needs_ref = pytest.mark.skipif(not os.path.exists(os.path.join(REF, "sw2dquads.py")), reason="reference tree not present")


class _NumpyCompat(types.ModuleType):
    def __getattr__(self, name):
        return np.float64 if name == "float" else getattr(np, name)

    @staticmethod
    def dtype(x, *a, **k):
        return np.dtype("float64" if isinstance(x, str) and x == "Float64" else x, *a, **k)


class _SetFinalTime(ast.NodeTransformer):
    def __init__(self, expr):
        self.expr, self.hits = expr, 0

    def visit_Assign(self, node):
        if len(node.targets) == 1 and isinstance(node.targets[0], ast.Name) and node.targets[0].id == "finalTime":
            self.hits += 1
            node.value = ast.parse(self.expr, mode="eval").body
        return node


def run_reference_script(name, final_time_expr, cwd, extra=None):
    """Executes /root/reference/<name> as __main__ in `cwd` (input/ linked in), `pyblitzdg` bound to the facade."""
    import blitzdg_b200.pyblitzdg as facade

    path = os.path.join(REF, name)
    tree = ast.parse(open(path).read(), filename=path)
    tr = _SetFinalTime(final_time_expr)
    tree = ast.fix_missing_locations(tr.visit(tree))
    assert tr.hits == 1
    compat = _NumpyCompat("numpy")
    saved = {k: sys.modules.get(k) for k in ("pyblitzdg", "matplotlib", "matplotlib.pyplot")}
    mpl = types.ModuleType("matplotlib")
    mpl.pyplot = types.ModuleType("matplotlib.pyplot")
    sys.modules.update({"pyblitzdg": facade, "matplotlib": mpl, "matplotlib.pyplot": mpl.pyplot})
    old_cwd, old_path = os.getcwd(), list(sys.path)
    if not os.path.exists(os.path.join(cwd, "input")):
        os.symlink(os.path.join(REF, "input"), os.path.join(cwd, "input"))
    ns = {"__name__": "__main__", "__file__": path}
    ns.update(extra or {})
    had_float = hasattr(np, "float")
    try:
        if not had_float:
            np.float = np.float64  # the modules a script imports (swhelpers/rhs.py:101) see the real numpy: NumPy 1.x alias
        os.chdir(cwd)
        sys.path.insert(0, REF)
        for node in tree.body:  # statement by statement: `np` is re-bound to the compat module right after the script imports numpy
            exec(compile(ast.Module(body=[node], type_ignores=[]), path, "exec"), ns)
            if isinstance(node, ast.Import) and any(a.name == "numpy" for a in node.names):
                ns["np"] = compat
    finally:
        os.chdir(old_cwd)
        sys.path[:] = old_path
        if not had_float and hasattr(np, "float"):
            del np.float
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    return ns


@needs_ref
def test_sw2dquads_driver_runs_unmodified_and_agrees_with_the_oracle_loop(tmp_path):
    from blitzdg_b200.dgcontext import as_context
    from blitzdg_b200.outputters import read_vtu
    from oracle.binding import Oracle

    dt = 0.45 * 0.000724295  # sw2dquads.py:174
    ns = run_reference_script("sw2dquads.py", "20.5*0.45*0.000724295", str(tmp_path))
    assert ns["step"] == 21 and abs(ns["t"] - 21 * dt) < 1e-15  # `while t < finalTime` takes the step that crosses it
    ctx = as_context(ns["ctx"])
    assert (ctx.N, ctx.Np, ctx.K, ctx.Nfaces) == (4, 25, 64, 4)
    # the same loop (sw2dquads.py:179-203) around the C oracle: per-face LF, wall BC, filtered RHS, midpoint predictor / corrector
    orc = Oracle(ctx, lf_mode=0, sources=0)
    x, y = ctx.x, ctx.y
    H = 10 * np.ones_like(x)
    h, hu, hv = H + np.exp(-10 * x * x - 10 * y * y), np.zeros_like(x), np.zeros_like(x)
    F = ctx.Filt
    for _ in range(21):
        R = [F @ r for r in orc.rhs(h, hu, hv)]
        R = [F @ r for r in orc.rhs(h + 0.5 * dt * R[0], hu + 0.5 * dt * R[1], hv + 0.5 * dt * R[2])]
        h, hu, hv = h + dt * R[0], hu + dt * R[1], hv + dt * R[2]
    for mine, theirs in ((h, ns["h"]), (hu, ns["hu"]), (hv, ns["hv"])):
        assert np.abs(mine - theirs).max() <= 1e-11 * max(1.0, np.abs(theirs).max())
    assert np.abs(ns["hu"]).max() > 1e-4  # the hump has started to move
    # outputs: step 0 and step 20 (sw2dquads.py:166, 213-217), three fields each, 16 sub-quads per element
    for stepno in (0, 20):
        for f in ("eta", "u", "v"):
            got = read_vtu(str(tmp_path / ("%s%07d.vtu" % (f, stepno))))
            assert got["ncells"] == 64 * 16 and got["scalars"] == f
    assert not (tmp_path / "eta0000021.vtu").exists()


@needs_ref
def test_sw2d_driver_with_coriolis_and_tracer_runs_unmodified(tmp_path):
    import blitzdg_b200.pyblitzdg as dg
    from blitzdg_b200.outputters import read_vtu

    # the driver's own dt (sw2d.py:212-214), computed here first so that finalTime can be given as 200 steps of it
    m = dg.MeshManager()
    m.readMesh(os.path.join(REF, "input", "R_8km_circle.msh"))
    nodes = dg.TriangleNodesProvisioner(6, m)
    ctx = nodes.dgContext()
    g = (1.0025 - 1.000) * 9.81
    h0 = 10.0 - 2.5 * (ctx.x / 8000.0)
    dt = 0.8 / np.max(((6 + 1) ** 2) * 0.5 * np.abs(ctx.Fscale.flatten("F")) * (np.sqrt(g * h0).flatten("F")[ctx.vmapM]))
    ns = run_reference_script("sw2d.py", repr(199.5 * float(dt)), str(tmp_path))
    assert ns["numSteps"] == 200 and ns["outputInterval"] == 1 and ns["step"] == 200 and abs(ns["dt"] - dt) < 1e-18
    c = ns["ctx"]
    assert (c.numLocalPoints, c.numElements, c.numFaces) == (28, 858, 3) and sorted(c.BCmap) == [3]
    # all-wall basin: the integrals of h and of hN are conserved by the scheme.  Quadrature weights of the nodal basis: M = (V V^T)^-1
    w = np.linalg.inv(c.V @ c.V.T).sum(axis=0)

    def integral(q):
        return float((w[:, None] * c.J * q).sum())

    H = ns["H"]
    eta0 = -2.5 * (c.x / 8000.0)
    N0 = np.exp(-((c.y - 2500.0) / 2e3) ** 2)
    assert abs(integral(ns["h"]) - integral(H + eta0)) <= 1e-12 * integral(H)
    assert abs(integral(ns["hN"]) - integral((H + eta0) * N0)) <= 1e-12 * integral(H)
    assert np.isfinite(ns["h"]).all() and np.abs(ns["hu"]).max() > 1e-3 and np.abs(ns["h"] - (H + eta0)).max() > 1e-6
    # the driver's own RHS (sw2d.py:30-146: per-face LF, Coriolis, tracer) on its end state against the C oracle the CUDA path is held to
    from blitzdg_b200.dgcontext import as_context
    from oracle.binding import Oracle

    orc = Oracle(as_context(c), lf_mode=0, sources=1, g=ns["g"], CD=0.0, f=ns["f"])
    orc.set_bathymetry(H, np.zeros_like(H), np.zeros_like(H))
    Rd = ns["sw2dComputeRHS"](ns["h"], ns["hu"], ns["hv"], ns["hN"], ns["g"], H, ns["f"], c)
    Ro = orc.rhs(ns["h"], ns["hu"], ns["hv"]) + [orc.rhs_tracer(ns["h"], ns["hu"], ns["hv"], ns["hN"])]
    for i in range(4):
        assert np.abs(Ro[i] - Rd[i]).max() <= 1e-11 * np.abs(Rd[i]).max(), i
    for stepno in (0, 1, 200):
        for f in ("eta", "u", "v", "N"):
            got = read_vtu(str(tmp_path / ("%s%07d.vtu" % (f, stepno))))
            assert got["ncells"] == 858 * 36 and got["scalars"] == f


@needs_ref
def test_reference_curved_rhs_takes_the_facade_contexts_as_they_are():
    """sw2dComputeRHS_curved(h, hu, hv, hN, zx, zy, g, H, f, CD, ctx, cub_ctx, gauss_ctx, curvedEls, J, gmapM, gmapP) with the three
    contexts straight from the facade's provisioner (sw2d_curved.py:51-53, 118, 126, 148) reproduces the stored outputs."""
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import make_curved_reference_vectors as gen

    import blitzdg_b200.pyblitzdg as dg
    from blitzdg_b200.synthetic import structured_arrays

    saved_float = getattr(np, "float", None)
    sys.path.insert(0, REF)
    try:
        if saved_float is None:
            np.float = np.float64  # swhelpers/rhs.py:101 (NumPy 1.x alias)
        from swhelpers.rhs import sw2dComputeRHS_curved

        fx = np.load(os.path.join(ROOT, "tests", "golden", "curved_reference_rhs.npz"))
        name = "curved_N2"
        nx, ny, N, ncub, ng = gen.CASES[name]
        etov, vert = structured_arrays("tri", nx, ny)
        mesh = dg.MeshManager()
        mesh.buildMesh(etov, vert)
        nodes = dg.TriangleNodesProvisioner(N, mesh)
        nodes.buildFilter(0.9 * N, 4)
        ctx = nodes.dgContext()
        xd, yd = gen.deformation(ctx.x, ctx.y)
        nodes.setCoordinates(xd, yd)
        gauss_ctx = nodes.buildGaussFaceNodes(ng)
        cub_ctx = nodes.buildCubatureVolumeMesh(ncub)
        assert np.array_equal(ctx.x, xd)  # the context reads through to its provisioner, as the reference's raw pointers do
        d = {k: fx["%s_%s" % (name, k)] for k in ("H", "h", "hu", "hv", "hN", "zx", "zy", "CD", "curved")}
        g, f = (float(v) for v in fx[name + "_params"][5:7])
        R = sw2dComputeRHS_curved(d["h"].copy(), d["hu"].copy(), d["hv"].copy(), d["hN"].copy(), d["zx"], d["zy"], g, d["H"], f, d["CD"], ctx, cub_ctx,
                                  gauss_ctx, d["curved"].tolist(), ctx.J, gauss_ctx.mapM, gauss_ctx.mapP)
        for i in range(4):
            want = fx["%s_RHS%d" % (name, i + 1)]
            assert np.abs(R[i] - want).max() <= 1e-13 * np.abs(want).max(), i
    finally:
        sys.path.remove(REF)
        if saved_float is None and hasattr(np, "float"):
            del np.float


def test_pyblitzdg_style_context_packs_like_the_mirror_context(coarse_box_mesh):
    """SW2dSolver / pack_context take the reference binding's property names (numLocalPoints, filter, ...): same C structure."""
    import blitzdg_b200.pyblitzdg as dg
    from blitzdg_b200 import TriangleNodesProvisioner, capi
    from blitzdg_b200.dgcontext import as_context

    m = dg.MeshManager()
    m.readMesh(os.path.join(ROOT, "tests", "golden", "coarse_box.msh"))
    nodes = dg.TriangleNodesProvisioner(4, m)
    ctx = nodes.dgContext()
    assert ctx.filter is None  # no buildFilter yet
    nodes.buildFilter(3.6, 4)
    assert ctx.filter.shape == (15, 15)  # seen through the context made earlier
    ref = TriangleNodesProvisioner(4, coarse_box_mesh)
    ref.buildFilter(3.6, 4)
    a, keep_a = capi.pack_context(ctx)
    b, keep_b = capi.pack_context(ref.get_DGContext())
    assert (a.N, a.Np, a.Nfp, a.Nfaces, a.K) == (b.N, b.Np, b.Nfp, b.Nfaces, b.K) == (4, 15, 5, 3, 40)
    for name, n in (("Dr", 225), ("Ds", 225), ("Lift", 225), ("Filt", 225), ("rx", 600), ("sy", 600), ("nx", 600), ("Fscale", 600)):
        assert np.array_equal(np.ctypeslib.as_array(getattr(a, name), (n,)), np.ctypeslib.as_array(getattr(b, name), (n,))), name
    assert np.array_equal(np.ctypeslib.as_array(a.vmapP, (600,)), np.ctypeslib.as_array(b.vmapP, (600,)))
    assert [a.bc_count[t] for t in range(9)] == [b.bc_count[t] for t in range(9)]
    # a bare object with pyblitzdg's property names (what the real binding returns) converts too
    bare = type("Ctx", (), {k: getattr(ctx, k) for k in ("numLocalPoints", "numFacePoints", "numElements", "numFaces", "filter", "Dr", "Ds", "Lift", "rx", "ry",
                                                         "sx", "sy", "nx", "ny", "Fscale", "vmapM", "vmapP", "BCmap", "J", "x", "y", "r", "s", "Fmask", "V")})()
    c = as_context(bare)
    assert (c.N, c.Np, c.K) == (4, 15, 40) and np.array_equal(c.vmapP, ref.get_DGContext().vmapP) and c.Filt.shape == (15, 15)
    assert m.bcType.shape == (40, 3) and m.elements.shape == (40, 3) and m.vertices.shape == (29, 3) and m.numElements == 40
    bt = m.bcType
    bt[bt == 3] = 2
    m.setBCType(bt)
    assert sorted(dg.TriangleNodesProvisioner(2, m).dgContext().BCmap) == [2]


def _write_channel_msh(path, nx=16, ny=4, Lx=8000.0, Ly=1000.0):
    """A gmsh 2.2 ASCII file of a structured triangle mesh of [0, Lx] x [0, Ly]: stands in for input/headlands_highres.msh,
    which the reference does not ship.  The driver's geometry tests (a wall curve with y > 200 between x = 3250 and 4750,
    periodic Out ends more than 1000 apart, a node within 19 of the origin) are met by this channel."""
    xs, ys = np.linspace(0.0, Lx, nx + 1), np.linspace(0.0, Ly, ny + 1)
    vid = lambda i, j: j * (nx + 1) + i + 1  # noqa: E731  (1-based)
    with open(path, "w") as fh:
        fh.write("$MeshFormat\n2.2 0 8\n$EndMeshFormat\n$Nodes\n%d\n" % ((nx + 1) * (ny + 1)))
        for j in range(ny + 1):
            for i in range(nx + 1):
                fh.write("%d %.17g %.17g 0\n" % (vid(i, j), xs[i], ys[j]))
        fh.write("$EndNodes\n$Elements\n%d\n" % (2 * nx * ny))
        e = 1
        for j in range(ny):
            for i in range(nx):
                a, b, c, d = vid(i, j), vid(i + 1, j), vid(i + 1, j + 1), vid(i, j + 1)
                fh.write("%d 2 2 0 1 %d %d %d\n%d 2 2 0 1 %d %d %d\n" % (e, a, b, c, e + 1, a, c, d))
                e += 2
        fh.write("$EndElements\n")


@needs_ref
def test_sw2d_curved_driver_runs_unmodified_on_a_channel_mesh(tmp_path):
    """sw2d_curved.py from its first line to its time loop: correctBCTable + setBCType, the boundary curve and its splines,
    meshhelpers/curved.py (VandermondeBuilder, Fmask, r, s), setCoordinates, buildGaussFaceNodes, makeMapsPeriodic on the nodal
    AND the Gauss maps, buildCubatureVolumeMesh, the drag field, three steps of the filtered midpoint loop around
    sw2dComputeRHS_curved, the .vtu files of step 0.  Then the driver's own state goes through the C-ABI marshalling
    (pack_curved) with its periodic Gauss maps."""
    from blitzdg_b200 import capi
    from blitzdg_b200.dgcontext import as_context
    from blitzdg_b200.outputters import read_vtu

    os.makedirs(tmp_path / "input")
    _write_channel_msh(str(tmp_path / "input" / "headlands_highres.msh"))
    ns = run_reference_script("sw2d_curved.py", "40.0", str(tmp_path))  # dt = 16.8 on this mesh: three steps
    assert ns["step"] == 3 and 40.0 < ns["t"] < 60.0 and 10.0 < ns["dt"] < 20.0
    ctx, cub, gauss = ns["ctx"], ns["cub_ctx"], ns["gauss_ctx"]
    assert (ctx.numLocalPoints, ctx.numElements) == (15, 128) and sorted(ctx.BCmap) == [2, 3]
    assert cub.NumCubaturePoints == 64 and gauss.NGauss == 10 and len(ns["curvedEls"]) > 0
    # periodic ends: every Out node found its partner on the other end, nodal and Gauss maps alike
    for mM, mP, ids, xf in ((ns["vmapM"], ns["vmapP"], ctx.BCmap[2], ns["xFlat"]), (ns["gmapM"], ns["gmapP"], gauss.BCmap[2], ns["gxFlat"])):
        ids = np.asarray(ids)
        assert (np.abs(xf[mM[ids]] - xf[mP[ids]]) > 7999.0).all()
    # the scheme conserves mass and tracer in the periodic, walled channel; the state moved
    w = cub.W
    mass0 = float((w * (cub.V @ (ns["H"] + ns["amp"] * np.exp(-((ns["y"] - ns["L"]) / ns["W"]) ** 2)))).sum())
    assert abs(float((w * (cub.V @ ns["h"])).sum()) - mass0) <= 1e-11 * mass0
    assert np.isfinite(ns["hu"]).all() and np.abs(ns["hv"]).max() > 0
    for f in ("eta", "u", "v", "B", "h", "vort"):
        assert read_vtu(str(tmp_path / ("%s0000000.vtu" % f)))["ncells"] == 128 * 16
    # the oracle the CUDA kernel is held to, on the driver's own end state with its periodic maps, Out faces and nodal drag
    from oracle.curved_numpy import rhs_curved

    args = (ns["h"], ns["hu"], ns["hv"], ns["hN"], ns["zx"], ns["zy"], ns["g"])
    had_float = hasattr(np, "float")
    try:
        if not had_float:
            np.float = np.float64  # swhelpers/rhs.py:99-101 (NumPy 1.x alias)
        Rd = ns["sw2dComputeRHS_curved"](*[a.copy() if hasattr(a, "copy") else a for a in args], ns["H"], ns["f"], ns["CD"], ctx, cub, gauss, ns["curvedEls"],
                                         ns["J"], ns["gmapM"], ns["gmapP"])
    finally:
        if not had_float and hasattr(np, "float"):
            del np.float
    Ro = rhs_curved(*args, ns["f"], ns["CD"], as_context(ctx), cub, gauss, ns["curvedEls"], ns["J"], ns["gmapM"], ns["gmapP"])
    for i in range(4):  # the jet is in geostrophic balance: RHS2 is what is left of terms 1e3 times larger, hence 1e-11 and not 1e-12
        assert np.abs(Ro[i] - Rd[i]).max() <= 1e-11 * np.abs(Rd[i]).max(), i
    # what the driver holds goes into the C ABI as it is
    c, keep = capi.pack_curved(as_context(ctx), cub, gauss, ns["curvedEls"], ns["J"], ns["gmapM"], ns["gmapP"])
    assert (c.Ncub, c.NGauss, c.n_curved) == (64, 10, len(ns["curvedEls"]))
