"""Output path: splitElements / VtkOutputter / CsvOutputter mirrors against a loop-level
restatement of the reference, and the device-fed output (bdg_sw2d_output_begin) against the same oracle."""
import os

import numpy as np
import pytest

from blitzdg_b200 import QuadNodesProvisioner, TriangleNodesProvisioner
from blitzdg_b200.initial_conditions import gaussian_hump
from blitzdg_b200.outputters import CsvOutputter, VtkOutputter, read_vtu, write_vtu
from blitzdg_b200.synthetic import provision, structured_mesh


def _prov(mesh, N, quad):
    return (QuadNodesProvisioner if quad else TriangleNodesProvisioner)(N, mesh)


@pytest.mark.parametrize("quad,N", [(False, 2), (False, 4), (True, 3)])
def test_split_elements_matches_reference_loops(coarse_box_mesh, coarse_quads_mesh, quad, N):
    from oracle.output_numpy import split_elements

    P = _prov(coarse_quads_mesh if quad else coarse_box_mesh, N, quad)
    x, y = P.get_xGrid(), P.get_yGrid()
    f = np.sin(2 * x) * np.cos(y) + 0.1 * x
    rout, sout, cells = P.splitTable()
    IM = P.computeInterpMatrix(rout, sout)
    assert len(rout) == P.NumLocalPoints and len(cells) == N * N
    want = split_elements(P.NumLocalPoints, N, quad, IM, x, y, f)
    got = P.splitElements(x, y, f)
    for a, b in zip(got, want):
        assert a.shape == b.shape == (4 if quad else 3, N * N * P.NumElements)
        assert np.abs(a - b).max() < 1e-13
    # the interpolation is exact for polynomials of the element's order: corner values = the polynomial at the corner
    g = 1.0 + 2.0 * x - 3.0 * y + (0.0 if quad else 1.0) * x * y
    xn, yn, gn = P.splitElements(x, y, g)
    assert np.abs(gn - (1.0 + 2.0 * xn - 3.0 * yn + (0.0 if quad else 1.0) * xn * yn)).max() < 1e-12
    # the sub-elements tile the element: summed areas = total mesh area (triangles: shoelace; quads: two triangles)
    def area(ix):
        return 0.5 * np.abs((xn[ix[1]] - xn[ix[0]]) * (yn[ix[2]] - yn[ix[0]]) - (xn[ix[2]] - xn[ix[0]]) * (yn[ix[1]] - yn[ix[0]]))
    tot = area((0, 1, 2)).sum() + (area((1, 3, 2)).sum() if quad else 0.0)
    assert abs(tot - 4.0) < 1e-9  # coarse_box covers [-1,1]^2


@pytest.mark.parametrize("quad,N,fmt", [(False, 3, "binary"), (False, 3, "ascii"), (True, 2, "binary"), (False, 1, "binary")])
def test_vtk_outputter_files(tmp_path, coarse_box_mesh, coarse_quads_mesh, quad, N, fmt):
    from oracle.output_numpy import insert_all_cells

    P = _prov(coarse_quads_mesh if quad else coarse_box_mesh, N, quad)
    x, y = P.get_xGrid(), P.get_yGrid()
    eta = np.exp(-x * x - y * y)
    out = VtkOutputter(P, directory=str(tmp_path), fmt=fmt)
    assert out.generateFileName("eta", 120) == "eta0000120.vtu"  # src/VtkOutputter.cpp:47-51
    out.writeFieldsToFiles({"eta": eta, "u": 2 * eta}, 120)
    assert sorted(os.listdir(tmp_path)) == ["eta0000120.vtu", "u0000120.vtu"]
    d = read_vtu(str(tmp_path / "eta0000120.vtu"))
    xs, ys, fs = (x, y, eta) if N == 1 else P.splitElements(x, y, eta)  # N = 1: written as is (VtkOutputter.hpp:55)
    pts, vals, cells = insert_all_cells(xs, ys, fs)
    assert d["scalars"] == "eta" and d["npoints"] == len(pts) and d["ncells"] == len(cells)
    assert np.array_equal(d["arrays"]["Points"], pts)
    assert np.array_equal(d["arrays"]["eta"], vals)
    assert np.array_equal(d["arrays"]["connectivity"].reshape(cells.shape), cells)
    assert np.array_equal(d["arrays"]["offsets"], (np.arange(len(cells)) + 1) * cells.shape[1])
    assert (d["arrays"]["types"] == (9 if quad else 5)).all()
    with pytest.raises(ValueError):
        write_vtu(str(tmp_path / "bad.vtu"), np.zeros((5, 2)), np.zeros((5, 2)), np.zeros((5, 2)), "f")


def test_csv_outputter(tmp_path):
    out = CsvOutputter(str(tmp_path))
    assert out.generateFileName("h", 7) == "h0000007.dat"  # src/CsvOutputter.cpp:46-50
    f = np.array([[1.0, 2.5, -3.25], [1e-7, 123456789.0, 0.1]])
    out.writeFieldsToFiles({"h": f}, 7)
    lines = open(tmp_path / "h0000007.dat").read().splitlines()
    assert lines == ["1 2.5 -3.25 ", "1e-07 1.23457e+08 0.1 "]  # ostream << double: 6 significant digits, trailing delimiter


# ------------------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
@pytest.mark.parametrize("shape,N,nx,ny,jit,bathy", [("tri", 4, 40, 20, 0.0, True), ("tri", 4, 23, 7, 0.2, False), ("quad", 3, 16, 8, 0.2, True),
                                                     ("tri", 8, 12, 6, 0.0, True), ("quad", 6, 9, 5, 0.0, True)])
def test_device_output_matches_host_path(tmp_path, shape, N, nx, ny, jit, bathy):
    """eta, u, v interpolated on the device from the live state = the reference's host-side derivation + IM product,
    before and after stepping, with the copies overlapping further steps; the written files equal the host path's."""
    from blitzdg_b200.outputters import DeviceOutput
    from blitzdg_b200.sw2d import SW2dSolver
    from oracle.output_numpy import derived_fields

    P = provision(structured_mesh(shape, nx, ny, jit), N)
    ctx = P.get_DGContext()
    H, h, hu, hv = gaussian_hump(ctx.x, ctx.y)
    rng = np.random.default_rng(2)
    hu = 0.3 * rng.standard_normal(h.shape)
    hv = 0.3 * rng.standard_normal(h.shape)
    sol = SW2dSolver(ctx)
    if bathy:
        H = H + 0.5 * ctx.x
        sol.set_bathymetry(H)
    sol.set_state(h, hu, hv)
    dev = DeviceOutput(sol, P, directory=str(tmp_path))
    IM = dev.IM

    def check(tstep, state):
        t, flds = dev.fields()
        assert t == tstep
        want = derived_fields(*state, H if bathy else 0.0 * H)
        for name in ("eta", "u", "v"):
            ref = IM @ want[name]
            assert np.abs(flds[name] - ref).max() <= 1e-12 * max(np.abs(ref).max(), 1.0), name

    dev.begin(0)
    check(0, (h, hu, hv))
    dt = 0.5 * sol.computeTimeStep()
    sol.step(5, dt, sync=False)
    dev.begin(5)            # snapshot after 5 steps ...
    sol.step(5, dt, sync=False)   # ... while the next interval is already running
    t, flds = dev.fields()
    snap = {k: v.copy() for k, v in flds.items()}
    names = dev.write()
    assert [os.path.basename(n) for n in names] == ["eta0000005.vtu", "u0000005.vtu", "v0000005.vtu"]
    sol.sync()
    # replay on a second handle to get the state at step 5
    sol2 = SW2dSolver(ctx)
    if bathy:
        sol2.set_bathymetry(H)
    sol2.set_state(h, hu, hv)
    sol2.step(5, dt)
    h5, hu5, hv5, _ = sol2.get_state()
    want = derived_fields(h5, hu5, hv5, H if bathy else 0.0 * H)
    for name in ("eta", "u", "v"):
        ref = IM @ want[name]
        assert np.abs(snap[name] - ref).max() <= 1e-12 * max(np.abs(ref).max(), 1.0), name
    # file written from the device-fed arrays == file the host path writes from the downloaded state
    VtkOutputter(P, directory=str(tmp_path / "..")).writeFieldToFile("host_eta.vtu", want["eta"], "eta")
    a, b = read_vtu(names[0]), read_vtu(str(tmp_path / ".." / "host_eta.vtu"))
    assert np.array_equal(a["arrays"]["connectivity"], b["arrays"]["connectivity"])
    assert np.abs(a["arrays"]["Points"] - b["arrays"]["Points"]).max() < 1e-13
    assert np.abs(a["arrays"]["eta"] - b["arrays"]["eta"]).max() < 1e-12
    # the time loop was not disturbed by the snapshot
    h10, _, _, _ = sol.get_state()
    sol2.step(5, dt)
    assert np.array_equal(h10, sol2.get_state()[0])
    sol.close()
    sol2.close()


@pytest.mark.gpu
def test_device_output_writer_thread(tmp_path):
    """Snapshots every 4 steps drained to .vtu files by the writer thread while the loop runs; each file equals what the
    host path writes from a replay of the same steps."""
    from blitzdg_b200.outputters import DeviceOutput
    from blitzdg_b200.sw2d import SW2dSolver

    P = provision(structured_mesh("tri", 30, 16), 4)
    ctx = P.get_DGContext()
    H, h, hu, hv = gaussian_hump(ctx.x, ctx.y)
    sol = SW2dSolver(ctx)
    sol.set_bathymetry(H)
    sol.set_state(h, hu, hv)
    dt = 0.5 * sol.computeTimeStep()
    dev = DeviceOutput(sol, P, directory=str(tmp_path))
    dev.start_writer()
    for count in range(4, 25, 4):
        sol.step(4, dt, sync=False)
        dev.begin(count)
    names = dev.finish()
    sol.sync()
    assert sorted(os.path.basename(n) for n in names) == sorted("%s%07d.vtu" % (f, c) for c in range(4, 25, 4) for f in ("eta", "u", "v"))
    sol2 = SW2dSolver(ctx)
    sol2.set_bathymetry(H)
    sol2.set_state(h, hu, hv)
    for count in range(4, 25, 4):
        sol2.step(4, dt)
        hh, hhu, hhv, _ = sol2.get_state()
        _, _, want = P.splitElements(ctx.x, ctx.y, hhu / hh)
        got = read_vtu(str(tmp_path / ("u%07d.vtu" % count)))["arrays"]["u"]
        assert np.abs(got - want.T.reshape(-1)).max() < 1e-12
    sol.close()
    sol2.close()
