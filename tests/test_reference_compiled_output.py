"""Output path against the REFERENCE'S OWN code, compiled verbatim from /root/reference by
oracle/make_ref_output.py (oracle/_ref/libsplit_ref.so): TriangleNodesProvisioner::splitElements, computeInterpMatrix,
computeVandermondeMatrix (+ evaluateSimplexPolynomial, rsToab) and JacobiBuilders::computeJacobiPolynomial.

Held to it: the host mirror (blitzdg_b200/triangle_nodes.py: the interpolation matrix the device output kernel is given, and
splitElements) and the loop-level oracle of the output path (oracle/output_numpy.py).  The library is prebuilt where the
reference tree exists and travels with the snapshot; absent library and absent reference => skip.
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "oracle", "_ref", "libsplit_ref.so")
_dp = C.POINTER(C.c_double)


def _p(a):
    return a.ctypes.data_as(_dp)


@pytest.fixture(scope="module")
def ref():
    if not os.path.exists(LIB):
        if not os.path.exists("/root/reference/src/TriangleNodesProvisioner.cpp"):
            pytest.skip("compiled reference not built and reference tree not present")
        subprocess.run([sys.executable, os.path.join(ROOT, "oracle", "make_ref_output.py")], check=True, capture_output=True)
    return C.CDLL(LIB)


@pytest.fixture(scope="module", params=[1, 2, 4, 7])
def nodes(request, coarse_box_mesh):
    from blitzdg_b200 import TriangleNodesProvisioner

    return TriangleNodesProvisioner(request.param, coarse_box_mesh)


def test_vandermonde_and_interpolation_matrix_match_the_compiled_reference(ref, nodes):
    N, Np = nodes.NOrder, nodes.NumLocalPoints
    rng = np.random.default_rng(N)
    # points inside the reference triangle, plus its three corners (s = 1 takes the a = -1 branch of rsToab, :398-401)
    a, b = rng.random(12), rng.random(12)
    r = np.concatenate([2 * a * (1 - b) - 1 + 0 * b, [-1.0, 1.0, -1.0]])
    s = np.concatenate([2 * b - 1, [-1.0, -1.0, 1.0]])
    r = np.minimum(r, -s)
    n = len(r)
    V = np.zeros((n, Np))
    ref.ref_vandermonde(N, n, _p(r), _p(s), _p(V))
    mine = nodes.computeVandermondeMatrix(N, r, s)
    assert np.abs(mine - V).max() <= 4e-15 * max(1.0, np.abs(V).max())
    IM = np.zeros((n, Np))
    Vinv = np.ascontiguousarray(nodes.Vinv)
    ref.ref_interp_matrix(N, n, _p(Vinv), _p(r), _p(s), _p(IM))
    assert np.abs(nodes.computeInterpMatrix(r, s) - IM).max() <= 1e-13 * max(1.0, np.abs(IM).max())
    # the nodal Vandermonde matrix itself: V(r_nodes) V^-1 = I through the reference's code
    I = np.zeros((Np, Np))
    rn, sn = np.ascontiguousarray(nodes.rGrid), np.ascontiguousarray(nodes.sGrid)
    ref.ref_interp_matrix(N, Np, _p(Vinv), _p(rn), _p(sn), _p(I))
    assert np.abs(I - np.eye(Np)).max() < 1e-11


def test_split_elements_matches_the_compiled_reference(ref, nodes):
    from oracle.output_numpy import split_elements

    N, Np, K = nodes.NOrder, nodes.NumLocalPoints, nodes.NumElements
    if N == 1:
        pytest.skip("N = 1 is written as is (VtkOutputter.hpp:55)")
    x, y = np.ascontiguousarray(nodes.xGrid), np.ascontiguousarray(nodes.yGrid)
    f = np.ascontiguousarray(np.sin(3 * x) * np.cos(2 * y) + 0.3 * x * y)
    tot = N * N * K
    want = [np.zeros((3, tot)) for _ in range(3)]
    Vinv = np.ascontiguousarray(nodes.Vinv)
    n = ref.ref_split_elements(N, K, _p(Vinv), _p(x), _p(y), _p(f), _p(want[0]), _p(want[1]), _p(want[2]))
    assert n == tot
    got = nodes.splitElements(x, y, f)
    rout, sout, _ = nodes.splitTable()
    orc = split_elements(Np, N, False, nodes.computeInterpMatrix(rout, sout), x, y, f)
    for g, o, w in zip(got, orc, want):
        assert g.shape == w.shape == o.shape
        scale = max(1.0, np.abs(w).max())
        assert np.abs(o - w).max() <= 1e-13 * scale   # the loop-level oracle of the output path
        assert np.abs(g - w).max() <= 1e-13 * scale   # the host mirror (BLAS summation order)
    # corner order of the sub-elements is the reference's: counter-clockwise linear triangles with positive area
    xa, ya = want[0], want[1]
    area = (xa[1] - xa[0]) * (ya[2] - ya[0]) - (xa[2] - xa[0]) * (ya[1] - ya[0])
    assert (area > 0).all()


@pytest.mark.parametrize("N", [1, 2, 4, 6])
def test_quad_output_path_follows_the_reference_as_executed(ref, coarse_quads_mesh, N):
    """QuadNodesProvisioner::computeVandermondeMatrix ends with Inverter.computeInverse(V, *Vinv) (src/QuadNodesProvisioner.cpp:90-91),
    so computeInterpMatrix (:707-719) overwrites Vinv with the inverse of the OUTPUT Vandermonde matrix and returns
    Vout * Vout^-1: the identity.  The reference's quad splitElements therefore emits the nodal values at the nodal positions,
    in nodal order, connected by the lattice table.  The host mirror reproduces that (blitzdg_b200/quad_nodes.py)."""
    from blitzdg_b200 import QuadNodesProvisioner
    from oracle.output_numpy import split_elements

    P = QuadNodesProvisioner(N, coarse_quads_mesh)
    Np, K = P.NumLocalPoints, P.NumElements
    rout, sout, _ = P.splitTable()
    rout, sout = np.ascontiguousarray(rout), np.ascontiguousarray(sout)
    Vinv = np.ascontiguousarray(P.Vinv)
    IM, after = np.zeros((Np, Np)), np.zeros((Np, Np))
    ref.ref_quad_interp_matrix(N, _p(Vinv), _p(rout), _p(sout), _p(IM), _p(after))
    assert np.abs(IM - np.eye(Np)).max() < 1e-12                      # what the reference's function returns
    assert np.abs(after - Vinv).max() > 1e-2                          # ... and it has clobbered Vinv on the way
    assert np.abs(P.computeInterpMatrix(rout, sout) - IM).max() < 1e-12
    assert np.array_equal(np.asarray(P.Vinv), Vinv)                   # the mirror leaves its own intact
    if N == 1:
        return
    x, y = np.ascontiguousarray(P.xGrid), np.ascontiguousarray(P.yGrid)
    f = np.ascontiguousarray(np.sin(3 * x) * np.cos(2 * y) + 0.3 * x * y)
    tot = N * N * K
    want = [np.zeros((4, tot)) for _ in range(3)]
    assert ref.ref_quad_split_elements(N, K, _p(Vinv), _p(x), _p(y), _p(f), _p(want[0]), _p(want[1]), _p(want[2])) == tot
    got = P.splitElements(x, y, f)
    orc = split_elements(Np, N, True, P.computeInterpMatrix(rout, sout), x, y, f)
    for g, o, w in zip(got, orc, want):
        assert g.shape == w.shape == o.shape
        assert np.abs(g - w).max() <= 1e-13 * max(1.0, np.abs(w).max())
        assert np.abs(o - w).max() <= 1e-13 * max(1.0, np.abs(w).max())
    # nodal values, not resampled ones: every emitted corner value is one of the element's nodal values
    k = 0
    assert np.abs(np.sort(np.unique(np.round(want[2][:, :N * N], 12))) - np.sort(np.unique(np.round(f[:, k], 12)))).max() < 1e-11


def test_csv_outputter_writes_the_bytes_the_reference_writes(ref, tmp_path):
    """src/CsvOutputter.cpp compiled as it stands: file names and file contents, byte for byte (ostream's default
    formatting of a double -- six significant digits, %g -- including exponents, negative zero, inf and nan)."""
    from blitzdg_b200.outputters import CsvOutputter

    if not hasattr(ref, "ref_csv_write_field"):
        pytest.skip("prebuilt library predates the CsvOutputter entry points")
    rng = np.random.default_rng(5)
    F = rng.standard_normal((7, 13)) * 10.0 ** rng.integers(-12, 12, size=(7, 13))
    F[0, :6] = [0.0, -0.0, 1.0, -1.0, 123456.0, 1234567.0]
    F[1, :6] = [1e-5, 1e-4, 0.1 + 0.2, 1e100, -1e-300, 5e-324]
    F[2, :3] = [np.inf, -np.inf, np.nan]
    F[3, :4] = [999999.5, 99999.95, 0.000123456789, 2.5]
    F = np.ascontiguousarray(F)
    mine = CsvOutputter(str(tmp_path))
    for delim in (" ", ","):
        want = tmp_path / ("ref_%d.dat" % ord(delim))
        ref.ref_csv_write_field(str(want).encode(), _p(F), F.shape[0], F.shape[1], C.c_char(delim.encode()))
        mine.writeFieldToFile("mine.dat", F, delim)
        assert (tmp_path / "mine.dat").read_bytes() == want.read_bytes()
    buf = C.create_string_buffer(64)
    for name, num in (("h", 7), ("eta", 0), ("hu", 1234567), ("v", 12345678)):
        ref.ref_csv_file_name(name.encode(), num, buf, 64)
        assert mine.generateFileName(name, num) == buf.value.decode()
    # writeFieldsToFiles: one file per field under the generated name, space-delimited
    cwd = os.getcwd()
    os.makedirs(tmp_path / "r")
    try:
        os.chdir(tmp_path / "r")
        ref.ref_csv_write_fields(b"h", _p(F), F.shape[0], F.shape[1], 42)
    finally:
        os.chdir(cwd)
    mine.writeFieldsToFiles({"h": F}, 42)
    assert (tmp_path / "h0000042.dat").read_bytes() == (tmp_path / "r" / "h0000042.dat").read_bytes()


def _read_record(path):
    """The flat record the recording VTK stand-in writes (oracle/ref_stub/vtk-7.1/vtk_recording.h)."""
    raw = open(path, "rb").read()
    npts, ncell, nconn, nlen = np.frombuffer(raw[:32], dtype="<i8")
    pos, out = 32, {}
    for name, dt, n in (("xyz", "<f8", 3 * npts), ("scalars", "<f8", npts), ("types", "<i8", ncell), ("offsets", "<i8", ncell), ("conn", "<i8", nconn)):
        out[name] = np.frombuffer(raw[pos:pos + 8 * n], dtype=dt)
        pos += 8 * n
    out["name"] = raw[pos:pos + nlen].decode()
    assert pos + nlen == len(raw)
    return out


@pytest.mark.parametrize("quad,N", [(False, 1), (False, 2), (False, 4), (True, 1), (True, 3)])
def test_vtk_outputter_hands_vtk_what_the_mirror_writes(ref, quad, N, coarse_box_mesh, coarse_quads_mesh, tmp_path):
    """include/VtkOutputter.hpp (writeFieldToFile, insertAllCells) and src/VtkOutputter.cpp (constructor, generateFileName,
    writeFieldsToFiles) compiled over recording VTK stand-ins: the points, point scalars, cell types, connectivity (the quad
    node swap included), active-scalars name and file name the reference hands to VTK are what the mirror's .vtu holds."""
    from blitzdg_b200 import QuadNodesProvisioner, TriangleNodesProvisioner
    from blitzdg_b200.outputters import VtkOutputter, read_vtu

    if not hasattr(ref, "ref_vtk_write_fields"):
        pytest.skip("prebuilt library predates the VtkOutputter entry points")
    P = (QuadNodesProvisioner if quad else TriangleNodesProvisioner)(N, coarse_quads_mesh if quad else coarse_box_mesh)
    x, y = np.ascontiguousarray(P.xGrid), np.ascontiguousarray(P.yGrid)
    f = np.ascontiguousarray(np.exp(-3 * x * x - 2 * y * y) + 0.2 * x * y)
    Vinv = np.ascontiguousarray(P.Vinv)
    name = C.create_string_buffer(64)
    cwd = os.getcwd()
    os.makedirs(tmp_path / "r")
    try:
        os.chdir(tmp_path / "r")
        ref.ref_vtk_write_fields(N, P.NumElements, int(quad), _p(Vinv), _p(x), _p(y), b"eta", _p(f), 120, name, 64)
    finally:
        os.chdir(cwd)
    mine = VtkOutputter(P, str(tmp_path))
    assert mine.generateFileName("eta", 120) == name.value.decode() == "eta0000120.vtu"
    mine.writeFieldsToFiles({"eta": f}, 120)
    got, want = read_vtu(str(tmp_path / "eta0000120.vtu")), _read_record(tmp_path / "r" / "eta0000120.vtu")
    nc = 4 if quad else 3
    ncell = P.NumElements * (N * N if N > 1 else 1)
    assert got["ncells"] == ncell == len(want["types"]) and got["npoints"] == nc * ncell == len(want["scalars"])
    assert got["scalars"] == want["name"] == "eta"
    assert np.array_equal(got["arrays"]["types"].astype(np.int64), want["types"]) and set(want["types"]) == {9 if quad else 5}
    assert np.array_equal(got["arrays"]["connectivity"], want["conn"]) and np.array_equal(got["arrays"]["offsets"], want["offsets"])
    assert np.abs(got["arrays"]["Points"].reshape(-1) - want["xyz"]).max() <= 1e-13
    assert np.abs(got["arrays"]["eta"] - want["scalars"]).max() <= 1e-13 * np.abs(want["scalars"]).max()
