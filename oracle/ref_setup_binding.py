"""TEST INFRASTRUCTURE -- ctypes access to oracle/_ref/libsetup_ref.so: the reference's OWN TriangleNodesProvisioner /
QuadNodesProvisioner builders (buildNodes, buildLift, buildPhysicalGrid, buildMaps, buildBCHash, buildFilter), compiled
verbatim by oracle/make_ref_setup.py.  Used to pin the host mirrors that fill DGContext2D
(tests/test_reference_compiled_setup.py); the product path never touches it."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(_HERE, "_ref", "libsetup_ref.so")
_dp, _ip = C.POINTER(C.c_double), C.POINTER(C.c_int)
_lib = None


class _Out(C.Structure):  # struct ref_tri_out of oracle/make_ref_setup.py
    _fields_ = [("r", _dp), ("s", _dp), ("Fmask", _ip), ("V", _dp), ("Vinv", _dp), ("Dr", _dp), ("Ds", _dp), ("Lift", _dp),
                ("x", _dp), ("y", _dp), ("J", _dp), ("rx", _dp), ("sx", _dp), ("ry", _dp), ("sy", _dp), ("nx", _dp), ("ny", _dp),
                ("Fscale", _dp), ("vmapM", _ip), ("vmapP", _ip), ("mapP", _ip), ("mapB", _ip), ("vmapB", _ip), ("nB", C.c_int),
                ("bcmap", _ip), ("Filter", _dp)]


def available():
    if not os.path.exists(LIB) and os.path.exists("/root/reference/src/TriangleNodesProvisioner.cpp"):
        import subprocess
        import sys

        subprocess.call([sys.executable, os.path.join(_HERE, "make_ref_setup.py")], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return os.path.exists(LIB)


def provision(mesh, N, quad=False, filt=None):
    """Runs the reference's constructor body on the mesh tables of `mesh` (EToV, Vert, EToE, EToF, BCType) at order N and
    returns its arrays by the reference's member names (row-major, the reference's shapes); filt = (Nc, s) also runs
    buildFilter.  The Gauss-Jacobi(1,1) points behind the GLL nodes come from the host mirror (LAPACK stand-in)."""
    global _lib
    from blitzdg_b200.jacobi import jacobi_gauss

    if _lib is None:
        _lib = C.CDLL(LIB)
    K, Nv = mesh.get_NumElements(), mesh.get_NumVerts()
    Nf = 4 if quad else 3
    Np, Nfp = ((N + 1) ** 2 if quad else (N + 1) * (N + 2) // 2), N + 1
    Nfn = Nf * Nfp

    def d(*sh):
        return np.zeros(sh)

    def i(*sh):
        return np.zeros(sh, dtype=np.int32)

    a = dict(r=d(Np), s=d(Np), Fmask=i(Nfp, Nf), V=d(Np, Np), Vinv=d(Np, Np), Dr=d(Np, Np), Ds=d(Np, Np), Lift=d(Np, Nfn),
             x=d(Np, K), y=d(Np, K), J=d(Np, K), rx=d(Np, K), sx=d(Np, K), ry=d(Np, K), sy=d(Np, K), nx=d(Nfn, K), ny=d(Nfn, K),
             Fscale=d(Nfn, K), vmapM=i(Nfn * K), vmapP=i(Nfn * K), mapP=i(Nfn * K), mapB=i(Nfn * K), vmapB=i(Nfn * K),
             bcmap=i(8 + Nfn * K), Filter=d(Np, Np))
    o = _Out()
    for k, v in a.items():
        setattr(o, k, v.ctypes.data_as(_dp if v.dtype == np.float64 else _ip))
    if filt is None:
        o.Filter = _dp()

    def ci(v):
        return np.ascontiguousarray(v, dtype=np.int32).reshape(-1)

    EToV, EToE, EToF, BC = ci(mesh.get_Elements()), ci(mesh.get_EToE()), ci(mesh.get_EToF()), ci(mesh.get_BCType())
    Vert = np.ascontiguousarray(mesh.get_Vertices(), dtype=np.float64).reshape(-1)
    if N >= 2:  # computeGaussLobottoPoints(0, 0, N) asks computeJacobiQuadWeights(1, 1, N - 2) for its N - 1 interior points
        gx, gw = (np.ascontiguousarray(v, dtype=np.float64) for v in jacobi_gauss(1.0, 1.0, N - 2))
    else:
        gx = gw = np.zeros(1)
    f = _lib.ref_provision_quads if quad else _lib.ref_provision_triangles
    rc = f(N, K, Nv, EToV.ctypes.data_as(_ip), Vert.ctypes.data_as(_dp), EToE.ctypes.data_as(_ip), EToF.ctypes.data_as(_ip),
           BC.ctypes.data_as(_ip), max(0, N - 1), gx.ctypes.data_as(_dp), gw.ctypes.data_as(_dp),
           C.c_double(filt[0] if filt else 0.0), int(filt[1]) if filt else 0, C.byref(o))
    assert rc == 0
    a["mapB"], a["vmapB"] = a["mapB"][:o.nB].copy(), a["vmapB"][:o.nB].copy()
    bc, pos = {}, 0
    for tag in range(1, 9):
        n = int(a["bcmap"][pos])
        pos += 1
        if n:
            bc[tag] = a["bcmap"][pos:pos + n].copy()
        pos += n
    a["BCmap"] = bc
    del a["bcmap"]
    if filt is None:
        del a["Filter"]
    return a
