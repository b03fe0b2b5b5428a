"""TEST INFRASTRUCTURE -- ctypes access to oracle/_ref/libmesh_ref.so: the reference's OWN MeshManager::buildConnectivity,
buildBCTable(tag) and counter-clockwise repair, compiled verbatim by oracle/make_ref_mesh.py.  Used to pin the host mirror's
mesh tables (tests/test_reference_compiled_setup.py); the product path never touches it."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(_HERE, "_ref", "libmesh_ref.so")
_dp, _ip = C.POINTER(C.c_double), C.POINTER(C.c_int)


def available():
    if not os.path.exists(LIB) and os.path.exists("/root/reference/src/MeshManager.cpp"):
        import subprocess
        import sys

        subprocess.call([sys.executable, os.path.join(_HERE, "make_ref_mesh.py")], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return os.path.exists(LIB)


def mesh_tables(etov, vert, tag=3):
    """etov [K][Nf] as read from the file (0-based, any orientation), vert [Nv][3] -> (EToV, EToE, EToF, BCType), each [K][Nf],
    as the tail of MeshManager::readMesh / buildMesh leaves them."""
    lib = C.CDLL(LIB)
    etov = np.ascontiguousarray(etov, dtype=np.int32)
    vert = np.ascontiguousarray(vert, dtype=np.float64)
    K, Nf = etov.shape
    out = [np.zeros((K, Nf), dtype=np.int32) for _ in range(4)]
    rc = lib.ref_mesh_tables(Nf, K, vert.shape[0], etov.ctypes.data_as(_ip), vert.ctypes.data_as(_dp), int(tag), *[o.ctypes.data_as(_ip) for o in out])
    assert rc == 0
    return out
