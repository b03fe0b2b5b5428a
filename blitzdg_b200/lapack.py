"""Dense LAPACK wrappers with the semantics of the reference's host solvers.

* DirectSolver.solve  -> dsgesv (mixed-precision LU + iterative refinement),
  /root/reference/src/DirectSolver.cpp:19-60
* DenseMatrixInverter -> dgetrf + dgetri, /root/reference/src/DenseMatrixinverter.cpp:20-58
* EigenSolver         -> dsyevd,          /root/reference/src/EigenSolver.cpp:20-73

LAPACK comes from the OpenBLAS that scipy bundles (LP64, `scipy_` symbol
prefix): there is no system LAPACK in this image.  Setup-time only.
"""
import ctypes
import glob
import os

import numpy as np

_lib = None


def _openblas():
    global _lib
    if _lib is None:
        import scipy

        d = os.path.join(os.path.dirname(scipy.__file__), "..", "scipy.libs")
        cands = sorted(glob.glob(os.path.join(d, "libscipy_openblas-*.so")))
        cands = [c for c in cands if "openblas64" not in os.path.basename(c)] or cands
        if not cands:
            raise RuntimeError("scipy's bundled OpenBLAS not found")
        _lib = ctypes.CDLL(cands[0])
    return _lib


def _i(v):
    return ctypes.byref(ctypes.c_int(v))


def solve(A, B):
    """X with A X = B, via dsgesv like DirectSolver::solve (row-major in, row-major out)."""
    A = np.asarray(A, dtype=np.float64)
    B = np.asarray(B, dtype=np.float64)
    n = A.shape[0]
    nrhs = B.shape[1]
    a = np.asfortranarray(A.copy())
    b = np.asfortranarray(B.copy())
    x = np.zeros((n, nrhs), order="F")
    ipiv = np.zeros(n, dtype=np.int32)
    work = np.zeros(n * nrhs)
    swork = np.zeros(n * (n + nrhs), dtype=np.float32)
    it = ctypes.c_int(0)
    info = ctypes.c_int(0)
    P = ctypes.c_void_p
    _openblas().scipy_dsgesv_(
        _i(n), _i(nrhs), P(a.ctypes.data), _i(n), P(ipiv.ctypes.data), P(b.ctypes.data), _i(n),
        P(x.ctypes.data), _i(n), P(work.ctypes.data), P(swork.ctypes.data), ctypes.byref(it), ctypes.byref(info))
    if info.value != 0:
        raise RuntimeError("dsgesv failed, info=%d" % info.value)
    return np.ascontiguousarray(x)


def inverse(A):
    """A^{-1} via dgetrf/dgetri like DenseMatrixInverter::computeInverse."""
    from scipy.linalg import lapack

    A = np.asarray(A, dtype=np.float64)
    lu, piv, info = lapack.dgetrf(A)
    if info != 0:
        raise RuntimeError("dgetrf failed, info=%d" % info)
    inv, info = lapack.dgetri(lu, piv)
    if info != 0:
        raise RuntimeError("dgetri failed, info=%d" % info)
    return np.ascontiguousarray(inv)


def eigh(A):
    """Eigenvalues (ascending) and eigenvectors via dsyevd like EigenSolver::solve."""
    from scipy.linalg import lapack

    w, v, info = lapack.dsyevd(np.asarray(A, dtype=np.float64), compute_v=1, lower=0)
    if info != 0:
        raise RuntimeError("dsyevd failed, info=%d" % info)
    return w, v
