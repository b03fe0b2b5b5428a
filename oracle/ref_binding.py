"""TEST INFRASTRUCTURE -- ctypes access to oracle/_ref/libsw2d_ref.so: the reference's OWN sw2d::computeRHS (both
drivers), computeTimeStep, buildSpongeCoeff and SSP-RK2 loop body compiled verbatim by oracle/make_ref.py.
Used to pin the oracle (tests/test_reference_compiled.py); the product path never touches it."""
import ctypes as C
import os

import numpy as np

from blitzdg_b200.capi import bdg_dgcontext2d, pack_context

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(_HERE, "_ref", "libsw2d_ref.so")
_dp = C.POINTER(C.c_double)
_lib = None


def available():
    if not os.path.exists(LIB) and os.path.exists("/root/reference/src/sw2d/main.cpp"):
        import subprocess
        import sys

        subprocess.call([sys.executable, os.path.join(_HERE, "make_ref.py")], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return os.path.exists(LIB)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(LIB)
        ctxp = C.POINTER(bdg_dgcontext2d)
        L.ref_sw2d_rhs_full.argtypes = [ctxp, C.c_double, C.c_double] + [_dp] * 6 + [C.c_double] + [_dp] * 3
        L.ref_sw2d_rhs_simple.argtypes = [ctxp, C.c_double] + [_dp] * 6
        L.ref_sw2d_timestep.argtypes = [ctxp, C.c_double] + [_dp] * 3
        L.ref_sw2d_timestep.restype = C.c_double
        L.ref_build_sponge.argtypes = [ctxp, _dp, _dp, C.c_double, C.c_double, _dp]
        L.ref_sw2d_ssprk2.argtypes = [ctxp, C.c_double, C.c_double] + [_dp] * 7 + [C.c_int, C.c_double, _dp]
        _lib = L
    return _lib


def _d(a):
    return a.ctypes.data_as(_dp) if a is not None else _dp()


def _c(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float64)


class CompiledReference:
    def __init__(self, ctx, CD=2.5e-3, f=1.0070e-4):
        self.ctx = ctx
        self.cctx, self._keep = pack_context(ctx)
        self.CD, self.f = CD, f
        self.H = self.Hx = self.Hy = None

    def set_bathymetry(self, H, Hx, Hy):
        self.H, self.Hx, self.Hy = _c(H), _c(Hx), _c(Hy)

    def rhs_full(self, h, hu, hv, t=0.0):
        h, hu, hv = _c(h), _c(hu), _c(hv)
        out = [np.empty_like(h) for _ in range(3)]
        rc = lib().ref_sw2d_rhs_full(C.byref(self.cctx), self.CD, self.f, _d(self.H), _d(self.Hx), _d(self.Hy), _d(h), _d(hu), _d(hv), t,
                                     _d(out[0]), _d(out[1]), _d(out[2]))
        assert rc == 0
        return out

    def rhs_simple(self, h, hu, hv, g=9.81):
        h, hu, hv = _c(h), _c(hu), _c(hv)
        out = [np.empty_like(h) for _ in range(3)]
        rc = lib().ref_sw2d_rhs_simple(C.byref(self.cctx), g, _d(h), _d(hu), _d(hv), _d(out[0]), _d(out[1]), _d(out[2]))
        assert rc == 0
        return out

    def timestep(self, h, hu, hv, CFL=0.25):
        h, hu, hv = _c(h), _c(hu), _c(hv)
        return lib().ref_sw2d_timestep(C.byref(self.cctx), CFL, _d(h), _d(hu), _d(hv))

    def build_sponge(self, x, y, strength, radInfl):
        x, y = _c(x), _c(y)
        sp = np.zeros_like(x)
        assert lib().ref_build_sponge(C.byref(self.cctx), _d(x), _d(y), strength, radInfl, _d(sp)) == 0
        return sp

    def step_ssprk2(self, h, hu, hv, nsteps, dt, t=0.0, sponge=None):
        tt = C.c_double(t)
        sponge = _c(sponge) if sponge is not None else np.zeros_like(h)
        rc = lib().ref_sw2d_ssprk2(C.byref(self.cctx), self.CD, self.f, _d(self.H), _d(self.Hx), _d(self.Hy), _d(sponge), _d(h), _d(hu),
                                   _d(hv), int(nsteps), float(dt), C.byref(tt))
        assert rc == 0
        return tt.value
