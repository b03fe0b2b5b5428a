"""TEST INFRASTRUCTURE -- ctypes access to oracle/liboracle.so (built by oracle/Makefile).

Importable only from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from blitzdg_b200.capi import bdg_dgcontext2d, bdg_sw2d_params, make_params, pack_context

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "liboracle.so")
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)
_lib = None


def build():
    subprocess.check_call(["make", "-C", _HERE, "-s"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            build()
        L = C.CDLL(_LIB)
        L.oracle_sw2d_rhs.argtypes = [C.POINTER(bdg_dgcontext2d), C.POINTER(bdg_sw2d_params), _dp, _dp, _dp, _dp, _dp, _dp,
                                      C.c_double, C.c_int, _dp, _dp, _dp, _dp]
        L.oracle_sw2d_timestep.argtypes = [C.POINTER(bdg_dgcontext2d), C.POINTER(bdg_sw2d_params), _dp, _dp, _dp, C.c_int, _dp]
        L.oracle_sw2d_timestep.restype = C.c_double
        L.oracle_sw2d_step.argtypes = [C.POINTER(bdg_dgcontext2d), C.POINTER(bdg_sw2d_params), _dp, _dp, _dp, _dp, _dp, _dp,
                                       _dp, _dp, _dp, C.c_int, C.c_double, _dp]
        L.oracle_sw2d_step_ssprk2.argtypes = [C.POINTER(bdg_dgcontext2d), C.POINTER(bdg_sw2d_params), _dp, _dp, _dp, _dp, _dp, _dp,
                                              _dp, C.c_int, C.c_double, _dp]
        L.oracle_build_sponge.argtypes = [C.POINTER(bdg_dgcontext2d), _dp, _dp, _ip, C.c_int, C.c_double, C.c_double, _dp]
        L.oracle_build_sponge.restype = None
        L.oracle_depth_to_H.argtypes = [_dp, C.c_size_t, C.c_int, C.c_int, _dp]
        L.oracle_depth_to_H.restype = None
        L.oracle_bed_slopes.argtypes = [C.POINTER(bdg_dgcontext2d), _dp, _dp, _dp]
        L.oracle_bed_slopes.restype = None
        L.oracle_sw2d_rhs_tracer.argtypes = [C.POINTER(bdg_dgcontext2d), C.POINTER(bdg_sw2d_params), _dp, _dp, _dp, _dp, _dp]
        L.oracle_lserk4_stage.argtypes = [C.c_int, C.c_double, C.c_size_t, _dp, _dp, _dp]
        L.oracle_lserk4_stage.restype = None
        L.oracle_lserk4_coeffs.argtypes = [_dp, _dp]
        L.oracle_full_to_vector.argtypes = [_dp, _dp, C.c_int, C.c_int]
        L.oracle_vector_to_full.argtypes = [_dp, _dp, C.c_int, C.c_int]
        L.oracle_apply_index_map.argtypes = [_dp, _ip, _dp, C.c_int]
        _lib = L
    return _lib


def _d(a):
    return a.ctypes.data_as(_dp) if a is not None else _dp()


def _c(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float64)


class Oracle:
    """CPU restatement bound to one DGContext2D + parameter set."""

    def __init__(self, ctx, **params):
        self.ctx = ctx
        self.cctx, self._keep = pack_context(ctx)
        self.prm = make_params(**params)
        self.H = self.Hx = self.Hy = None

    def set_bathymetry(self, H, Hx, Hy):
        self.H, self.Hx, self.Hy = _c(H), _c(Hx), _c(Hy)

    def rhs(self, h, hu, hv, t=0.0, ghost=None):
        h, hu, hv = _c(h), _c(hu), _c(hv)
        out = [np.empty_like(h) for _ in range(3)]
        ng = 0
        if ghost is not None:
            ghost = _c(ghost)
            ng = ghost.shape[1]
        lib().oracle_sw2d_rhs(C.byref(self.cctx), C.byref(self.prm), _d(self.H), _d(self.Hx), _d(self.Hy), _d(h), _d(hu),
                              _d(hv), t, ng, _d(ghost), _d(out[0]), _d(out[1]), _d(out[2]))
        return out

    def rhs_tracer(self, h, hu, hv, hN):
        """RHS4 of the tracer equation (swhelpers/rhs.py:206-297), per-face LF."""
        h, hu, hv, hN = _c(h), _c(hu), _c(hv), _c(hN)
        out = np.empty_like(h)
        lib().oracle_sw2d_rhs_tracer(C.byref(self.cctx), C.byref(self.prm), _d(h), _d(hu), _d(hv), _d(hN), _d(out))
        return out

    def step_with_tracer(self, h, hu, hv, hN, res, nsteps, dt):
        """In-place nsteps LSERK4 steps of flow + tracer (res: 4 residual arrays); every stage evaluates both RHS on the
        same pre-stage state, as sw2d.py:129-140 does."""
        for _ in range(nsteps):
            for st in range(5):
                R = self.rhs(h, hu, hv)
                R4 = self.rhs_tracer(h, hu, hv, hN)
                for rhs, r, q in zip(R + [R4], res, (h, hu, hv, hN)):
                    self.stage(st, dt, rhs, r, q)

    def timestep(self, h, hu, hv, use_abs=False):
        h, hu, hv = _c(h), _c(hu), _c(hv)
        mx = C.c_double(0)
        dt = lib().oracle_sw2d_timestep(C.byref(self.cctx), C.byref(self.prm), _d(h), _d(hu), _d(hv), int(use_abs), C.byref(mx))
        return dt, mx.value

    def step(self, h, hu, hv, res, nsteps, dt, t=0.0):
        """In-place nsteps LSERK4 steps; returns the new time."""
        tt = C.c_double(t)
        lib().oracle_sw2d_step(C.byref(self.cctx), C.byref(self.prm), _d(self.H), _d(self.Hx), _d(self.Hy), _d(h), _d(hu), _d(hv),
                               _d(res[0]), _d(res[1]), _d(res[2]), nsteps, dt, C.byref(tt))
        return tt.value

    @staticmethod
    def stage(stage, dt, rhs, res, q):
        lib().oracle_lserk4_stage(stage, dt, rhs.size, _d(rhs), _d(res), _d(q))

    # ---- the rest of the src/sw2d driver ----
    def step_ssprk2(self, h, hu, hv, nsteps, dt, t=0.0, sponge=None):
        """In-place nsteps steps of the driver's SSP-RK2 + sponge loop; returns the new time."""
        tt = C.c_double(t)
        sponge = _c(sponge)
        lib().oracle_sw2d_step_ssprk2(C.byref(self.cctx), C.byref(self.prm), _d(self.H), _d(self.Hx), _d(self.Hy), _d(sponge),
                                      _d(h), _d(hu), _d(hv), nsteps, dt, C.byref(tt))
        return tt.value

    def build_sponge(self, x, y, out_nodes, strength, radInfl):
        x, y = _c(x), _c(y)
        out_nodes = np.ascontiguousarray(out_nodes, dtype=np.int32)
        sp = np.zeros_like(x)
        lib().oracle_build_sponge(C.byref(self.cctx), _d(x), _d(y), out_nodes.ctypes.data_as(_ip), out_nodes.size, strength, radInfl, _d(sp))
        return sp

    def bed_slopes(self, H):
        H = _c(H)
        Hx, Hy = np.empty_like(H), np.empty_like(H)
        lib().oracle_bed_slopes(C.byref(self.cctx), _d(H), _d(Hx), _d(Hy))
        return Hx, Hy

    @staticmethod
    def depth_to_H(values, Np, K):
        values = np.ascontiguousarray(values, dtype=np.float64)
        H = np.empty((Np, K))
        lib().oracle_depth_to_H(_d(values), values.size, Np, K, _d(H))
        return H


# ---- extended-precision arbiter (oracle/sw2d_arbiter.c) -------------------------------------------------------------
_arb = {}


def arbiter_lib(quad=False):
    name = "libarbiter_q.so" if quad else "libarbiter.so"
    if name not in _arb:
        path = os.path.join(_HERE, name)
        if not os.path.exists(path):
            subprocess.check_call(["make", "-C", _HERE, "-s", name], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        L = C.CDLL(path)
        u8 = C.POINTER(C.c_uint8)
        L.arbiter_sw2d_rhs.argtypes = [C.POINTER(bdg_dgcontext2d), C.POINTER(bdg_sw2d_params), u8, _dp, _dp, _dp, _dp, _dp, _dp,
                                       C.c_double, C.c_int, _dp, C.c_double, C.c_int, C.c_int, _dp, _dp, _dp]
        L.arbiter_sw2d_max_speed.argtypes = [C.POINTER(bdg_dgcontext2d), C.POINTER(bdg_sw2d_params), u8, _dp, _dp, _dp, _dp,
                                             C.c_double, C.c_int, _dp, C.c_int, C.c_int]
        L.arbiter_sw2d_max_speed.restype = C.c_double
        _arb[name] = L
    return _arb[name]


class Arbiter:
    """Extended-precision evaluation of the same RHS from the same fp64 inputs: the yardstick for "which fp64 result is
    closer to the exact value".  rhs(..., elements=(k0, k1)) evaluates a range only (other columns of the result are NaN);
    `threads` ranges run concurrently (ctypes releases the GIL)."""

    def __init__(self, ctx, quad=False, **params):
        self.ctx = ctx
        self.cctx, self._keep = pack_context(ctx)
        self.prm = make_params(**params)
        self.L = arbiter_lib(quad)
        self.H = self.Hx = self.Hy = None
        Nfn = ctx.Nfaces * ctx.Nfp
        self.bctag = np.zeros(Nfn * ctx.K, dtype=np.uint8)
        for tag, ids in ctx.BCmap.items():
            self.bctag[np.asarray(ids, dtype=np.int64)] = int(tag)

    def set_bathymetry(self, H, Hx, Hy):
        self.H, self.Hx, self.Hy = _c(H), _c(Hx), _c(Hy)

    def _ranges(self, elements, threads):
        k0, k1 = elements if elements is not None else (0, self.ctx.K)
        threads = max(1, min(int(threads), k1 - k0))
        cuts = [k0 + (k1 - k0) * i // threads for i in range(threads + 1)]
        return [(cuts[i], cuts[i + 1]) for i in range(threads)]

    def _run(self, fn, ranges):
        import threading

        res = [None] * len(ranges)

        def work(i):
            res[i] = fn(*ranges[i])

        ths = [threading.Thread(target=work, args=(i,)) for i in range(len(ranges))]
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        return res

    def rhs(self, h, hu, hv, t=0.0, ghost=None, elements=None, threads=None):
        h, hu, hv = _c(h), _c(hu), _c(hv)
        threads = threads or (os.cpu_count() or 1)
        ng = 0
        if ghost is not None:
            ghost = _c(ghost)
            ng = ghost.shape[1]
        tags = self.bctag.ctypes.data_as(C.POINTER(C.c_uint8))
        ranges = self._ranges(elements, threads)
        lam = 0.0
        if self.prm.lf_mode == 1:  # global LF speed: over ALL elements, whatever range is asked for
            allr = self._ranges(None, threads)
            lam = max(self._run(lambda a, b: self.L.arbiter_sw2d_max_speed(C.byref(self.cctx), C.byref(self.prm), tags, _d(self.H), _d(h),
                                                                           _d(hu), _d(hv), t, ng, _d(ghost), a, b), allr))
        out = [np.full_like(h, np.nan) for _ in range(3)]
        rcs = self._run(lambda a, b: self.L.arbiter_sw2d_rhs(C.byref(self.cctx), C.byref(self.prm), tags, _d(self.H), _d(self.Hx), _d(self.Hy),
                                                             _d(h), _d(hu), _d(hv), t, ng, _d(ghost), lam, a, b, _d(out[0]), _d(out[1]),
                                                             _d(out[2])), ranges)
        assert all(r == 0 for r in rcs)
        return out
