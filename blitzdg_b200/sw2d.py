"""Host-side mirror of the reference's sw2d driver interface on top of the C ABI.

Names follow /root/reference/src/sw2d/SW2d.hpp:15-55 (physParams, numParams, fields,
computeRHS, computeTimeStep) and the driver loops (src/sw2d/main.cpp:192-244 with the LSERK4 stage
idiom of src/advec1d/main.cpp:92-102).  All arithmetic runs in the CUDA library; this file only
marshals host arrays ([Np][K], row-major) across the boundary.  No CPU fallback.
"""
import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import capi


class NumericalInstability(RuntimeError):
    """The reference throws std::runtime_error("A numerical instability has occurred!")."""


@dataclass
class physParams:  # src/sw2d/SW2d.hpp:15-22, literals src/sw2d/main.cpp:49-51
    g: float = 9.81
    CD: float = 2.5e-3
    f: float = 1.0070e-4
    initTime: float = 0.0
    finalTime: float = 24 * 3600.0


@dataclass
class numParams:  # src/sw2d/SW2d.hpp:24-30
    N: int = 4
    CFL: float = 0.25
    filterPercent: float = 0.90
    filterOrder: int = 4
    outputInterval: int = 20


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double)) if a is not None else C.POINTER(C.c_double)()


class SW2dSolver:
    """One handle = one GPU.  mode 'simple' = sw2d-simple physics (per-face LF, no sources);
    'full' = src/sw2d physics (global LF, star states, bathymetry/drag/Coriolis, Out BC)."""

    def __init__(self, ctx, phys=None, num=None, mode="simple", lf_mode=None, sources=None, use_filter=False,
                 halo=None, device=0, variant=0):
        from .dgcontext import as_context

        self.lib = capi.load()
        ctx = as_context(ctx)  # a pyblitzdg-style context (property names of the reference's Python binding) is converted here
        self.ctx = ctx
        self.phys = phys or physParams()
        self.num = num or numParams(N=ctx.N)
        full = mode == "full"
        lf = (1 if full else 0) if lf_mode is None else int(lf_mode)
        src = (1 if full else 0) if sources is None else int(sources)
        self.prm = capi.make_params(g=self.phys.g, CD=self.phys.CD, f=self.phys.f, CFL=self.num.CFL, lf_mode=lf,
                                    sources=src, filter=int(use_filter))
        self._cctx, self._keep = capi.pack_context(ctx)
        self._chalo, self._keep_h = (capi.pack_halo(halo) if halo is not None else (None, None))
        h = C.c_void_p()
        rc = self.lib.bdg_sw2d_create(C.byref(h), C.byref(self._cctx), C.byref(self.prm),
                                      C.byref(self._chalo) if self._chalo is not None else None, int(device))
        if rc != capi.BDG_OK:
            raise RuntimeError("bdg_sw2d_create failed (%d): %s" % (rc, self.lib.bdg_last_error(None).decode()))
        self.h = h
        self.shape = (ctx.Np, ctx.K)
        if variant:
            self.set_variant(variant)

    def _ck(self, rc):
        if rc == capi.BDG_E_UNSTABLE:
            raise NumericalInstability(self.lib.bdg_last_error(self.h).decode())
        if rc != capi.BDG_OK:
            raise RuntimeError("blitzdg_b200 error %d: %s" % (rc, self.lib.bdg_last_error(self.h).decode()))

    def close(self):
        if getattr(self, "h", None):
            self.lib.bdg_sw2d_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # fields --------------------------------------------------------------------------------------
    def _in(self, a):
        a = np.ascontiguousarray(a, dtype=np.float64)
        assert a.shape == self.shape, (a.shape, self.shape)
        return a

    def set_bathymetry(self, H, Hx=None, Hy=None):
        H = self._in(H)
        Hx = self._in(Hx) if Hx is not None else np.zeros(self.shape)
        Hy = self._in(Hy) if Hy is not None else np.zeros(self.shape)
        self._ck(self.lib.bdg_sw2d_set_bathymetry(self.h, _dp(H), _dp(Hx), _dp(Hy)))

    def set_state(self, h, hu, hv, t=0.0):
        h, hu, hv = self._in(h), self._in(hu), self._in(hv)
        self._ck(self.lib.bdg_sw2d_set_state(self.h, _dp(h), _dp(hu), _dp(hv), float(t)))

    def get_state(self, out=None):
        h, hu, hv = out if out is not None else (np.empty(self.shape), np.empty(self.shape), np.empty(self.shape))
        t = C.c_double(0)
        self._ck(self.lib.bdg_sw2d_get_state(self.h, _dp(h), _dp(hu), _dp(hv), C.byref(t)))
        return h, hu, hv, t.value

    # streaming variants (batches of independent states): copies overlap the steps of the current state
    def stage_state(self, h, hu, hv, t=0.0):
        """h, hu, hv must be C-contiguous float64 [Np][K] (pinned for real overlap) and stay alive until commit_state ran."""
        for a in (h, hu, hv):
            assert a.dtype == np.float64 and a.flags.c_contiguous and a.shape == self.shape
        self._staged = (h, hu, hv)
        self._ck(self.lib.bdg_sw2d_stage_state(self.h, _dp(h), _dp(hu), _dp(hv), float(t)))

    def commit_state(self):
        self._ck(self.lib.bdg_sw2d_commit_state(self.h))

    def get_state_async(self, out):
        for a in out:
            assert a.dtype == np.float64 and a.flags.c_contiguous and a.shape == self.shape
        self._ck(self.lib.bdg_sw2d_get_state_async(self.h, _dp(out[0]), _dp(out[1]), _dp(out[2])))

    def get_state_wait(self):
        self._ck(self.lib.bdg_sw2d_get_state_wait(self.h))

    # hot path ------------------------------------------------------------------------------------
    def computeRHS(self):
        """sw2d::computeRHS on the current device state; returns (RHS1, RHS2, RHS3) as [Np][K]."""
        R = [np.empty(self.shape) for _ in range(3)]
        self._ck(self.lib.bdg_sw2d_rhs(self.h, _dp(R[0]), _dp(R[1]), _dp(R[2])))
        return R

    def computeTimeStep(self):
        dt = C.c_double(0)
        self._ck(self.lib.bdg_sw2d_timestep(self.h, C.byref(dt)))
        return dt.value

    def step(self, nsteps, dt=0.0, sync=True):
        self._ck(self.lib.bdg_sw2d_step(self.h, int(nsteps), float(dt)))
        if sync:
            self.sync()

    def set_sponge(self, spongeCoeff):
        """spongeCoeff [Np][K] of the src/sw2d driver (main.cpp:184-186); None removes it."""
        a = self._in(spongeCoeff) if spongeCoeff is not None else None
        self._ck(self.lib.bdg_sw2d_set_sponge(self.h, _dp(a)))

    def step_ssprk2(self, nsteps, dt=0.0, sync=True):
        """The src/sw2d driver's own loop body (main.cpp:192-244): SSP-RK2 + sponge relaxation, dt <= 0: adaptive."""
        self._ck(self.lib.bdg_sw2d_step_ssprk2(self.h, int(nsteps), float(dt)))
        if sync:
            self.sync()

    # output path (blitzdg_b200/outputters.py::DeviceOutput drives it) ----------------
    def set_output(self, interp):
        """interp [n_out][Np]: the IM of splitElements (NodesProvisioner.computeInterpMatrix)."""
        a = np.ascontiguousarray(interp, dtype=np.float64)
        assert a.ndim == 2 and a.shape[1] == self.shape[0], a.shape
        self._n_out = a.shape[0]
        self._ck(self.lib.bdg_sw2d_set_output(self.h, _dp(a), a.shape[0]))

    def output_begin(self, eta, u, v):
        """eta, u, v: C-contiguous float64 [n_out][K] host arrays (pinned for asynchronous copies) or None."""
        for a in (eta, u, v):
            assert a is None or (a.dtype == np.float64 and a.flags.c_contiguous and a.shape == (self._n_out, self.shape[1]))
        self._ck(self.lib.bdg_sw2d_output_begin(self.h, _dp(eta), _dp(u), _dp(v)))

    def output_wait(self):
        self._ck(self.lib.bdg_sw2d_output_wait(self.h))

    def positivityPreservingLimiter2D(self):
        """swhelpers/limiters.py:3-21 on the device state, in place (the Python drivers call it after a stage)."""
        self._ck(self.lib.bdg_sw2d_limit_positivity(self.h))

    # tracer: the 4th conserved field hN of the Python drivers (sw2d.py:30-33,129-140; swhelpers/rhs.py:206-297) -----------
    def set_tracer(self, hN):
        """hN [Np][K] (None removes it); step() / step_ssprk2() then advance it with the flow."""
        a = self._in(hN) if hN is not None else None
        self._ck(self.lib.bdg_sw2d_set_tracer(self.h, _dp(a)))

    def get_tracer(self):
        out = np.empty(self.shape)
        self._ck(self.lib.bdg_sw2d_get_tracer(self.h, _dp(out)))
        return out

    def computeRHS4(self):
        """RHS4 of sw2dComputeRHS (swhelpers/rhs.py:275-296) on the current device state."""
        out = np.empty(self.shape)
        self._ck(self.lib.bdg_sw2d_rhs_tracer(self.h, _dp(out)))
        return out

    # over-integrated RHS on curved triangles (swhelpers/rhs.py:6-176; contexts: blitzdg_b200/curved.py) --------------------
    def set_curved(self, cub_ctx, gauss_ctx, curvedEls, J=None, gmapM=None, gmapP=None):
        """Uploads the cubature volume mesh, the Gauss face nodes and the list of curved elements (None removes them)."""
        if cub_ctx is None:
            self._ck(self.lib.bdg_sw2d_set_curved(self.h, None))
            return
        c, keep = capi.pack_curved(self.ctx, cub_ctx, gauss_ctx, curvedEls, J, gmapM, gmapP)
        self._ck(self.lib.bdg_sw2d_set_curved(self.h, C.byref(c)))
        del keep  # copied by the library

    def computeRHS_curved(self, CD=None):
        """sw2dComputeRHS_curved on the device state: (RHS1, RHS2, RHS3, RHS4) as [Np][K].  hN = the tracer (zero if none),
        zx, zy = -Hx, -Hy of set_bathymetry, CD = nodal array [Np][K] or None for the parameters' scalar."""
        R = [np.empty(self.shape) for _ in range(4)]
        cd = self._in(np.broadcast_to(np.asarray(CD, dtype=np.float64), self.shape)) if CD is not None else None
        self._ck(self.lib.bdg_sw2d_rhs_curved(self.h, _dp(cd), _dp(R[0]), _dp(R[1]), _dp(R[2]), _dp(R[3])))
        return R

    def step_curved(self, nsteps, dt, CD=None):
        """The curved driver's midpoint predictor / corrector loop on the device (sw2d_curved.py:240-262); the RHS is filtered
        between the stages when the solver was created with use_filter=True."""
        cd = self._in(np.broadcast_to(np.asarray(CD, dtype=np.float64), self.shape)) if CD is not None else None
        self._ck(self.lib.bdg_sw2d_step_curved(self.h, int(nsteps), float(dt), _dp(cd)))

    def sync(self):
        self._ck(self.lib.bdg_sw2d_sync(self.h))

    def check(self):
        """The drivers' blow-up guard; raises NumericalInstability like the reference's throw."""
        m = C.c_double(0)
        self._ck(self.lib.bdg_sw2d_check(self.h, C.byref(m)))
        return m.value

    # plumbing ------------------------------------------------------------------------------------
    def set_variant(self, v):
        self._ck(self.lib.bdg_sw2d_set_variant(self.h, int(v)))

    @property
    def variant(self):
        return self.lib.bdg_sw2d_get_variant(self.h)

    @property
    def launch_count(self):
        return self.lib.bdg_sw2d_launch_count(self.h)

    @property
    def stream(self):
        return self.lib.bdg_sw2d_stream(self.h)

    @property
    def algorithmic_bytes_per_stage(self):
        return self.lib.bdg_sw2d_algorithmic_bytes_per_stage(self.h)

    def comm_init(self, unique_id_bytes):
        buf = C.create_string_buffer(bytes(unique_id_bytes), capi.BDG_NCCL_ID_BYTES)
        self._ck(self.lib.bdg_sw2d_comm_init(self.h, buf))

    def halo_pack_to_host(self, n_send):
        np_ = 4 if self.prm.sources else 3
        buf = np.empty(np_ * n_send)
        self._ck(self.lib.bdg_sw2d_halo_pack_to_host(self.h, _dp(buf)))
        return buf

    def halo_set_ghosts(self, buf):
        buf = np.ascontiguousarray(buf, dtype=np.float64)
        self._ck(self.lib.bdg_sw2d_halo_set_ghosts(self.h, _dp(buf)))


def nccl_unique_id():
    lib = capi.load()
    buf = C.create_string_buffer(capi.BDG_NCCL_ID_BYTES)
    rc = lib.bdg_nccl_unique_id(buf)
    if rc != capi.BDG_OK:
        raise RuntimeError("ncclGetUniqueId failed")
    return buf.raw


# ---- setup steps of the src/sw2d driver around the solver (host side, run once) ------------------------------
def readDepthData(depthFile, Np, K):
    """sw2d::readDepthData (src/sw2d/main.cpp:486-515): one depth per row, element-major (k outer, n inner),
    floored at 150 m; rows missing at the end count as 0 (-> 150).  Returns H [Np][K]."""
    vals = []
    with open(depthFile) as fh:
        for lineno, line in enumerate(fh, 1):
            tok = line.strip().rstrip(",")
            if not tok:
                continue
            try:
                vals.append(float(tok))
            except ValueError:
                raise RuntimeError("readDepthData: %s:%d: not a number: %r" % (depthFile, lineno, tok))
    if len(vals) > Np * K:
        raise RuntimeError("readDepthData: %s holds %d values, the mesh has %d nodes" % (depthFile, len(vals), Np * K))
    d = np.zeros(Np * K)
    d[:len(vals)] = vals
    return np.ascontiguousarray(np.maximum(d, 150.0).reshape(K, Np).T)


def bedSlopes(ctx, H):
    """Hx, Hy of the driver (src/sw2d/main.cpp:129-133): filtered nodal gradient of H."""
    from .dgcontext import as_context

    ctx = as_context(ctx)
    dr, ds = ctx.Dr @ H, ctx.Ds @ H
    Hx, Hy = ctx.rx * dr + ctx.sx * ds, ctx.ry * dr + ctx.sy * ds
    if getattr(ctx, "Filt", None) is not None:
        Hx, Hy = ctx.Filt @ Hx, ctx.Filt @ Hy
    return np.ascontiguousarray(Hx), np.ascontiguousarray(Hy)


def buildSpongeCoeff(ctx, spongeStrength, radInfl, x=None, y=None):
    """sw2d::buildSpongeCoeff (src/sw2d/main.cpp:517-553): strength*(1 - d/radInfl) where d < radInfl is the distance
    from a volume node to the closest node of an Out face.  The reference scans every (node, Out node) pair; here a
    k-d tree answers the same nearest-neighbour query, so meshes of 10^6 elements are provisioned in seconds."""
    from scipy.spatial import cKDTree

    from .bctypes import BCTag
    from .dgcontext import as_context

    ctx = as_context(ctx)
    x = ctx.x if x is None else x
    y = ctx.y if y is None else y
    Np, K = x.shape
    out = np.asarray(ctx.BCmap.get(int(BCTag.Out), ()), dtype=np.int64)
    coeff = np.zeros((Np, K))
    if out.size == 0:
        return coeff  # the reference throws out_of_range here (bcHash.at); an empty sponge is the useful answer
    xv, yv = x.T.reshape(-1), y.T.reshape(-1)  # element-major vectors, the index space of vmapM
    o = np.asarray(ctx.vmapM)[out]
    tree = cKDTree(np.column_stack([xv[o], yv[o]]))
    d, _ = tree.query(np.column_stack([xv, yv]))
    c = np.where(d < radInfl, spongeStrength * (1.0 - d / radInfl), 0.0)
    return np.ascontiguousarray(c.reshape(K, Np).T)
