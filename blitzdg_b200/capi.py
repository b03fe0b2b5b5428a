"""ctypes binding of the C ABI declared in include/blitzdg_b200.h.

This is the stub a reference maintainer would add on the Python side (pyblitzdg users); the C++
side is shown in INTEGRATION.md.  The shared library is built in-tree by __graft_entry__.build()
(blitzdg_b200/csrc/libblitzdg_b200.so).  There is no CPU fallback: if the library is missing the
import of the product path fails loudly.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# BDG_B200_LIB selects another build of the SAME sources (tools/build_variant.sh: kernel experiments, the phase profiler)
LIB_PATH = os.environ.get("BDG_B200_LIB") or os.path.join(_HERE, "csrc", "libblitzdg_b200.so")

BDG_OK, BDG_E_INVALID, BDG_E_CUDA, BDG_E_UNSTABLE, BDG_E_NCCL, BDG_E_UNSUPPORTED = range(6)
BDG_NOALLOC_MAX_FACE_NODES = 1 << 28  # include/blitzdg_b200.h
BDG_NCCL_ID_BYTES = 128

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)


class bdg_dgcontext2d(C.Structure):
    _fields_ = [
        ("N", C.c_int32), ("Np", C.c_int32), ("Nfp", C.c_int32), ("Nfaces", C.c_int32), ("K", C.c_int32),
        ("Dr", _dp), ("Ds", _dp), ("Lift", _dp), ("Filt", _dp),
        ("rx", _dp), ("sx", _dp), ("ry", _dp), ("sy", _dp),
        ("nx", _dp), ("ny", _dp), ("Fscale", _dp),
        ("vmapM", _ip), ("vmapP", _ip),
        ("bc_face_nodes", _ip * 9), ("bc_count", C.c_int32 * 9),
    ]


class bdg_sw2d_params(C.Structure):
    _fields_ = [
        ("g", C.c_double), ("CD", C.c_double), ("f", C.c_double), ("CFL", C.c_double),
        ("lf_mode", C.c_int32), ("sources", C.c_int32), ("filter", C.c_int32), ("reserved", C.c_int32),
        ("tide_amp", C.c_double), ("tide_period", C.c_double), ("tide_ramp", C.c_double),
    ]


class bdg_halo(C.Structure):
    _fields_ = [
        ("rank", C.c_int32), ("nranks", C.c_int32), ("n_ghost", C.c_int32), ("n_neighbors", C.c_int32),
        ("nbr_rank", _ip), ("send_offset", _ip), ("send_nodes", _ip), ("recv_offset", _ip),
        ("n_interior", C.c_int32),
    ]


class bdg_curved2d(C.Structure):
    """include/blitzdg_b200.h: cubature volume mesh + Gauss face nodes as sw2dComputeRHS_curved reads them."""
    _fields_ = [
        ("Ncub", C.c_int32), ("NGauss", C.c_int32),
        ("cubV", _dp), ("cubDr", _dp), ("cubDs", _dp),
        ("cubW", _dp), ("cub_rx", _dp), ("cub_sx", _dp), ("cub_ry", _dp), ("cub_sy", _dp),
        ("MMChol", _dp), ("curvedEls", _ip), ("n_curved", C.c_int32), ("n_gmapW", C.c_int32),
        ("gInterp", _dp), ("gW", _dp), ("gnx", _dp), ("gny", _dp),
        ("gmapM", _ip), ("gmapP", _ip), ("gmapW", _ip), ("V", _dp), ("J", _dp),
    ]


def _d(a):
    return a.ctypes.data_as(_dp) if a is not None else _dp()


def _i(a):
    return a.ctypes.data_as(_ip) if a is not None else _ip()


def make_params(g=9.81, CD=2.5e-3, f=1.0070e-4, CFL=0.25, lf_mode=0, sources=0, filter=0,
                tide_amp=3.0, tide_period=3600 * 12.42, tide_ramp=0.15 / 3600):
    """Defaults are the literals of the reference drivers (src/sw2d/main.cpp:49-60,281-283, SW2d.hpp:16)."""
    return bdg_sw2d_params(g, CD, f, CFL, int(lf_mode), int(sources), int(filter), 0, tide_amp, tide_period, tide_ramp)


def pack_curved(ctx, cub, gauss, curvedEls, J=None, gmapM=None, gmapP=None):
    """(DGContext2D, cubature context, Gauss face context, curved element list) -> (bdg_curved2d, keepalive list).
    gmapM / gmapP default to the Gauss context's own maps (sw2d_curved.py passes periodic ones)."""
    keep = []

    def dd(a):
        b = np.ascontiguousarray(a, dtype=np.float64)
        keep.append(b)
        return _d(b)

    def ii(a):
        b = np.ascontiguousarray(a, dtype=np.int32).reshape(-1)
        keep.append(b)
        return _i(b), len(b)

    s = bdg_curved2d()
    s.Ncub, s.NGauss = int(cub.V.shape[0]), int(gauss.Interp.shape[0]) // 3
    s.cubV, s.cubDr, s.cubDs = dd(cub.V), dd(cub.Dr), dd(cub.Ds)
    s.cubW, s.cub_rx, s.cub_sx, s.cub_ry, s.cub_sy = dd(cub.W), dd(cub.rx), dd(cub.sx), dd(cub.ry), dd(cub.sy)
    s.curvedEls, s.n_curved = ii(curvedEls)
    s.MMChol = dd(cub.MMChol) if s.n_curved else _dp()
    s.gInterp, s.gW, s.gnx, s.gny = dd(gauss.Interp), dd(gauss.W), dd(gauss.nx), dd(gauss.ny)
    s.gmapM, _ = ii(gauss.mapM if gmapM is None else gmapM)
    s.gmapP, _ = ii(gauss.mapP if gmapP is None else gmapP)
    s.gmapW, s.n_gmapW = ii(gauss.BCmap.get(3, np.zeros(0, np.int32)))
    s.V, s.J = dd(ctx.V), dd(ctx.J if J is None else J)
    return s, keep


def pack_context(ctx):
    """DGContext2D -> (bdg_dgcontext2d, keepalive list).  Arrays are made C-contiguous float64/int32.  Also accepts a context
    with the property names of the reference's Python binding (pyblitzdg), see dgcontext.as_context."""
    from .dgcontext import as_context

    ctx = as_context(ctx)
    keep = []

    def dd(a):
        if a is None:
            return None
        b = np.ascontiguousarray(a, dtype=np.float64)
        keep.append(b)
        return b

    def ii(a):
        b = np.ascontiguousarray(a, dtype=np.int32)
        keep.append(b)
        return b

    s = bdg_dgcontext2d()
    s.N, s.Np, s.Nfp, s.Nfaces, s.K = ctx.N, ctx.Np, ctx.Nfp, ctx.Nfaces, ctx.K
    for name in ("Dr", "Ds", "Lift", "rx", "sx", "ry", "sy", "nx", "ny", "Fscale"):
        setattr(s, name, _d(dd(getattr(ctx, name))))
    s.Filt = _d(dd(ctx.Filt))
    s.vmapM = _i(ii(ctx.vmapM))
    s.vmapP = _i(ii(ctx.vmapP))
    for tag in range(9):
        lst = ctx.BCmap.get(tag)
        if lst is not None and len(lst):
            arr = ii(lst)
            s.bc_face_nodes[tag] = _i(arr)
            s.bc_count[tag] = len(arr)
        else:
            s.bc_face_nodes[tag] = _ip()
            s.bc_count[tag] = 0
    return s, keep


def pack_halo(h):
    """dict(rank,nranks,n_ghost,nbr_rank,send_offset,send_nodes,recv_offset,n_interior) -> bdg_halo."""
    keep = []

    def ii(a):
        b = np.ascontiguousarray(a, dtype=np.int32)
        keep.append(b)
        return b

    s = bdg_halo()
    s.rank, s.nranks, s.n_ghost = int(h["rank"]), int(h["nranks"]), int(h["n_ghost"])
    s.n_neighbors = len(h["nbr_rank"])
    s.nbr_rank = _i(ii(h["nbr_rank"]))
    s.send_offset = _i(ii(h["send_offset"]))
    s.send_nodes = _i(ii(h["send_nodes"]))
    s.recv_offset = _i(ii(h["recv_offset"]))
    s.n_interior = int(h.get("n_interior", 0))
    return s, keep


_lib = None

EXPORTS = [
    "bdg_sw2d_create", "bdg_sw2d_destroy", "bdg_last_error", "bdg_sw2d_set_bathymetry", "bdg_sw2d_set_state",
    "bdg_sw2d_get_state", "bdg_sw2d_rhs", "bdg_sw2d_timestep", "bdg_sw2d_step", "bdg_sw2d_sync", "bdg_sw2d_check",
    "bdg_sw2d_set_variant", "bdg_sw2d_get_variant", "bdg_sw2d_launch_count", "bdg_sw2d_stream",
    "bdg_sw2d_algorithmic_bytes_per_stage", "bdg_nccl_unique_id", "bdg_sw2d_comm_init",
    "bdg_sw2d_halo_pack_to_host", "bdg_sw2d_halo_set_ghosts", "bdg_element_order", "bdg_version",
    "bdg_sw2d_set_sponge", "bdg_sw2d_step_ssprk2", "bdg_sw2d_set_output", "bdg_sw2d_output_begin", "bdg_sw2d_output_wait",
    "bdg_sw2d_limit_positivity", "bdg_sw2d_set_tracer", "bdg_sw2d_get_tracer", "bdg_sw2d_rhs_tracer", "bdg_sw2d_stage_state", "bdg_sw2d_commit_state", "bdg_sw2d_get_state_async", "bdg_sw2d_get_state_wait",
    "bdg_sw2d_timestep_noalloc", "bdg_sw2d_set_curved", "bdg_sw2d_rhs_curved", "bdg_sw2d_step_curved",
]


def load():
    """dlopen the product library; raises (never falls back) when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("blitzdg_b200 CUDA library missing: %s (run __graft_entry__.build()); there is no CPU fallback" % LIB_PATH)
    # NCCL: resolve against the copy torch bundles (same soname), loaded globally first
    try:
        import nvidia.nccl  # type: ignore

        ncclp = os.path.join(list(nvidia.nccl.__path__)[0], "lib", "libnccl.so.2")
        if os.path.exists(ncclp):
            C.CDLL(ncclp, mode=C.RTLD_GLOBAL)
    except Exception:
        pass
    L = C.CDLL(LIB_PATH)
    vp = C.c_void_p
    L.bdg_sw2d_create.argtypes = [C.POINTER(vp), C.POINTER(bdg_dgcontext2d), C.POINTER(bdg_sw2d_params), C.POINTER(bdg_halo), C.c_int]
    L.bdg_sw2d_destroy.argtypes = [vp]
    L.bdg_sw2d_destroy.restype = None
    L.bdg_last_error.argtypes = [vp]
    L.bdg_last_error.restype = C.c_char_p
    L.bdg_sw2d_set_bathymetry.argtypes = [vp, _dp, _dp, _dp]
    L.bdg_sw2d_set_state.argtypes = [vp, _dp, _dp, _dp, C.c_double]
    L.bdg_sw2d_get_state.argtypes = [vp, _dp, _dp, _dp, _dp]
    L.bdg_sw2d_rhs.argtypes = [vp, _dp, _dp, _dp]
    L.bdg_sw2d_timestep.argtypes = [vp, _dp]
    L.bdg_sw2d_step.argtypes = [vp, C.c_int32, C.c_double]
    L.bdg_sw2d_step_ssprk2.argtypes = [vp, C.c_int32, C.c_double]
    L.bdg_sw2d_set_sponge.argtypes = [vp, _dp]
    L.bdg_sw2d_set_output.argtypes = [vp, _dp, C.c_int32]
    L.bdg_sw2d_output_begin.argtypes = [vp, _dp, _dp, _dp]
    L.bdg_sw2d_output_wait.argtypes = [vp]
    L.bdg_sw2d_limit_positivity.argtypes = [vp]
    L.bdg_sw2d_set_tracer.argtypes = [vp, _dp]
    L.bdg_sw2d_get_tracer.argtypes = [vp, _dp]
    L.bdg_sw2d_rhs_tracer.argtypes = [vp, _dp]
    L.bdg_sw2d_stage_state.argtypes = [vp, _dp, _dp, _dp, C.c_double]
    L.bdg_sw2d_commit_state.argtypes = [vp]
    L.bdg_sw2d_get_state_async.argtypes = [vp, _dp, _dp, _dp]
    L.bdg_sw2d_get_state_wait.argtypes = [vp]
    L.bdg_sw2d_sync.argtypes = [vp]
    L.bdg_sw2d_check.argtypes = [vp, _dp]
    L.bdg_sw2d_set_variant.argtypes = [vp, C.c_int32]
    L.bdg_sw2d_get_variant.argtypes = [vp]
    L.bdg_sw2d_launch_count.argtypes = [vp]
    L.bdg_sw2d_launch_count.restype = C.c_int64
    L.bdg_sw2d_stream.argtypes = [vp]
    L.bdg_sw2d_stream.restype = vp
    L.bdg_sw2d_algorithmic_bytes_per_stage.argtypes = [vp]
    L.bdg_sw2d_algorithmic_bytes_per_stage.restype = C.c_double
    L.bdg_nccl_unique_id.argtypes = [vp]
    L.bdg_sw2d_comm_init.argtypes = [vp, vp]
    L.bdg_sw2d_halo_pack_to_host.argtypes = [vp, _dp]
    L.bdg_sw2d_halo_set_ghosts.argtypes = [vp, _dp]
    L.bdg_element_order.argtypes = [_ip, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _ip]
    L.bdg_version.restype = C.c_int
    L.bdg_sw2d_set_curved.argtypes = [vp, C.POINTER(bdg_curved2d)]
    L.bdg_sw2d_rhs_curved.argtypes = [vp, _dp, _dp, _dp, _dp, _dp]
    L.bdg_sw2d_step_curved.argtypes = [vp, C.c_int32, C.c_double, _dp]
    L.bdg_sw2d_timestep_noalloc.argtypes = [C.c_int32] * 6 + [C.c_double, C.c_double, vp, vp, vp, vp, vp, vp, vp, C.c_char_p, C.c_int32]
    _lib = L
    return L
