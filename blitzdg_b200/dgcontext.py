"""DGContext2D: the read-only operator / geometry / map contract of the hot path.

Mirrors /root/reference/include/DGContext2D.hpp:9-196.  All arrays are kept in
the reference's native layouts:

  Dr, Ds, Filt, V, Vinv : float64 [Np][Np]
  Lift                  : float64 [Np][Nfaces*Nfp]
  x, y, J, rx, ry, sx, sy : float64 [Np][K]   (row-major: element index fastest)
  nx, ny, Fscale        : float64 [Nfaces*Nfp][K]
  Fmask                 : int32   [Nfp][Nfaces]
  vmapM, vmapP          : int32   [Nfaces*Nfp*K], values n + Np*k
  BCmap                 : dict tag -> int32 list of face-node indices in [0, Nfaces*Nfp*K)
"""
from dataclasses import dataclass, field
from typing import Dict, Optional

import numpy as np


@dataclass
class DGContext2D:
    N: int
    Np: int
    Nfp: int
    Nfaces: int
    K: int
    Dr: np.ndarray
    Ds: np.ndarray
    Lift: np.ndarray
    rx: np.ndarray
    ry: np.ndarray
    sx: np.ndarray
    sy: np.ndarray
    nx: np.ndarray
    ny: np.ndarray
    Fscale: np.ndarray
    vmapM: np.ndarray
    vmapP: np.ndarray
    BCmap: Dict[int, np.ndarray] = field(default_factory=dict)
    J: Optional[np.ndarray] = None
    x: Optional[np.ndarray] = None
    y: Optional[np.ndarray] = None
    r: Optional[np.ndarray] = None
    s: Optional[np.ndarray] = None
    Fmask: Optional[np.ndarray] = None
    V: Optional[np.ndarray] = None
    Vinv: Optional[np.ndarray] = None
    Filt: Optional[np.ndarray] = None

    # accessor names of the reference (DGContext2D.hpp:82-196)
    def order(self):
        return self.N

    def numLocalPoints(self):
        return self.Np

    def numFacePoints(self):
        return self.Nfp

    def numElements(self):
        return self.K

    def numFaces(self):
        return self.Nfaces

    def filter(self):
        return self.Filt

    def fscale(self):
        return self.Fscale

    def lift(self):
        return self.Lift

    def bcmap(self):
        return self.BCmap

    def jacobian(self):
        return self.J


def as_context(ctx):
    """Accepts what a caller holds for the context and returns a DGContext2D of this package:
      * a DGContext2D of this package: returned as it is;
      * a pyblitzdg-style context -- the PROPERTY names of the reference's Python binding (src/pyblitzdg/pyblitzdg.cpp:160-187:
        numLocalPoints, numFacePoints, numElements, numFaces, filter, Dr, Ds, Lift, rx .. sy, nx, ny, Fscale, vmapM, vmapP,
        BCmap, J, x, y, r, s, Fmask, V), whether it comes from pyblitzdg itself or from blitzdg_b200.pyblitzdg: converted (the
        binding has no property for the order; N = numFacePoints - 1)."""
    if isinstance(ctx, DGContext2D):
        return ctx
    inner = getattr(ctx, "_mirror", None)
    if isinstance(inner, DGContext2D):
        return inner
    if callable(getattr(ctx, "numLocalPoints", None)) or not hasattr(ctx, "numLocalPoints"):
        if hasattr(ctx, "Np") and hasattr(ctx, "K"):
            return ctx  # duck-typed holder with this package's field names (tests use such views)
        raise TypeError("as_context: neither a blitzdg_b200 DGContext2D nor an object with pyblitzdg's DGContext2D properties")
    c = np.ascontiguousarray

    def opt(name):
        try:
            v = getattr(ctx, name)
        except Exception:
            return None
        return None if v is None else np.asarray(v)

    Nfp = int(ctx.numFacePoints)
    bc = {int(k): np.asarray(v, dtype=np.int32).reshape(-1) for k, v in dict(ctx.BCmap).items()}
    filt = opt("filter")
    return DGContext2D(
        N=Nfp - 1, Np=int(ctx.numLocalPoints), Nfp=Nfp, Nfaces=int(ctx.numFaces), K=int(ctx.numElements),
        Dr=c(ctx.Dr, dtype=np.float64), Ds=c(ctx.Ds, dtype=np.float64), Lift=c(ctx.Lift, dtype=np.float64),
        rx=c(ctx.rx, dtype=np.float64), ry=c(ctx.ry, dtype=np.float64), sx=c(ctx.sx, dtype=np.float64), sy=c(ctx.sy, dtype=np.float64),
        nx=c(ctx.nx, dtype=np.float64), ny=c(ctx.ny, dtype=np.float64), Fscale=c(ctx.Fscale, dtype=np.float64),
        vmapM=c(np.asarray(ctx.vmapM).reshape(-1), dtype=np.int32), vmapP=c(np.asarray(ctx.vmapP).reshape(-1), dtype=np.int32), BCmap=bc,
        J=opt("J"), x=opt("x"), y=opt("y"), r=opt("r"), s=opt("s"), Fmask=opt("Fmask"), V=opt("V"),
        Filt=None if filt is None or filt.size == 0 else c(filt, dtype=np.float64))
