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
