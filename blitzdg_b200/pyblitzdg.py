"""`import blitzdg_b200.pyblitzdg as dg` -- the surface of the reference's Python binding for the 2-D shallow-water path.

The reference's Python drivers (sw2d.py, sw2dquads.py, sw2d_curved.py) are written against `pyblitzdg`, the boost.python
module of src/pyblitzdg/pyblitzdg.cpp:97-193.  This module offers the same class names, method names and PROPERTY names for
the classes those drivers touch, over the host mirrors of this package, so that a driver keeps its set-up lines and only
replaces its RHS / time loop by the device path:

    MeshManager                 readMesh, buildMesh, partitionMesh, setBCType; numElements, vertices [Nv][3],
                                elements [K][Nf], bcType [K][Nf], vertexPartitionMap, elementPartitionMap      (:100-112)
    TriangleNodesProvisioner    (N, mesh); buildFilter, setCoordinates, buildGaussFaceNodes, buildCubatureVolumeMesh,
                                dgContext                                                                        (:113-118)
    QuadNodesProvisioner        (N, mesh); buildFilter, dgContext                                                (:120-122)
    DGContext2D                 numLocalPoints, numFacePoints, numElements, numFaces, filter, r, s, x, y, Fscale, Fmask,
                                gather, scatter, J, rx, ry, sx, sy, nx, ny, Dr, Ds, Lift, vmapM, vmapP, BCmap, V  (:160-187)
    GaussFaceContext2D / CubatureContext2D   the dataclasses of blitzdg_b200.curved (same property names, :124-158)
    VtkOutputter                (nodes); writeFieldToFile, writeFieldsToFiles                                    (:189-192)
    VandermondeBuilder          buildVandermondeMatrix(r, includeInverse, order) -> (V, Vinv) or (V,)            (:92-93)
    LSERK4                      numStages, rk4a, rk4b;   BCType.Wall / Dirichlet / Neuman                        (:79-99)

Like the reference's accessors (src/DGContext2D.cpp:24-176), every array property returns a fresh copy.  A DGContext2D of
this module goes straight into blitzdg_b200.sw2d.SW2dSolver (dgcontext.as_context).  `from pyblitzdg import pyblitzdg as dg`
(the spelling of sw2d.py) works too: the module carries itself under that name.  Out of scope, as in SURVEY.md section 2: the
1-D provisioner, the Poisson operator, the sparse solvers.
"""
import sys

import numpy as np

from .bctypes import BCTag  # noqa: F401
from .curved import (CubatureContext2D, GaussFaceContext2D, build_cubature_volume_mesh,  # noqa: F401
                     build_gauss_face_nodes, set_coordinates)
from .mesh_manager import MeshManager as _MeshManager
from .outputters import VtkOutputter as _VtkOutputter
from .quad_nodes import QuadNodesProvisioner as _QuadNodesProvisioner
from .triangle_nodes import TriangleNodesProvisioner as _TriangleNodesProvisioner

pyblitzdg = sys.modules[__name__]  # `from pyblitzdg import pyblitzdg as dg`


class MeshManager(_MeshManager):
    @property
    def numElements(self):
        return self.get_NumElements()

    @property
    def vertices(self):  # MeshManager::get_Vertices_numpy (src/MeshManager.cpp:637-642): [NumVerts][Dim]
        return np.array(self.get_Vertices(), dtype=np.float64).reshape(self.get_NumVerts(), -1)

    @property
    def elements(self):  # get_Elements_numpy (:630-635): [NumElements][NumFaces]
        return np.array(self.get_Elements(), dtype=np.int32).reshape(self.get_NumElements(), self.get_NumFaces())

    @property
    def bcType(self):  # get_BCType_numpy (:644-649)
        return np.array(self.get_BCType(), dtype=np.int32).reshape(self.get_NumElements(), self.get_NumFaces())

    def setBCType(self, bcType):  # set_BCType_numpy (:651-654): the [NumElements][NumFaces] array, row by row
        a = np.asarray(bcType)
        if a.size != self.get_NumElements() * self.get_NumFaces():
            raise ValueError("setBCType: %d entries for %d x %d faces" % (a.size, self.get_NumElements(), self.get_NumFaces()))
        self.set_BCType(np.ascontiguousarray(a, dtype=np.int32).reshape(-1))

    @property
    def vertexPartitionMap(self):
        return np.array(self.get_VertexPartitionMap(), dtype=np.int32)

    @property
    def elementPartitionMap(self):
        return np.array(self.get_ElementPartitionMap(), dtype=np.int32)


class DGContext2D:
    """The property set of src/pyblitzdg/pyblitzdg.cpp:160-187.  Like the reference's context -- raw pointers into the arrays of
    the provisioner that made it (include/DGContext2D.hpp:16-39) -- it reads THROUGH to its provisioner: buildFilter or
    setCoordinates after dgContext() are seen.  `_mirror` is the blitzdg_b200.dgcontext.DGContext2D the device path takes."""

    _PROV_NAME = {"x": "xGrid", "y": "yGrid", "r": "rGrid", "s": "sGrid", "filter": "Filter"}

    def __init__(self, mirror=None, provisioner=None):
        self._fixed, self._prov = mirror, provisioner

    @property
    def _mirror(self):
        return self._prov.get_DGContext() if self._prov is not None else self._fixed

    def _get(self, name, dtype):
        if self._prov is not None:
            v = getattr(self._prov, self._PROV_NAME.get(name, name), None)
        else:
            v = getattr(self._fixed, "Filt" if name == "filter" else name, None)
        return None if v is None else np.array(v, dtype=dtype)

    numLocalPoints = property(lambda self: int(self._prov.NumLocalPoints if self._prov is not None else self._fixed.Np))
    numFacePoints = property(lambda self: int(self._prov.NumFacePoints if self._prov is not None else self._fixed.Nfp))
    numElements = property(lambda self: int(self._prov.NumElements if self._prov is not None else self._fixed.K))
    numFaces = property(lambda self: int(self._prov.NumFaces if self._prov is not None else self._fixed.Nfaces))

    @property
    def BCmap(self):  # DGContext2D::bcmap_numpy: dict tag -> list of face-node ids
        src = self._prov.BCmap if self._prov is not None else self._fixed.BCmap
        return {int(k): [int(i) for i in v] for k, v in src.items()}

    @property
    def gather(self):
        return self._unique()[0]

    @property
    def scatter(self):
        return self._unique()[1]

    def _unique(self):
        """uniquetol over all nodes (src/TriangleNodesProvisioner.cpp:1011-1019): gather = one representative per distinct point,
        scatter = node -> index of its representative.  Sort-based, tolerance 1e-9 like the reference."""
        pts = np.stack([self.x.T.reshape(-1), self.y.T.reshape(-1)], axis=1)
        q = np.round(pts / 1e-9).astype(np.int64)
        _, first, inv = np.unique(q, axis=0, return_index=True, return_inverse=True)
        return first.astype(np.int32), np.asarray(inv).reshape(-1).astype(np.int32)


def _array_property(name, dtype):
    def get(self):
        return self._get(name, dtype)

    get.__name__ = name
    return property(get)


for _n in ("filter", "r", "s", "x", "y", "Fscale", "J", "rx", "ry", "sx", "sy", "nx", "ny", "Dr", "Ds", "Lift", "V"):
    setattr(DGContext2D, _n, _array_property(_n, np.float64))
for _n in ("Fmask", "vmapM", "vmapP"):
    setattr(DGContext2D, _n, _array_property(_n, np.int32))


class _Provisioner2D:
    def dgContext(self):
        return DGContext2D(provisioner=self)


class TriangleNodesProvisioner(_Provisioner2D, _TriangleNodesProvisioner):
    def setCoordinates(self, x, y):  # TriangleNodesProvisioner::setCoordinates_numpy (src/TriangleNodesProvisioner.cpp:1266-1316)
        set_coordinates(self, x, y)

    def buildGaussFaceNodes(self, NGauss):  # (:207-381)
        return build_gauss_face_nodes(self, int(NGauss))

    def buildCubatureVolumeMesh(self, NCubature, rule=None):  # (:81-203); rule: the caller's own cubature points, see curved.py
        return build_cubature_volume_mesh(self, int(NCubature), rule=rule)


class QuadNodesProvisioner(_Provisioner2D, _QuadNodesProvisioner):
    pass


class VtkOutputter(_VtkOutputter):
    def __init__(self, nodesProvisioner, directory=".", fmt="binary"):
        super().__init__(nodesProvisioner, directory, fmt)


class VandermondeBuilder:
    """VandermondeBuilders::buildVandermondeMatrix_numpy (include/VandermondeBuilders.hpp:76-104): the 1-D orthonormal Legendre
    Vandermonde matrix V [len(r)][order + 1] (order = -1: square) and, on request, its inverse; returned as a tuple."""

    def buildVandermondeMatrix(self, r, includeInverse=True, order=-1):
        from . import lapack
        from .jacobi import vandermonde_1d

        r = np.ascontiguousarray(r, dtype=np.float64).reshape(-1)
        V = np.ascontiguousarray(vandermonde_1d(r, ncols=None if order < 0 else int(order) + 1))
        if not includeInverse:
            return (V,)
        return V, lapack.inverse(V)


class _LSERK4:
    numStages = 5

    @property
    def rk4a(self):
        from . import lserk4

        return [float(a) for a in lserk4.rk4a]

    @property
    def rk4b(self):
        from . import lserk4

        return [float(b) for b in lserk4.rk4b]


LSERK4 = _LSERK4()


class _BCType:  # struct BCType of src/pyblitzdg/pyblitzdg.cpp:52-56, injected into the module as an instance
    Dirichlet, Neuman, Wall = 6, 7, 3


BCType = _BCType()
