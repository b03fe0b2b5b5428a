"""QuadNodesProvisioner: tensor-product GLL nodal-DG operators on bilinear quadrilaterals.

Host-side mirror of /root/reference/src/QuadNodesProvisioner.cpp (citations per
method).  Node id (N+1)*j+i has r = gll(j), s = gll(i) (:218-223); faces
s=-1, r=+1, s=+1, r=-1 with node lists in increasing node-id order (:228-266).
"""
import numpy as np

from . import jacobi, lapack
from .triangle_nodes import TriangleNodesProvisioner, NodeTol


class QuadNodesProvisioner(TriangleNodesProvisioner):
    NumFaces = 4

    def __init__(self, N, mesh, build=True):
        self.NOrder = int(N)
        self.Mesh2D = mesh
        self.NumLocalPoints = (N + 1) * (N + 1)
        self.NumFacePoints = N + 1
        self.NumElements = mesh.get_NumElements() if mesh is not None else 0
        self.Filter = None
        self.BCmap = {}
        if build and mesh is not None:
            self.buildNodes()
            self.buildLift()
            self.buildPhysicalGrid()
            self.buildMaps()

    # basis (:76-112): mode (i,j), i outer in s, j inner in r
    def computeVandermondeMatrix(self, N, r, s):
        V = np.zeros((len(r), (N + 1) * (N + 1)))
        c = 0
        for i in range(N + 1):
            for j in range(N + 1):
                V[:, c] = jacobi.jacobi_polynomial(s, 0.0, 0.0, i) * jacobi.jacobi_polynomial(r, 0.0, 0.0, j)
                c += 1
        return V

    def computeGradVandermondeMatrix(self, N, r, s):
        n = (N + 1) * (N + 1)
        V2Dr, V2Ds = np.zeros((len(r), n)), np.zeros((len(r), n))
        c = 0
        for i in range(N + 1):
            for j in range(N + 1):
                v1dr = jacobi.grad_jacobi(r, 0.0, 0.0, j)
                v1ds = jacobi.grad_jacobi(s, 0.0, 0.0, i)
                v1s = jacobi.jacobi_polynomial(s, 0.0, 0.0, i)
                v1r = jacobi.jacobi_polynomial(r, 0.0, 0.0, j)
                V2Dr[:, c] = v1dr * v1s
                V2Ds[:, c] = v1r * v1ds
                c += 1
        return V2Dr, V2Ds

    def buildNodes(self):
        N = self.NOrder
        g = jacobi.jacobi_gauss_lobatto(0.0, 0.0, N)
        r = np.repeat(g, N + 1)  # id = (N+1)*j + i -> r = g[j]
        s = np.tile(g, N + 1)    #                     s = g[i]
        self.rGrid, self.sGrid = r, s
        f1 = np.nonzero(np.abs(s + 1) < NodeTol)[0]
        f2 = np.nonzero(np.abs(r - 1) < NodeTol)[0]
        f3 = np.nonzero(np.abs(s - 1) < NodeTol)[0]
        f4 = np.nonzero(np.abs(r + 1) < NodeTol)[0]
        self.Fmask = np.stack([f1, f2, f3, f4], axis=1).astype(np.int32)

    def buildLift(self):
        Np, Nfp = self.NumLocalPoints, self.NumFacePoints
        r, s, Fm = self.rGrid, self.sGrid, self.Fmask
        E = np.zeros((Np, 4 * Nfp))
        for f, coord in enumerate((r, s, r, s)):
            V1D = jacobi.vandermonde_1d(coord[Fm[:, f]])
            massEdge = lapack.inverse(V1D @ V1D.T)
            E[Fm[:, f], f * Nfp:(f + 1) * Nfp] = massEdge
        self.V = self.computeVandermondeMatrix(self.NOrder, r, s)
        self.Vinv = lapack.inverse(self.V)
        self.Lift = np.ascontiguousarray((self.V @ self.V.T) @ E)

    def buildPhysicalGrid(self):
        mesh = self.Mesh2D
        K = self.NumElements = mesh.get_NumElements()
        etov = mesh.get_Elements().reshape(K, 4)
        vert = mesh.get_Vertices().reshape(-1, 3)
        VX, VY = vert[:, 0], vert[:, 1]
        va, vb, vc, vd = etov[:, 0], etov[:, 1], etov[:, 2], etov[:, 3]
        self.buildOperators()
        r, s = self.rGrid[:, None], self.sGrid[:, None]
        # bilinear map (:358-362)
        self.xGrid = (0.25 * ((1 - r) * (1 - s)) * VX[va][None, :] + 0.25 * ((1 + r) * (1 - s)) * VX[vb][None, :]
                      + 0.25 * ((1 + r) * (1 + s)) * VX[vc][None, :] + 0.25 * ((1 - r) * (1 + s)) * VX[vd][None, :])
        self.yGrid = (0.25 * ((1 - r) * (1 - s)) * VY[va][None, :] + 0.25 * ((1 + r) * (1 - s)) * VY[vb][None, :]
                      + 0.25 * ((1 + r) * (1 + s)) * VY[vc][None, :] + 0.25 * ((1 - r) * (1 + s)) * VY[vd][None, :])
        self._metrics_and_normals()

    def _metrics_and_normals(self):
        Dr, Ds, x, y = self.Dr, self.Ds, self.xGrid, self.yGrid
        xr, yr, xs, ys = Dr @ x, Dr @ y, Ds @ x, Ds @ y
        J = xr * ys - xs * yr
        self.J = J
        self.rx, self.ry, self.sx, self.sy = ys / J, -xs / J, -yr / J, xr / J
        Nfp = self.NumFacePoints
        fm = self.Fmask.T.reshape(-1)
        fxr, fxs, fyr, fys = xr[fm], xs[fm], yr[fm], ys[fm]
        nx, ny = np.empty_like(fxr), np.empty_like(fxr)
        f1, f2, f3, f4 = (slice(i * Nfp, (i + 1) * Nfp) for i in range(4))
        nx[f1], ny[f1] = fyr[f1], -fxr[f1]    # bottom (:414-432)
        nx[f2], ny[f2] = fys[f2], -fxs[f2]    # right
        nx[f3], ny[f3] = -fyr[f3], fxr[f3]    # top
        nx[f4], ny[f4] = -fys[f4], fxs[f4]    # left
        norm = np.sqrt(nx * nx + ny * ny)
        self.nx, self.ny = nx / norm, ny / norm
        self.Fscale = norm / J[fm]
        self.Fx, self.Fy = x[fm], y[fm]

    def buildFilter(self, Nc, s):
        """Restates the reference's (defective) quad filter: only the first (N+1)(N+2)/2 diagonal
        entries are filled, the rest stay 0 (QuadNodesProvisioner.cpp:185-196)."""
        N, Np = self.NOrder, self.NumLocalPoints
        alpha = -np.log(np.finfo(np.float64).eps)
        d = np.zeros(Np)
        c = 0
        for i in range(N + 1):
            for j in range(N - i + 1):
                if (i + j) >= Nc:
                    k = (float(i + j) - Nc) / (float(N) - Nc)
                    d[c] = np.exp(-alpha * k ** s)
                else:
                    d[c] = 1.0
                c += 1
        self.Filter = np.ascontiguousarray(self.V @ (np.diag(d) @ self.Vinv))
        return self.Filter

    def computeInterpMatrix(self, rout, sout):
        """QuadNodesProvisioner::computeInterpMatrix AS THE REFERENCE EXECUTES IT (src/QuadNodesProvisioner.cpp:707-719): its
        computeVandermondeMatrix (:76-92) ends with `Inverter.computeInverse(V, *Vinv)`, so the call that evaluates the basis at
        the output points overwrites Vinv with the inverse of THAT matrix, and IM = Vout * Vinv = Vout * Vout^-1 -- the identity
        up to rounding (found by compiling the reference's own functions, tests/test_reference_compiled_output.py).  The quad
        output path therefore writes the NODAL values at the nodal (Gauss-Lobatto) positions, in nodal order, connected by the
        lattice table below; nothing is resampled to equispaced points.  Reproduced as executed, except that this mirror leaves
        its own Vinv intact (the reference's stays clobbered for whatever runs next).  Only defined for as many output points
        as nodes, which is how splitElements calls it."""
        Vout = self.computeVandermondeMatrix(self.NOrder, np.asarray(rout, float), np.asarray(sout, float))
        if Vout.shape[0] != Vout.shape[1]:
            raise ValueError("the reference inverts the output Vandermonde matrix: it needs NumLocalPoints output points")
        return np.ascontiguousarray(Vout @ lapack.inverse(Vout))

    def splitTable(self):
        """Equispaced (N+1)x(N+1) output lattice, one linear quad (v1,v2,v3,v4) per cell (src/QuadNodesProvisioner.cpp:728-770)."""
        N = self.NumFacePoints - 1
        rout = np.array([-1.0 + 2.0 * m / N for n in range(N + 1) for m in range(N + 1)])
        sout = np.array([-1.0 + 2.0 * n / N for n in range(N + 1) for m in range(N + 1)])
        c = np.arange((N + 1) * (N + 1)).reshape(N + 1, N + 1)
        cells = [(c[n, m], c[n, m + 1], c[n + 1, m], c[n + 1, m + 1]) for n in range(N) for m in range(N)]
        return rout, sout, np.array(cells, dtype=np.int64)

