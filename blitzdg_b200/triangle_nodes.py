"""TriangleNodesProvisioner: nodal-DG operators, geometry and maps on straight-sided triangles.

Host-side mirror of /root/reference/src/TriangleNodesProvisioner.cpp (same
method names and array conventions); citations per method.  Vectorised over
elements so that 10^6-10^7-element meshes provision in seconds (the
reference's per-node Python-free loops and quadratic uniquetol do not, F11).
"""
import numpy as np

from . import jacobi, lapack
from .dgcontext import DGContext2D

NodeTol = 1e-5
_pi = 3.1415926535897932


class TriangleNodesProvisioner:
    NumFaces = 3

    def __init__(self, N, mesh, build=True):
        """ctor builds everything (TriangleNodesProvisioner.cpp:39-79)."""
        self.NOrder = int(N)
        self.Mesh2D = mesh
        self.NumLocalPoints = (N + 1) * (N + 2) // 2
        self.NumFacePoints = N + 1
        self.NumElements = mesh.get_NumElements() if mesh is not None else 0
        self.Filter = None
        self.BCmap = {}
        if build and mesh is not None:
            self.buildNodes()
            self.buildLift()
            self.buildPhysicalGrid()
            self.buildMaps()

    # ------------------------------------------------------------ basis (:383-454)
    @staticmethod
    def rsToab(r, s):
        r, s = np.asarray(r, float), np.asarray(s, float)
        a = np.full_like(r, -1.0)
        m = s != 1.0
        a[m] = 2.0 * (1.0 + r[m]) / (1.0 - s[m]) - 1.0
        return a, s.copy()

    @staticmethod
    def xyTors(x, y):
        L1 = (np.sqrt(3.0) * y + 1.0) / 3.0
        L2 = (-3.0 * x - np.sqrt(3.0) * y + 2.0) / 6.0
        L3 = (3.0 * x - np.sqrt(3.0) * y + 2.0) / 6.0
        return -L2 + L3 - L1, -L2 - L3 + L1

    @staticmethod
    def evaluateSimplexPolynomial(a, b, i, j):
        h1 = jacobi.jacobi_polynomial(a, 0.0, 0.0, i)
        h2 = jacobi.jacobi_polynomial(b, 2.0 * i + 1.0, 0.0, j)
        return np.sqrt(2.0) * h1 * h2 * (1.0 - b) ** i

    @staticmethod
    def evaluateGradSimplex(a, b, id, jd):
        """(:642-676)"""
        fa = jacobi.jacobi_polynomial(a, 0.0, 0.0, id)
        gb = jacobi.jacobi_polynomial(b, 2.0 * id + 1.0, 0.0, jd)
        dfa = jacobi.grad_jacobi(a, 0.0, 0.0, id)
        dgb = jacobi.grad_jacobi(b, 2.0 * id + 1.0, 0.0, jd)
        dpdr = dfa * gb
        if id > 1:
            dpdr = dpdr * (0.5 * (1.0 - b)) ** (id - 1)
        dpds = dfa * (gb * (0.5 * (1 + a)))
        if id > 1:
            dpds = dpds * (0.5 * (1.0 - b)) ** (id - 1)
        tmp = dgb * (0.5 * (1.0 - b)) ** id
        if id > 0:
            tmp = tmp - 0.5 * id * gb * (0.5 * (1 - b)) ** (id - 1)
        dpds = dpds + fa * tmp
        return 2.0 ** (id + 0.5) * dpdr, 2.0 ** (id + 0.5) * dpds

    def computeVandermondeMatrix(self, N, r, s):
        a, b = self.rsToab(r, s)
        V = np.zeros((len(r), (N + 1) * (N + 2) // 2))
        c = 0
        for i in range(N + 1):
            for j in range(N - i + 1):
                V[:, c] = self.evaluateSimplexPolynomial(a, b, i, j)
                c += 1
        return V

    def computeGradVandermondeMatrix(self, N, r, s):
        a, b = self.rsToab(r, s)
        n = (N + 1) * (N + 2) // 2
        V2Dr, V2Ds = np.zeros((len(r), n)), np.zeros((len(r), n))
        c = 0
        for i in range(N + 1):
            for j in range(N - i + 1):
                V2Dr[:, c], V2Ds[:, c] = self.evaluateGradSimplex(a, b, i, j)
                c += 1
        return V2Dr, V2Ds

    @staticmethod
    def computeDifferentiationMatrices(V2Dr, V2Ds, V):
        """Dr = (V^T \\ V2Dr^T)^T through dsgesv (:456-491)."""
        Dr = lapack.solve(V.T, V2Dr.T).T
        Ds = lapack.solve(V.T, V2Ds.T).T
        return np.ascontiguousarray(Dr), np.ascontiguousarray(Ds)

    # ------------------------------------------------------------ nodes (:549-640, :678-728)
    def computeWarpFactor(self, r):
        N = self.NOrder
        Np1 = N + 1
        r = np.asarray(r, float)
        req = -1.0 + 2.0 * np.arange(Np1) / (Np1 - 1.0)
        rLGL = jacobi.jacobi_gauss_lobatto(0.0, 0.0, N)
        Veq = jacobi.vandermonde_1d(req)
        P = np.zeros((Np1, len(r)))
        for i in range(Np1):
            P[i, :] = jacobi.jacobi_polynomial(r, 0.0, 0.0, i)
        L = lapack.solve(Veq.T, P)
        warp = L.T @ (rLGL - req)
        zf = (np.abs(r) < 1.0 - 1e-10).astype(float)
        sf = 1.0 - (zf * r) * (zf * r)
        return warp / sf + warp * (zf - 1.0)

    def computeEquilateralNodes(self):
        N = self.NOrder
        alphaOptimal = [0.0000, 0.0000, 1.4152, 0.1001, 0.2751, 0.9800, 1.0999,
                        1.2832, 1.3648, 1.4773, 1.4959, 1.5743, 1.5770, 1.6223, 1.6258]
        alpha = alphaOptimal[N - 1] if N < 16 else 2.0 / 3.0
        Np = (N + 1) * (N + 2) // 2
        L1, L3 = np.zeros(Np), np.zeros(Np)
        c = 0
        for n in range(1, N + 2):
            for m in range(1, N + 3 - n):
                L1[c] = (n - 1.0) / N
                L3[c] = (m - 1.0) / N
                c += 1
        L2 = 1.0 - L1 - L3
        x = -L2 + L3
        y = (-L2 - L3 + 2 * L1) / np.sqrt(3.0)
        blend1, blend2, blend3 = 4 * L2 * L3, 4 * L1 * L3, 4 * L1 * L2
        warpf1 = self.computeWarpFactor(L3 - L2)
        warpf2 = self.computeWarpFactor(L1 - L3)
        warpf3 = self.computeWarpFactor(L2 - L1)
        a2 = alpha * alpha
        warp1 = blend1 * warpf1 * (1 + a2 * L1 * L1)
        warp2 = blend2 * warpf2 * (1 + a2 * L2 * L2)
        warp3 = blend3 * warpf3 * (1 + a2 * L3 * L3)
        x = x + 1 * warp1 + np.cos(2 * _pi / 3) * warp2 + np.cos(4 * _pi / 3) * warp3
        y = y + 0 * warp1 + np.sin(2 * _pi / 3) * warp2 + np.sin(4 * _pi / 3) * warp3
        return x, y

    def buildNodes(self):
        x, y = self.computeEquilateralNodes()
        self.rGrid, self.sGrid = self.xyTors(x, y)
        r, s = self.rGrid, self.sGrid
        Nfp = self.NumFacePoints
        f1 = np.nonzero(np.abs(s + 1) < NodeTol)[0]
        f2 = np.nonzero(np.abs(r + s) < NodeTol)[0]
        f3 = np.nonzero(np.abs(r + 1) < NodeTol)[0]
        assert len(f1) == Nfp and len(f2) == Nfp and len(f3) == Nfp
        self.Fmask = np.stack([f1, f2, f3], axis=1).astype(np.int32)  # [Nfp][3]

    # ------------------------------------------------------------ lift (:1060-1138)
    def buildLift(self):
        Np, Nfp = self.NumLocalPoints, self.NumFacePoints
        r, s, Fm = self.rGrid, self.sGrid, self.Fmask
        E = np.zeros((Np, 3 * Nfp))
        for f, coord in enumerate((r, r, s)):
            V1D = jacobi.vandermonde_1d(coord[Fm[:, f]])
            massEdge = lapack.inverse(V1D @ V1D.T)
            E[Fm[:, f], f * Nfp:(f + 1) * Nfp] = massEdge
        self.V = self.computeVandermondeMatrix(self.NOrder, r, s)
        self.Vinv = lapack.inverse(self.V)
        self.Lift = np.ascontiguousarray((self.V @ self.V.T) @ E)

    # ------------------------------------------------------------ geometry (:730-893)
    def buildOperators(self):
        r, s = self.rGrid, self.sGrid
        self.V = self.computeVandermondeMatrix(self.NOrder, r, s)
        self.Vinv = lapack.inverse(self.V)
        V2Dr, V2Ds = self.computeGradVandermondeMatrix(self.NOrder, r, s)
        self.Dr, self.Ds = self.computeDifferentiationMatrices(V2Dr, V2Ds, self.V)

    def buildPhysicalGrid(self):
        mesh = self.Mesh2D
        K = self.NumElements = mesh.get_NumElements()
        etov = mesh.get_Elements().reshape(K, 3)
        vert = mesh.get_Vertices().reshape(-1, 3)
        VX, VY = vert[:, 0], vert[:, 1]
        va, vb, vc = etov[:, 0], etov[:, 1], etov[:, 2]
        self.buildOperators()
        r, s = self.rGrid[:, None], self.sGrid[:, None]
        self.xGrid = 0.5 * (-(r + s) * VX[va][None, :] + (1 + r) * VX[vb][None, :] + (1 + s) * VX[vc][None, :])
        self.yGrid = 0.5 * (-(r + s) * VY[va][None, :] + (1 + r) * VY[vb][None, :] + (1 + s) * VY[vc][None, :])
        self._metrics_and_normals()

    def _metrics_and_normals(self):
        Dr, Ds, x, y = self.Dr, self.Ds, self.xGrid, self.yGrid
        xr, yr, xs, ys = Dr @ x, Dr @ y, Ds @ x, Ds @ y
        J = xr * ys - xs * yr
        self.J = J
        self.rx, self.ry, self.sx, self.sy = ys / J, -xs / J, -yr / J, xr / J
        Nfp = self.NumFacePoints
        fm = self.Fmask.T.reshape(-1)  # face-major list of local node ids: [f*Nfp + n]
        fxr, fxs, fyr, fys = xr[fm], xs[fm], yr[fm], ys[fm]
        nx, ny = np.empty_like(fxr), np.empty_like(fxr)
        f1, f2, f3 = slice(0, Nfp), slice(Nfp, 2 * Nfp), slice(2 * Nfp, 3 * Nfp)
        nx[f1], ny[f1] = fyr[f1], -fxr[f1]
        nx[f2], ny[f2] = fys[f2] - fyr[f2], -fxs[f2] + fxr[f2]
        nx[f3], ny[f3] = -fys[f3], fxs[f3]
        norm = np.sqrt(nx * nx + ny * ny)
        self.nx, self.ny = nx / norm, ny / norm
        self.Fscale = norm / J[fm]
        self.Fx, self.Fy = x[fm], y[fm]

    # ------------------------------------------------------------ maps (:895-1057)
    def buildMaps(self):
        mesh = self.Mesh2D
        K, Np, Nfp, Nf = self.NumElements, self.NumLocalPoints, self.NumFacePoints, self.NumFaces
        E2E = mesh.get_EToE().reshape(K, Nf).astype(np.int64)
        E2F = mesh.get_EToF().reshape(K, Nf).astype(np.int64)
        etov = mesh.get_Elements().reshape(K, Nf)
        vert = mesh.get_Vertices().reshape(-1, 3)
        fm = self.Fmask.astype(np.int64)  # [Nfp][Nf]
        x, y = self.xGrid, self.yGrid  # [Np][K]
        v1 = etov
        v2 = np.roll(etov, -1, axis=1)
        refd_all = np.hypot(vert[v1, 0] - vert[v2, 0], vert[v1, 1] - vert[v2, 1])  # [K][Nf]
        vmapM = np.empty((K, Nf, Nfp), dtype=np.int32)
        vmapP = np.empty((K, Nf, Nfp), dtype=np.int32)
        mapP = np.empty((K, Nf, Nfp), dtype=np.int32)
        CH = 1 << 17  # elements per chunk: bounds the [K][Nf][Nfp][Nfp] distance tensor
        for c0 in range(0, K, CH):
            c1 = min(K, c0 + CH)
            karr = np.arange(c0, c1, dtype=np.int64)
            e2e, e2f, refd = E2E[c0:c1], E2F[c0:c1], refd_all[c0:c1]
            # vmapM3[k][f][n] = Fmask[n][f] + Np*k
            vM = fm.T[None, :, :] + Np * karr[:, None, None]
            xM = x[fm.T[None, :, :], karr[:, None, None]]  # [k][Nf][Nfp]
            yM = y[fm.T[None, :, :], karr[:, None, None]]
            nodeP = fm.T[e2f]  # local ids on the neighbour's face: [k][Nf][Nfp]
            xP = x[nodeP, e2e[:, :, None]]
            yP = y[nodeP, e2e[:, :, None]]
            # distance matrix [k][Nf][n][nP]; the reference keeps the LAST nP that matches (loop order, :957-968)
            dx = xM[:, :, :, None] - xP[:, :, None, :]
            dy = yM[:, :, :, None] - yP[:, :, None, :]
            match = np.hypot(dx, dy) < (refd * NodeTol)[:, :, None, None]
            has = match.any(axis=3)
            last = Nfp - 1 - np.argmax(match[:, :, :, ::-1], axis=3)
            nP = np.take_along_axis(nodeP, last, axis=2)
            vmapM[c0:c1] = vM
            vmapP[c0:c1] = np.where(has, nP + Np * e2e[:, :, None], 0)
            mapP[c0:c1] = np.where(has, last + e2f[:, :, None] * Nfp + e2e[:, :, None] * Nf * Nfp, 0)
        self.vmapM = vmapM.reshape(-1).astype(np.int32)
        self.vmapP = vmapP.reshape(-1).astype(np.int32)
        self.mapP = mapP.reshape(-1).astype(np.int32)
        self.mapB = np.nonzero(self.vmapP == self.vmapM)[0].astype(np.int32)
        self.vmapB = self.vmapM[self.mapB]
        self.BCmap = {}
        self.buildBCHash(mesh.get_BCType())

    def buildBCHash(self, bcType=None):
        """Appends to the existing hash like the reference (:1028-1057); BCmap[tag] = face-node ids."""
        if bcType is None:
            bcType = self.Mesh2D.get_BCType()
        bc = np.repeat(np.asarray(bcType, dtype=np.int32).reshape(-1), self.NumFacePoints)
        for tag in np.unique(bc):
            if tag == 0:
                continue
            ids = np.nonzero(bc == tag)[0].astype(np.int32)
            prev = self.BCmap.get(int(tag))
            self.BCmap[int(tag)] = ids if prev is None else np.concatenate([prev, ids])

    # ------------------------------------------------------------ filter (:515-547)
    def buildFilter(self, Nc, s):
        N, Np = self.NOrder, self.NumLocalPoints
        alpha = -np.log(np.finfo(np.float64).eps)
        d = np.ones(Np)
        c = 0
        for i in range(N + 1):
            for j in range(N - i + 1):
                if (i + j) >= Nc:
                    k = (float(i + j) - Nc) / (float(N) - Nc)
                    d[c] = np.exp(-alpha * k ** s)
                c += 1
        self.Filter = np.ascontiguousarray(self.V @ (np.diag(d) @ self.Vinv))
        return self.Filter

    # ------------------------------------------------------------ output support
    def computeInterpMatrix(self, rout, sout):
        """IM = Vout * V^-1 (src/TriangleNodesProvisioner.cpp:1140-1152, src/QuadNodesProvisioner.cpp:707-719)."""
        Vout = self.computeVandermondeMatrix(self.NOrder, np.asarray(rout, float), np.asarray(sout, float))
        return np.ascontiguousarray(Vout @ self.Vinv)

    def splitTable(self):
        """Equispaced output nodes and the linear sub-elements of one element (src/TriangleNodesProvisioner.cpp:1158-1204):
        node (n,m) has r = -1 + 2m/N, s = -1 + 2n/N; triangles (v1,v2,v3) and (v2,v4,v3) per lattice cell.
        Returns rout, sout, localE2V [numLocal][3]."""
        N = self.NumFacePoints - 1
        counter = -np.ones((N + 2, N + 2), dtype=np.int64)
        rout, sout = [], []
        for n in range(N + 1):
            for m in range(N + 1 - n):
                counter[n, m] = len(rout)
                rout.append(-1.0 + 2.0 * m / N)
                sout.append(-1.0 + 2.0 * n / N)
        cells = []
        for n in range(N + 1):
            for m in range(N - n):
                v1, v2, v3, v4 = counter[n, m], counter[n, m + 1], counter[n + 1, m], counter[n + 1, m + 1]
                cells.append((v1, v2, v3))
                if v4 >= 0:
                    cells.append((v2, v4, v3))
        return np.array(rout), np.array(sout), np.array(cells, dtype=np.int64)

    def splitElements(self, x, y, field, interpolated=False):
        """splitElements of the reference (src/TriangleNodesProvisioner.cpp:1154-1264, src/QuadNodesProvisioner.cpp:721-845):
        x, y, field [Np][K] -> xnew, ynew, fieldnew [3 or 4][numLocal*K], every linear sub-element with its own corner
        values.  interpolated=True: `field` already holds IM*field (the device output path), only x, y are interpolated."""
        rout, sout, cells = self.splitTable()
        IM = self.computeInterpMatrix(rout, sout)
        K = field.shape[1]
        fi = field if interpolated else IM @ field
        idx = cells.T  # [corners][numLocal]

        def corners(a):  # -> [corners][numLocal*K], element-major like E2Vnew
            return np.ascontiguousarray(a[idx].transpose(0, 2, 1).reshape(idx.shape[0], -1))

        return corners(IM @ x), corners(IM @ y), corners(fi)

    # ------------------------------------------------------------ getters
    def get_NumLocalPoints(self):
        return self.NumLocalPoints

    def get_NumFacePoints(self):
        return self.NumFacePoints

    def get_NumElements(self):
        return self.NumElements

    def get_Dr(self):
        return self.Dr

    def get_Ds(self):
        return self.Ds

    def get_Lift(self):
        return self.Lift

    def get_Filter(self):
        return self.Filter

    def get_xGrid(self):
        return self.xGrid

    def get_yGrid(self):
        return self.yGrid

    def get_rx(self):
        return self.rx

    def get_ry(self):
        return self.ry

    def get_sx(self):
        return self.sx

    def get_sy(self):
        return self.sy

    def get_nx(self):
        return self.nx

    def get_ny(self):
        return self.ny

    def get_Fscale(self):
        return self.Fscale

    def get_vmapM(self):
        return self.vmapM

    def get_vmapP(self):
        return self.vmapP

    def get_vmapB(self):
        return self.vmapB

    def get_mapB(self):
        return self.mapB

    def get_bcMap(self):
        return self.BCmap

    def get_DGContext(self):
        """(:1364-1396) -- here arrays are owned (contiguous copies), not raw pointers."""
        c = np.ascontiguousarray
        return DGContext2D(
            N=self.NOrder, Np=self.NumLocalPoints, Nfp=self.NumFacePoints, Nfaces=self.NumFaces, K=self.NumElements,
            Dr=c(self.Dr), Ds=c(self.Ds), Lift=c(self.Lift),
            rx=c(self.rx), ry=c(self.ry), sx=c(self.sx), sy=c(self.sy),
            nx=c(self.nx), ny=c(self.ny), Fscale=c(self.Fscale),
            vmapM=c(self.vmapM), vmapP=c(self.vmapP), BCmap=dict(self.BCmap),
            J=c(self.J), x=c(self.xGrid), y=c(self.yGrid), r=self.rGrid, s=self.sGrid,
            Fmask=self.Fmask, V=self.V, Vinv=self.Vinv, Filt=self.Filter)
