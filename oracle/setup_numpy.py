"""TEST INFRASTRUCTURE -- restatement of the reference's connectivity / '+' map construction (small meshes only).

build_connectivity follows src/MeshManager.cpp:383-481: the vertex-to-face incidence matrix VToF, FToF = VToF^T VToF
(scipy.sparse stands in for CXSparse cs_multiply/cs_transpose: integer-valued products), entries equal to 2 off the
diagonal are shared faces.  build_vmapP follows src/TriangleNodesProvisioner.cpp:895-1020: for every connected face
pair the all-pairs distance table of the face nodes, matched where the distance is below NodeTol.
Used by tests/ to check the sort-based O(n log n) host mirror (blitzdg_b200/mesh_manager.py, triangle_nodes.py).
"""
import numpy as np
from scipy import sparse

NodeTol = 1e-5


def build_connectivity(EToV, K, Nf, Nv):
    E2V = np.asarray(EToV).reshape(K, Nf)
    rows, cols = [], []
    for k in range(K):
        for f in range(Nf):
            g = k * Nf + f
            rows += [E2V[k, f], E2V[k, (f + 1) % Nf]]
            cols += [g, g]
    VToF = sparse.csc_matrix((np.ones(len(rows)), (rows, cols)), shape=(Nv, K * Nf))
    FToF = (VToF.T @ VToF).tocsc()
    E2E = np.repeat(np.arange(K), Nf).reshape(K, Nf).copy()
    E2F = np.tile(np.arange(Nf), K).reshape(K, Nf).copy()
    for j in range(K * Nf):
        for p in range(FToF.indptr[j], FToF.indptr[j + 1]):
            i = FToF.indices[p]
            if i != j and abs(FToF.data[p] - 2.0) < NodeTol:
                E2E[i // Nf, i % Nf] = j // Nf
                E2F[i // Nf, i % Nf] = j % Nf
    return E2E.reshape(-1), E2F.reshape(-1)


def build_vmapP(x, y, Fmask, E2E, E2F, Np, Nfp, Nf, K):
    """x, y [Np][K]; Fmask [Nfp][Nf].  Returns vmapM, vmapP (length Nf*Nfp*K, element-major, face, face node)."""
    xv, yv = x.T.reshape(-1), y.T.reshape(-1)
    E2E, E2F = E2E.reshape(K, Nf), E2F.reshape(K, Nf)
    vmapM = np.zeros((K, Nf, Nfp), dtype=np.int64)
    for k in range(K):
        for f in range(Nf):
            for n in range(Nfp):
                vmapM[k, f, n] = Fmask[n, f] + k * Np
    vmapP = vmapM.copy()
    for k1 in range(K):
        for f1 in range(Nf):
            k2, f2 = E2E[k1, f1], E2F[k1, f1]
            vM, vP = vmapM[k1, f1], vmapM[k2, f2]
            for i in range(Nfp):
                for j in range(Nfp):
                    d = (xv[vM[i]] - xv[vP[j]]) ** 2 + (yv[vM[i]] - yv[vP[j]]) ** 2
                    if np.sqrt(d) < NodeTol:
                        vmapP[k1, f1, i] = vP[j]
    return vmapM.reshape(-1), vmapP.reshape(-1)
