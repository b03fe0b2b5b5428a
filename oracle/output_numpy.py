"""TEST INFRASTRUCTURE -- loop-level restatement of the reference's output path (small cases only).

Follows src/TriangleNodesProvisioner.cpp:1154-1264 / src/QuadNodesProvisioner.cpp:721-845 (splitElements),
include/VtkOutputter.hpp:111-141 (insertAllCells) and src/sw2d/main.cpp:199-210 (the derived fields), one loop per
reference loop.  Used by tests/ only, as the checker of blitzdg_b200/outputters.py and of the device output kernel.
"""
import numpy as np


def split_elements(Np, N, quad, IM, x, y, field):
    K = field.shape[1]
    counter = -np.ones((N + 2, N + 2), dtype=int)
    count = 0
    for n in range(N + 1):
        for m in range(N + 1 if quad else N + 2 - (n + 1)):
            counter[n, m] = count
            count += 1
    localE2V = []
    for n in range(N if quad else N + 1):
        for m in range(N if quad else N + 1 - (n + 1)):
            v1, v2, v3, v4 = counter[n, m], counter[n, m + 1], counter[n + 1, m], counter[n + 1, m + 1]
            if quad:
                localE2V.append([v1, v2, v3, v4])
            else:
                localE2V.append([v1, v2, v3])
                if v4 >= 0:
                    localE2V.append([v2, v4, v3])
    E2Vnew = []
    for k in range(K):
        shift = k * Np
        for row in localE2V:
            E2Vnew.append([v + shift for v in row])

    def contract(A, F):  # sum(IM(ii,kk)*F(kk,jj),kk): sequential, no FMA
        out = np.zeros((A.shape[0], F.shape[1]))
        for i in range(A.shape[0]):
            for j in range(F.shape[1]):
                acc = 0.0
                for k in range(A.shape[1]):
                    acc = acc + A[i, k] * F[k, j]
                out[i, j] = acc
        return out

    rx, ry, rf = contract(IM, x), contract(IM, y), contract(IM, field)
    xV, yV, fV = rx.T.reshape(-1), ry.T.reshape(-1), rf.T.reshape(-1)  # fullToVector(.., false)
    nc, tot = len(localE2V[0]), len(E2Vnew)
    xnew, ynew, fnew = np.zeros((nc, tot)), np.zeros((nc, tot)), np.zeros((nc, tot))
    for i in range(tot):
        for c in range(nc):
            xnew[c, i] = xV[E2Vnew[i][c]]
            ynew[c, i] = yV[E2Vnew[i][c]]
            fnew[c, i] = fV[E2Vnew[i][c]]
    return xnew, ynew, fnew


def insert_all_cells(x, y, field):
    """points [Np*K][3], point scalars, cells (node ids) exactly as insertAllCells builds them."""
    Np, K = field.shape
    pts, vals, cells = [], [], []
    nodeId = 0
    for k in range(K):
        nodes = []
        for n in range(Np):
            pts.append((x[n, k], y[n, k], 0.0))
            vals.append(field[n, k])
            nodes.append(nodeId)
            nodeId += 1
        if Np == 4:
            nodes[1], nodes[3] = nodes[3], nodes[1]
            nodes[1], nodes[2] = nodes[2], nodes[1]
        cells.append(nodes)
    return np.array(pts), np.array(vals), np.array(cells)


def derived_fields(h, hu, hv, H):
    return {"eta": h - H, "u": hu / h, "v": hv / h}


def positivity_limiter(h, hu, hv):
    """Restatement of swhelpers/limiters.py:3-21, element by element: theta from the element's mean and minimum depth
    (minimum floored at 1e-3, theta capped at 1), then every field contracted towards its element mean."""
    Np, K = h.shape
    out = [np.empty_like(h), np.empty_like(h), np.empty_like(h)]
    for k in range(K):
        hmin = max(h[:, k].min(), 1e-3)
        hmean = h[:, k].mean()
        theta = min(hmean / (hmean - hmin + 1e-4), 1.0)
        for o, f in zip(out, (h, hu, hv)):
            mean = f[:, k].mean()
            o[:, k] = mean + theta * (f[:, k] - mean)
    return out
