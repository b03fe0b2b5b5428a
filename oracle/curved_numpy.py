"""TEST INFRASTRUCTURE -- CPU restatement of the reference's over-integrated (curved-element) shallow-water RHS,
sw2dComputeRHS_curved (reference swhelpers/rhs.py:6-176, fluxes swhelpers/flux.py:1-23).

Pinned: tests/test_curved_rhs.py holds it to outputs of the reference's own function (imported unmodified from
/root/reference by tests/golden/make_curved_reference_vectors.py -> tests/golden/curved_reference_rhs.npz).

Only tests/ may import this.  Element by element and face by face, where the reference works on whole arrays -- the
same values in a different loop structure, so that a transcription slip in either would show:
  * volume part (rhs.py:8-47): the four conserved fields are interpolated to the cubature points, the fluxes are
    evaluated THERE, and tested against the gradients of the nodal basis:  MM*RHS_i = Dr_c^T (W (rx F_i + ry G_i)) +
    Ds_c^T (W (sx F_i + sy G_i));
  * surface part (rhs.py:49-139): traces at the Gauss nodes of the faces (Interp), wall reflection of the '+' momentum
    (the '+' velocities of the wave speed are taken BEFORE the reflection, :80-87), local Lax-Friedrichs flux with the
    maximum speed over the Gauss nodes of a face (:98-104), weighted and tested:  MM*RHS_i -= Interp^T (W_g flux_i);
  * mass matrix (rhs.py:141-164): straight elements  RHS = (V V^T)(MMRHS / J),  curved elements two triangular solves
    with the element's upper Cholesky factor;
  * sources at the nodes (rhs.py:166-176), as written: RHS2 += f hv - CD |u| u,  RHS3 -= f hu - CD |u| v  (the drag
    enters RHS3 with a plus sign in the reference; kept),  RHS2 -= g h zx,  RHS3 -= g h zy.
"""
import numpy as np


def _fluxes(h, hu, hv, hN, g):
    u, v = hu / h, hv / h
    half = 0.5 * g * h * h
    return (hu, hu * u + half, hv * u, hN * u), (hv, hu * v, hv * v + half, hN * v)


def rhs_curved(h, hu, hv, hN, zx, zy, g, f, CD, ctx, cub, gauss, curvedEls, J, gmapM, gmapP):
    """Returns (RHS1..RHS4), each [Np][K].  ctx: DGContext2D (V), cub / gauss: blitzdg_b200.curved contexts."""
    Np, K = h.shape
    NG3 = gauss.Interp.shape[0]
    Ng = NG3 // 3
    q = (h, hu, hv, hN)
    # Gauss-node values of every element first: the '+' side of a face node lives in the neighbour
    gq = [(gauss.Interp @ a).flatten("F") for a in q]
    wall = np.zeros(NG3 * K, dtype=bool)
    if 3 in gauss.BCmap:
        wall[np.asarray(gauss.BCmap[3], dtype=np.int64)] = True
    mminv = ctx.V @ ctx.V.T
    curved = set(int(k) for k in curvedEls)
    CDa = np.broadcast_to(np.asarray(CD, dtype=np.float64), h.shape)
    out = [np.empty((Np, K)) for _ in range(4)]
    for k in range(K):
        cq = [cub.V @ a[:, k] for a in q]
        F, G = _fluxes(cq[0], cq[1], cq[2], cq[3], g)
        Wc = cub.W[:, k]
        mm = [cub.Dr.T @ (Wc * (cub.rx[:, k] * F[i] + cub.ry[:, k] * G[i])) + cub.Ds.T @ (Wc * (cub.sx[:, k] * F[i] + cub.sy[:, k] * G[i]))
              for i in range(4)]
        flux = [np.empty(NG3) for _ in range(4)]
        for face in range(3):
            ids = np.arange(face * Ng, (face + 1) * Ng) + NG3 * k
            iM, iP = gmapM[ids], gmapP[ids]
            M = [gq[i][iM] for i in range(4)]
            P = [gq[i][iP].copy() for i in range(4)]
            uM, vM, uP, vP = M[1] / M[0], M[2] / M[0], P[1] / P[0], P[2] / P[0]
            nxg, nyg = gauss.nx[face * Ng:(face + 1) * Ng, k], gauss.ny[face * Ng:(face + 1) * Ng, k]
            wl = wall[ids]
            dn = M[1] * nxg + M[2] * nyg
            P[1] = np.where(wl, M[1] - 2 * nxg * dn, P[1])
            P[2] = np.where(wl, M[2] - 2 * nyg * dn, P[2])
            FM, GM = _fluxes(M[0], M[1], M[2], M[3], g)
            FP, GP = _fluxes(P[0], P[1], P[2], P[3], g)
            lam = max(np.max(np.sqrt(uM * uM + vM * vM) + np.sqrt(g * M[0])), np.max(np.sqrt(uP * uP + vP * vP) + np.sqrt(g * P[0])))
            for i in range(4):
                flux[i][face * Ng:(face + 1) * Ng] = 0.5 * ((FM[i] + FP[i]) * nxg + (GM[i] + GP[i]) * nyg + lam * (M[i] - P[i]))
        for i in range(4):
            b = mm[i] - gauss.Interp.T @ (gauss.W[:, k] * flux[i])
            if k in curved:
                U = cub.MMChol[:, :, k]
                out[i][:, k] = np.linalg.solve(U, np.linalg.solve(U.T, b))
            else:
                out[i][:, k] = mminv @ (b / J[:, k])
    u, v = hu / h, hv / h
    nrm = np.hypot(u, v)
    out[1] += f * hv - CDa * nrm * u
    out[2] -= f * hu - CDa * nrm * v
    out[1] -= g * h * zx
    out[2] -= g * h * zy
    return tuple(out)
