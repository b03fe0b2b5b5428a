"""TEST INFRASTRUCTURE -- independent NumPy transcription of the sw2d RHS (second derivation).

Written from the reference's Python drivers (/root/reference/sw2dquads.py:25-123, sw2d.py:12-146:
Fortran-order flattening, fancy-index gathers, np.dot contractions) plus, for the full mode, the
star-state / source lines of /root/reference/src/sw2d/main.cpp:355-368,460-483.  It exists to catch a
mis-transcription in oracle/sw2d_oracle.c: the two must agree to ~1e-14 (SURVEY.md §8c self-check ii).
Only tests/ may import this module.
"""
import numpy as np

PI = 3.1415926535897932


def fluxes(h, hu, hv, g):
    F2 = (hu * hu) / h + 0.5 * g * h * h
    G2 = (hu * hv) / h
    G3 = (hv * hv) / h + 0.5 * g * h * h
    return (hu, F2, G2), (hv, G2, G3)


def compute_rhs(ctx, h, hu, hv, g=9.81, lf_mode=0, sources=False, H=None, Hx=None, Hy=None,
                CD=2.5e-3, f=1.0070e-4, t=0.0, tide_amp=3.0, tide_period=3600 * 12.42, tide_ramp=0.15 / 3600,
                use_filter=False, hN=None):
    """hN: optional tracer of the Python drivers' 4th equation (swhelpers/flux.py:18-20, swhelpers/rhs.py:206-296):
    F4 = hN*u, G4 = hN*v, same LF speed; a fourth array is then returned."""
    Nfp, Nf, K = ctx.Nfp, ctx.Nfaces, ctx.K
    vM, vP = ctx.vmapM, ctx.vmapP
    flat = lambda a: a.flatten("F")
    nx, ny = flat(ctx.nx), flat(ctx.ny)
    hM, hP = flat(h)[vM], flat(h)[vP]
    huM, huP = flat(hu)[vM], flat(hu)[vP]
    hvM, hvP = flat(hv)[vM], flat(hv)[vP]
    mapW = ctx.BCmap.get(3, np.zeros(0, dtype=np.int32))
    if len(mapW):
        dn = huM[mapW] * nx[mapW] + hvM[mapW] * ny[mapW]
        if sources:
            hP[mapW] = hM[mapW]
        huP[mapW] = huM[mapW] - 2 * nx[mapW] * dn
        hvP[mapW] = hvM[mapW] - 2 * ny[mapW] * dn
    if sources:
        HM, HP = flat(H)[vM], flat(H)[vP]
        mapO = ctx.BCmap.get(2, np.zeros(0, dtype=np.int32))
        if len(mapO):
            om = 2.0 * PI / tide_period
            huP[mapO] = huM[mapO]
            hvP[mapO] = hvM[mapO]
            hP[mapO] = HM[mapO] + tide_amp * np.cos(om * t) * 0.5 * (np.tanh(tide_ramp * (t - tide_period)) + 1)
        bM, bP = -HM, -HP
        bmax = np.maximum(bP, bM)
        hMs = np.maximum(0.0, hM + bM - bmax)
        hPs = np.maximum(0.0, hP + bP - bmax)
        huM, hvM = hMs * (huM / hMs), hMs * (hvM / hMs)
        huP, hvP = hPs * (huP / hPs), hPs * (hvP / hPs)
        hM, hP = hMs, hPs
    dh, dhu, dhv = hM - hP, huM - huP, hvM - hvP
    (F1M, F2M, F3M), (G1M, G2M, G3M) = fluxes(hM, huM, hvM, g)
    (F1P, F2P, F3P), (G1P, G2P, G3P) = fluxes(hP, huP, hvP, g)
    (F1, F2, F3), (G1, G2, G3) = fluxes(h, hu, hv, g)
    spdM = np.sqrt((huM / hM) ** 2 + (hvM / hM) ** 2) + np.sqrt(g * hM)
    spdP = np.sqrt((huP / hP) ** 2 + (hvP / hP) ** 2) + np.sqrt(g * hP)
    spd = np.maximum(spdM, spdP)
    if lf_mode == 1:
        lam = np.full_like(spd, spd.max())
    else:
        lam = np.repeat(spd.reshape(Nf * K, Nfp).max(axis=1), Nfp)
    dF = [0.5 * ((F1M - F1P) * nx + (G1M - G1P) * ny - lam * dh),
          0.5 * ((F2M - F2P) * nx + (G2M - G2P) * ny - lam * dhu),
          0.5 * ((F3M - F3P) * nx + (G3M - G3P) * ny - lam * dhv)]
    Fs, Gs = [F1, F2, F3], [G1, G2, G3]
    if hN is not None:
        hNM, hNP = flat(hN)[vM], flat(hN)[vP]
        F4M, G4M = hNM * (huM / hM), hNM * (hvM / hM)
        F4P, G4P = hNP * (huP / hP), hNP * (hvP / hP)
        dF.append(0.5 * ((F4M - F4P) * nx + (G4M - G4P) * ny - lam * (hNM - hNP)))
        Fs.append(hN * (hu / h))
        Gs.append(hN * (hv / h))
    out = []
    for Fc, Gc, dFc in zip(Fs, Gs, dF):
        R = -(ctx.rx * (ctx.Dr @ Fc) + ctx.sx * (ctx.Ds @ Fc))
        R += -(ctx.ry * (ctx.Dr @ Gc) + ctx.sy * (ctx.Ds @ Gc))
        R += ctx.Lift @ (ctx.Fscale * dFc.reshape((Nfp * Nf, K), order="F"))
        out.append(R)
    if sources:
        u, v = hu / h, hv / h
        nu = np.sqrt(u * u + v * v)
        out[1] += g * h * Hx - CD * u * nu + f * hv
        out[2] += g * h * Hy - CD * v * nu - f * hu
    if use_filter:
        out = [ctx.Filt @ R for R in out]
    return out
