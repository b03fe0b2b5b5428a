"""Debug: two METIS ranks emulated on ONE GPU (ghosts moved through the pack/set_ghosts hooks): where do the fused and the
two-launch kernels disagree?"""
import ctypes as C
import sys

import numpy as np

sys.path.insert(0, ".")
from blitzdg_b200 import capi
from blitzdg_b200.initial_conditions import gaussian_hump
from blitzdg_b200.partition import build_local, make_partition
from blitzdg_b200.sw2d import SW2dSolver
from blitzdg_b200.synthetic import provision, structured_mesh

nx, ny, P = (int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (96, 80, 2)
mesh = structured_mesh("tri", nx, ny)
ctx = provision(mesh, 4).get_DGContext()
H, h, hu, hv = gaussian_hump(ctx.x, ctx.y)
rng = np.random.default_rng(21)
hu = 0.2 * rng.standard_normal(h.shape)
hv = 0.2 * rng.standard_normal(h.shape)
ep, _ = make_partition(mesh, P, "metis", seed=5)
parts = [build_local(mesh, ep, r, 4) for r in range(P)]
sols = []
for c, halo, info in parts:
    s = SW2dSolver(c, halo=halo)
    el = info["elements"]
    s.set_state(h[:, el], hu[:, el], hv[:, el])
    sols.append(s)
# ghosts through the host: rank r's ghost g has key (owner, gelem, node); read straight from the global state
for s, (c, halo, info) in zip(sols, parts):
    gk = info["ghost_keys"]
    Np, K = c.Np, ctx.K
    gn, ge = gk % Np, (gk // Np) % K
    # halo_set_ghosts wants the peers' send-buffer layout: per neighbour segment, 3 planes of cnt values
    ro = halo["recv_offset"]
    buf = np.zeros(3 * len(gk))
    for i in range(len(halo["nbr_rank"])):
        a, b = ro[i], ro[i + 1]
        cnt = b - a
        for f, arr in enumerate((h, hu, hv)):
            buf[3 * a + f * cnt: 3 * a + (f + 1) * cnt] = arr[gn[a:b], ge[a:b]]
    s.halo_set_ghosts(buf)
lib = capi.load()
for r, (s, (c, halo, info)) in enumerate(zip(sols, parts)):
    R = {}
    for var in (1, 2):
        s.set_variant(var)
        # set_variant invalidates nothing on a comm-less handle; ghosts stay
        R[var] = s.computeRHS()
    d = max(np.abs(R[1][f] - R[2][f]).max() for f in range(3))
    scale = max(np.abs(R[1][f]).max() for f in range(3))
    print("rank", r, "K", c.K, "n_interior", halo["n_interior"], "max|v1-v2|", d, "scale", scale)
    bad = np.nonzero(np.max([np.abs(R[1][f] - R[2][f]).max(axis=0) for f in range(3)], axis=0) > 1e-9 * scale)[0]
    print("  differing elements:", len(bad), bad[:40])
    if len(bad):
        K, Np, Nfp = c.K, c.Np, c.Nfp
        vP = np.ascontiguousarray(c.vmapP, dtype=np.int32)
        perm = np.zeros(K, np.int32)
        lib.bdg_element_order(vP.ctypes.data_as(C.POINTER(C.c_int32)), Np, Nfp, 3, K, halo["n_interior"], perm.ctypes.data_as(C.POINTER(C.c_int32)))
        inv = np.empty(K, np.int64)
        inv[perm] = np.arange(K)
        print("  device positions:", sorted(inv[bad].tolist())[:40], " tiles:", sorted(set((inv[bad] // 32).tolist()))[:20], "of", (K + 31) // 32)
        print("  cut elements among them:", int((bad >= halo["n_interior"]).sum()))
    s.close()
