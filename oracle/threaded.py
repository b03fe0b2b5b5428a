"""TEST INFRASTRUCTURE -- the CPU oracle over a WHOLE mesh, domain-decomposed over host threads.

The reference is single-threaded; to give it "all the host threads it can use" on the arm's own
mesh (bench.py --impl reference) and to evaluate the oracle over the BASELINE-size meshes with bounded memory
(tests/test_fullsize_parity.py), the mesh is cut into strips (blitzdg_b200.partition.build_local), each strip owns an
Oracle instance on its local DGContext2D, and the '+' traces on strip boundaries are gathered from the global state
before every RHS evaluation -- the CPU twin of the multi-GPU halo exchange.  The arithmetic of an element does not
depend on the decomposition (tests/test_halo_gloo.py).  Simple mode (per-face LF) only: the global-LF maximum would
need a reduction across strips inside oracle_sw2d_rhs.

Only tests/ and bench.py's cpu_baseline / --impl reference legs import this.
"""
import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np

from blitzdg_b200.lserk4 import rk4a, rk4b
from blitzdg_b200.partition import build_local, strip_partition

from .binding import Oracle


class StripOracle:
    def __init__(self, mesh, N, nparts=None, threads=None, max_part_elements=250_000, parts=None, **oracle_params):
        K, Nf = mesh.get_NumElements(), mesh.get_NumFaces()
        self.K, self.N = K, N
        self.threads = int(threads or os.cpu_count() or 1)
        if nparts is None:
            nparts = max(self.threads, -(-K // max_part_elements))
        nparts = max(1, min(int(nparts), K // 64 if K >= 128 else 1))
        self.nparts = nparts
        cy = mesh.get_Vertices().reshape(-1, 3)[mesh.get_Elements().reshape(K, Nf), 1].mean(axis=1)
        self.epart = strip_partition(cy, nparts)
        self.which = list(range(nparts)) if parts is None else list(parts)
        self.pool = ThreadPoolExecutor(self.threads)

        def make(r):
            lctx, halo, info = build_local(mesh, self.epart, r, N)
            gk = info["ghost_keys"]
            Np = lctx.Np
            return dict(ctx=lctx, orc=Oracle(lctx, **oracle_params), el=info["elements"], gnode=gk % Np, gelem=(gk // Np) % K)

        self.parts = list(self.pool.map(make, self.which)) if nparts > 1 else None
        if nparts == 1:
            from blitzdg_b200.synthetic import provision

            ctx = provision(mesh, N).get_DGContext()
            self.parts = [dict(ctx=ctx, orc=Oracle(ctx, **oracle_params), el=np.arange(K), gnode=np.zeros(0, np.int64), gelem=np.zeros(0, np.int64))]
        self.Np = self.parts[0]["ctx"].Np
        self.elements = int(sum(len(p["el"]) for p in self.parts))
        self.stage = 0

    def close(self):
        self.pool.shutdown()

    def set_state(self, h, hu, hv):
        self.q = [np.array(a, dtype=np.float64, order="C") for a in (h, hu, hv)]
        for p in self.parts:
            p["q"] = [np.ascontiguousarray(a[:, p["el"]]) for a in self.q]
            p["res"] = [np.zeros_like(a) for a in p["q"]]
        self.stage = 0

    def _ghost(self, p):
        g = np.zeros((4, len(p["gnode"])))
        for c in range(3):
            g[c] = self.q[c][p["gnode"], p["gelem"]]
        return g

    def rhs(self):
        """One RHS of the current state: [3] arrays [Np][K], NaN in strips that were not selected."""
        R = [np.full(self.q[0].shape, np.nan) for _ in range(3)]

        def work(p):
            Rl = p["orc"].rhs(p["q"][0], p["q"][1], p["q"][2], ghost=self._ghost(p))
            for c in range(3):
                R[c][:, p["el"]] = Rl[c]

        list(self.pool.map(work, self.parts))
        return R

    def stages(self, n, dt):
        """n RHS+stage evaluations continuing the LSERK4 stage cycle (include/LSERK4.hpp:16-27, src/advec1d/main.cpp:92-102)."""
        assert len(self.which) == self.nparts, "stepping needs every strip"
        for _ in range(n):
            s = self.stage

            def work(p):
                Rl = p["orc"].rhs(p["q"][0], p["q"][1], p["q"][2], ghost=self._ghost(p))
                for c in range(3):
                    Oracle.stage(s, dt, Rl[c], p["res"][c], p["q"][c])

            list(self.pool.map(work, self.parts))  # every strip has read its ghosts before any strip publishes

            def publish(p):
                for c in range(3):
                    self.q[c][:, p["el"]] = p["q"][c]

            list(self.pool.map(publish, self.parts))
            self.stage = (s + 1) % 5

    def step(self, nsteps, dt):
        self.stages(5 * nsteps, dt)

    def get_state(self):
        return self.q


def lserk4_coefficients():
    return np.asarray(rk4a), np.asarray(rk4b)
