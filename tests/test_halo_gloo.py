"""world_size-2 (and 4) CPU test of the multi-rank path's host logic over torch.distributed/gloo.

Each rank provisions its element partition (blitzdg_b200.partition.build_local), packs the face traces its
neighbours need in send-list order, exchanges them with isend/irecv exactly as the NCCL path does on the GPU
(capi.cu halo_begin), and evaluates the RHS of its elements with the CPU oracle reading ghost slots.  The
result must equal the single-rank oracle on the same global state: per-face mode is bitwise independent of P.
"""
import os
import socket

import numpy as np
import pytest

from blitzdg_b200.initial_conditions import gaussian_hump
from blitzdg_b200.partition import build_local, make_partition
from blitzdg_b200.synthetic import provision, structured_mesh


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _global_problem(shape, N, how="strips"):
    nx, ny = (24, 16) if how == "metis" else (12, 8)
    mesh = structured_mesh(shape, nx, ny, jitter=0.15 if shape == "quad" else 0.0)
    ctx = provision(mesh, N).get_DGContext()
    H, h, hu, hv = gaussian_hump(ctx.x, ctx.y)
    rng = np.random.default_rng(5)
    hu = 0.3 * rng.standard_normal(h.shape)
    hv = 0.3 * rng.standard_normal(h.shape)
    return mesh, ctx, h, hu, hv


def _worker(rank, world, port, shape, N, how, out):
    import torch
    import torch.distributed as dist

    from oracle.binding import Oracle

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mesh, ctx, h, hu, hv = _global_problem(shape, N, how)
        epart, _ = make_partition(mesh, world, how)
        lctx, halo, info = build_local(mesh, epart, rank, N)
        el = info["elements"]
        q = [np.ascontiguousarray(a[:, el]) for a in (h, hu, hv)]
        flat = [a.T.reshape(-1) for a in q]  # element-major m = n + Np*k
        ghost = np.zeros((4, halo["n_ghost"]))
        reqs, bufs = [], []
        for i, nb in enumerate(halo["nbr_rank"]):
            s0, s1 = halo["send_offset"][i], halo["send_offset"][i + 1]
            r0, r1 = halo["recv_offset"][i], halo["recv_offset"][i + 1]
            send = torch.from_numpy(np.stack([f[halo["send_nodes"][s0:s1]] for f in flat]))
            recv = torch.empty((3, r1 - r0), dtype=torch.float64)
            reqs.append(dist.isend(send, int(nb)))
            reqs.append(dist.irecv(recv, int(nb)))
            bufs.append((r0, r1, recv, send))
        for r in reqs:
            r.wait()
        for r0, r1, recv, _ in bufs:
            ghost[:3, r0:r1] = recv.numpy()
        R = Oracle(lctx).rhs(q[0], q[1], q[2], ghost=ghost)
        Rg = Oracle(ctx).rhs(h, hu, hv)
        err = max(np.abs(R[c] - Rg[c][:, el]).max() for c in range(3))
        scale = max(np.abs(Rg[c]).max() for c in range(3))
        t = torch.tensor([err / scale, float(halo["n_interior"]), float(len(el))], dtype=torch.float64)
        gathered = [torch.zeros(3, dtype=torch.float64) for _ in range(world)]
        dist.all_gather(gathered, t)
        if rank == 0:
            out.put([g.tolist() for g in gathered])
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("shape,N,world,how", [("tri", 4, 2, "strips"), ("tri", 3, 4, "rcb"), ("quad", 3, 2, "strips"),
                                                ("tri", 4, 4, "metis"), ("quad", 3, 3, "metis")])
def test_partitioned_rhs_equals_single_rank_over_gloo(shape, N, world, how):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, shape, N, how, out)) for r in range(world)]
    for p in procs:
        p.start()
    res = out.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    K = sum(r[2] for r in res)
    assert K == (24 * 16 if how == "metis" else 12 * 8) * (2 if shape == "tri" else 1)
    for err, n_int, kloc in res:
        assert err < 1e-14, res          # same arithmetic per element: independent of the partition
        assert 0 < n_int < kloc          # interior-first renumbering leaves work to overlap with the exchange
