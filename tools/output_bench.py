#!/usr/bin/env python
"""Output path on the bench workload (1 M triangles, N=4): cost of a snapshot every outputInterval (20) steps.
  a) device-fed: bdg_sw2d_output_begin (derive + interpolate on device, async copies to pinned memory)
  b) host path as the reference does it: download h, hu, hv, derive eta/u/v and apply IM on the host (numpy)
Prints one JSON line; run on the GPU box: python tools/output_bench.py"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from blitzdg_b200.initial_conditions import gaussian_hump  # noqa: E402
from blitzdg_b200.outputters import DeviceOutput  # noqa: E402
from blitzdg_b200.sw2d import SW2dSolver  # noqa: E402
from blitzdg_b200.synthetic import provision, structured_mesh  # noqa: E402

nx, ny, N = (int(a) for a in (sys.argv[1:4] + ["1000", "500", "4"][len(sys.argv) - 1:]))
P = provision(structured_mesh("tri", nx, ny), N)
ctx = P.get_DGContext()
H, h, hu, hv = gaussian_hump(ctx.x, ctx.y)
sol = SW2dSolver(ctx)
sol.set_bathymetry(H)
sol.set_state(h, hu, hv)
dt = 0.5 * sol.computeTimeStep() * (0.65 / sol.num.CFL)
dev = DeviceOutput(sol, P)
stream = torch.cuda.ExternalStream(sol.stream)
INTERVAL, REPS = 20, 10


def loop(snapshot):
    sol.sync()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(REPS):
        sol.step(INTERVAL, dt, sync=False)
        snapshot(i)
    sol.sync()
    dev.sol.output_wait()
    return (time.perf_counter() - t0) / REPS * 1e3


def snap_device(i):
    if i:
        dev.fields()  # previous snapshot has arrived (a writer thread would consume it here)
    dev.begin(i)


def snap_host(i):
    a, b, c, _ = sol.get_state()
    eta, u, v = a - H, b / a, c / a
    return dev.IM @ eta, dev.IM @ u, dev.IM @ v


for _ in range(2):
    loop(lambda i: None)
base = loop(lambda i: None)
devms = loop(snap_device)
hostms = loop(snap_host)
# the interpolation kernel alone
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
sol.sync()
bufs = [t.numpy() for t in dev._pin[0]]
t0 = time.perf_counter()
sol.output_begin(*bufs)
sol.output_wait()
one = (time.perf_counter() - t0) * 1e3
nbytes = 3 * dev.n_out * ctx.K * 8
print(json.dumps({"workload": "%d triangles N=%d, snapshot of eta,u,v every %d LSERK4 steps" % (ctx.K, N, INTERVAL),
                  "ms_per_interval_no_output": base, "ms_per_interval_device_output": devms, "ms_per_interval_host_output": hostms,
                  "device_output_overhead_pct": 100 * (devms / base - 1), "host_output_overhead_pct": 100 * (hostms / base - 1),
                  "single_snapshot_begin_to_arrival_ms": one, "snapshot_bytes_to_host": nbytes,
                  "snapshot_d2h_GBps": nbytes / (one * 1e-3) / 1e9}))
