#!/usr/bin/env python
"""bench.py -- sw2d fp64 DOF-stage-updates/s on B200 (BASELINE.json metric), one JSON line.

  python bench.py --gpus N --steps K --warmup W            # the CUDA path (this repo)
  python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (oracle port) on host cores

A "step" is one output interval of the reference driver: OUTPUT_INTERVAL LSERK4 time steps (5 fused
RHS+stage launches each) over the synthetic mesh.  N=1 workload: BASELINE configs[1], 1M structured
triangles, N=4; N>1: the same 1M elements per GPU (weak scaling), horizontal strips of a 1000 x 500N grid,
face-trace halos over NCCL send/recv.  DSU/s = 3*Np*K*5*lserk_steps / seconds.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

OUTPUT_INTERVAL = 20  # reference numParams.outputInterval (src/sw2d/main.cpp:60)
METRIC = "sw2d fp64 DOF-stage-updates/s"
UNIT = "DSU/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--order", type=int, default=4)
    ap.add_argument("--shape", default="tri", choices=["tri", "quad"])
    ap.add_argument("--nx", type=int, default=1000)
    ap.add_argument("--ny", type=int, default=500, help="cells in y PER GPU")
    ap.add_argument("--variant", type=int, default=0, help="0 auto, 1 generic, 2 fused")
    ap.add_argument("--mode", default="simple", choices=["simple", "full"])
    ap.add_argument("--stepper", default="lserk4", choices=["lserk4", "ssprk2"],
                    help="ssprk2 = the src/sw2d driver loop (SURVEY.md 8f rank 1): with --mode full the top edge becomes an open "
                         "boundary with tide, sloping 150-200 m bathymetry and a sponge behind the open boundary")
    ap.add_argument("--block", default="", help="element order: '' = cell-major, 'BXxBY' = blocks of BX x BY cells (locality study)")
    ap.add_argument("--shuffle", type=int, default=None, help="seed: random element / vertex numbering (unstructured-numbering study)")
    ap.add_argument("--jitter", type=float, default=0.0, help="interior vertex jitter in cell widths")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,enforced.power.limit")

    def __init__(self, gpu):
        self.gpu = gpu
        self.proc = None
        self.f = None

    def start(self):
        try:
            self.f = tempfile.NamedTemporaryFile("w+", delete=False)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.proc.wait()
        self.f.flush()
        rows = [r.strip().split(", ") for r in open(self.f.name) if r.strip()]
        os.unlink(self.f.name)
        sm, mx, reasons, pw, plim = [], None, set(), [], None
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            try:
                sm.append(float(r[1]))
                mx = float(r[2])
                for nm, val in zip(names, r[4:8]):
                    if val.strip().lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                continue
            try:  # board power is informative only: never let it cost the clocks line
                pw.append(float(r[3]))
                plim = float(r[8])
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm),
                "power_w": float(np.median(pw)) if pw else None, "power_max_w": float(np.max(pw)) if pw else None,
                "power_limit_w": plim}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def build_problem(args, rank, world):
    """Global strip mesh -> this rank's DGContext (+ halo) and initial state, all through the host setup mirror."""
    from blitzdg_b200.initial_conditions import gaussian_hump
    from blitzdg_b200.partition import build_local, strip_partition
    from blitzdg_b200.synthetic import provision, structured_mesh

    block = tuple(int(x) for x in args.block.split("x")) if args.block else None
    mesh = structured_mesh(args.shape, args.nx, args.ny * world, jitter=args.jitter, block=block, shuffle=args.shuffle)
    if world == 1:
        ctx = provision(mesh, args.order).get_DGContext()
        halo = None
    else:
        K, Nf = mesh.get_NumElements(), mesh.get_NumFaces()
        cy = mesh.get_Vertices().reshape(-1, 3)[mesh.get_Elements().reshape(K, Nf), 1].mean(axis=1)
        epart = strip_partition(cy, world)
        ctx, halo, _ = build_local(mesh, epart, rank, args.order)
    H, h, hu, hv = gaussian_hump(ctx.x, ctx.y)
    if args.stepper == "ssprk2" and args.mode == "full":
        H, h = coastal_setup(ctx, h - H)
    return ctx, halo, (H, h, hu, hv), mesh.get_NumElements()


def coastal_setup(ctx, eta):
    """The coastal workload of --stepper ssprk2 --mode full: boundary faces on the top edge (y = max) become Out faces
    (tidal forcing, reference src/sw2d/main.cpp:348-353), depth 150-200 m sloping in x.  Returns H, h."""
    wall = np.asarray(ctx.BCmap[3]).reshape(-1, ctx.Nfp)
    yv = ctx.y.T.reshape(-1)[np.asarray(ctx.vmapM)[wall]]
    top = (yv > ctx.y.max() - 1e-9).all(axis=1)
    ctx.BCmap = {3: wall[~top].reshape(-1), 2: wall[top].reshape(-1)}
    H = 175.0 + 25.0 * ctx.x
    return H, H + eta


def cpu_oracle_rate(args, threads, target_seconds=12.0):
    """Times the CPU oracle (port of the reference arithmetic + sweep structure) on a bounded sample of the
    workload: `threads` independent single-threaded instances (the reference has no threading), each on its own
    250x100 slab (50k elements) of the same mesh family.  Returns (DSU/s aggregate, sample description)."""
    from blitzdg_b200.initial_conditions import gaussian_hump
    from blitzdg_b200.synthetic import provision, structured_mesh
    from oracle.binding import Oracle

    nxs, nys = (250, 100) if args.shape == "tri" else (250, 200)
    mesh = structured_mesh(args.shape, nxs, nys)
    ctx = provision(mesh, args.order).get_DGContext()
    H, h, hu, hv = gaussian_hump(ctx.x, ctx.y)
    full = args.mode == "full"
    rk2 = args.stepper == "ssprk2"
    Hx = Hy = np.zeros_like(H)
    sponge = None
    if rk2 and full:
        from blitzdg_b200.sw2d import bedSlopes, buildSpongeCoeff

        H, h = coastal_setup(ctx, h - H)
        Hx, Hy = bedSlopes(ctx, H)
        sponge = buildSpongeCoeff(ctx, 10000.0, 0.1)
    orcs = []
    for _ in range(threads):
        o = Oracle(ctx, CFL=0.65, lf_mode=1 if full else 0, sources=1 if full else 0)
        if full:
            o.set_bathymetry(H, Hx, Hy)
        orcs.append((o, h.copy(), hu.copy(), hv.copy(), [np.zeros_like(h) for _ in range(3)]))
    dt = 0.5 * orcs[0][0].timestep(h, hu, hv, use_abs=True)[0]
    t0 = time.perf_counter()
    def advance(i, n):
        o, a, b, c, r = orcs[i]
        if rk2:
            o.step_ssprk2(a, b, c, n, dt, sponge=sponge)
        else:
            o.step(a, b, c, r, n, dt)

    advance(0, 1)  # calibrate (also warm-up)
    one = time.perf_counter() - t0
    nsteps = max(1, int(target_seconds / max(one, 1e-3)))

    def work(i):
        advance(i, nsteps)

    ths = [threading.Thread(target=work, args=(i,)) for i in range(threads)]
    t0 = time.perf_counter()
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    el = time.perf_counter() - t0
    dsu = 3.0 * ctx.Np * ctx.K * (2 if rk2 else 5) * nsteps * threads
    sample = "%d thread(s) x %d " % (threads, nsteps) + ("SSP-RK2" if rk2 else "LSERK4") + " steps on a %dx%d %s slab (%d elements, N=%d), %.1f s" % (
        nxs, nys, args.shape, ctx.K, args.order, el)
    return dsu / el, sample, el / nsteps


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    vals = []
    info = None
    for _ in range(max(1, min(args.warmup, 1))):
        cpu_oracle_rate(args, cores, target_seconds=2.0)
    per = max(2.0, min(15.0, 120.0 / max(1, args.steps)))
    t_all = time.perf_counter()
    for _ in range(args.steps):
        v, sample, _ = cpu_oracle_rate(args, cores, target_seconds=per)
        vals.append(v)
        info = sample
    value = float(np.mean(vals))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * (time.perf_counter() - t_all) / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, 1, None),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": info},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(args, world, variant):
    fam = "triangles" if args.shape == "tri" else "quads"
    Kper = args.nx * args.ny * (2 if args.shape == "tri" else 1)
    rk2 = args.stepper == "ssprk2"
    if rk2:
        return {
            "workload": "sw2d driver loop (SSP-RK2 + sponge relaxation, reference src/sw2d/main.cpp:192-244) on synthetic structured %s mesh, "
                        "%d elements per GPU, N=%d, %s" % (fam, Kper, args.order,
                        "full physics: global LF, star states, sources, open top edge with tide, 150-200 m bathymetry, sponge" if args.mode == "full"
                        else "sw2d-simple physics, all-wall BC, no sponge"),
            "elements_per_gpu": Kper, "order": args.order, "ssprk2_steps_per_step": OUTPUT_INTERVAL, "stages_per_step": 2 * OUTPUT_INTERVAL,
            "partition": "single GPU" if world == 1 else "%d horizontal strips, NCCL send/recv halo" % world,
            "l2": "working set (>=1 GB of state per GPU) exceeds the 126 MB L2; no explicit flush", "kernel_variant": variant,
        }
    return {
        "workload": "sw2d on synthetic structured %s mesh, %d elements per GPU (%dx%d cells), N=%d, LSERK4, per-face LF, all-wall BC"
                    % (fam, Kper, args.nx, args.ny, args.order) if args.mode == "simple" else
                    "sw2d full physics (global LF, star states, sources) on synthetic %s mesh, %d elements per GPU, N=%d" % (fam, Kper, args.order),
        "elements_per_gpu": Kper, "order": args.order, "lserk_steps_per_step": OUTPUT_INTERVAL, "stages_per_step": 5 * OUTPUT_INTERVAL,
        "partition": "single GPU" if world == 1 else "%d horizontal strips; face-trace halo per stage by direct NVLink peer stores from the sender's kernel + epoch flags (NCCL send/recv fallback)" % world,
        "l2": "working set (>=1 GB of state per GPU) exceeds the 126 MB L2; no explicit flush",
        "kernel_variant": variant,
    }


def run_b200(args):
    import torch

    from blitzdg_b200.sw2d import SW2dSolver, nccl_unique_id

    rank, world, local = dist_env()
    if world != args.gpus:
        if rank == 0 and world == 1 and args.gpus > 1:
            print("bench.py: --gpus %d needs torchrun (WORLD_SIZE=%d)" % (args.gpus, world), file=sys.stderr)
            sys.exit(2)
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    ctx, halo, (H, h, hu, hv), K_global = build_problem(args, rank, world)
    sol = SW2dSolver(ctx, mode=args.mode, halo=halo, device=local)
    if args.variant:
        sol.set_variant(args.variant)
    rk2 = args.stepper == "ssprk2"
    nst = 2 if rk2 else 5  # RHS evaluations per time step
    has_sponge = False
    if rk2 and args.mode == "full":
        from blitzdg_b200.sw2d import bedSlopes, buildSpongeCoeff

        Hx, Hy = bedSlopes(ctx, H)
        sol.set_bathymetry(H, Hx, Hy)
        sol.set_sponge(buildSpongeCoeff(ctx, 10000.0, 0.1))
        has_sponge = True
    else:
        sol.set_bathymetry(H, np.zeros_like(H), np.zeros_like(H))  # flat bottom; also feeds eta = h - H of the blow-up guard
    if world > 1:
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt = torch.frombuffer(bytearray(nccl_unique_id()), dtype=torch.uint8).cuda()
        dist.broadcast(idt, 0)
        sol.comm_init(bytes(idt.cpu().numpy().tobytes()))
    sol.set_state(h, hu, hv)
    dt = 0.5 * sol.computeTimeStep() * (0.65 / sol.num.CFL)  # reference formula (sw2d-simple CFL 0.65), halved for LSERK4 margin
    if rk2:
        dt = sol.computeTimeStep()  # the src/sw2d driver's own CFL 0.25 (main.cpp:56)
    stream = torch.cuda.ExternalStream(sol.stream, device=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        sol.sync()

    def timed(fn, steps, tail=None):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            fn()
        if tail is not None:
            tail()  # host-side join of side-stream work: e1 is recorded, and reached, after it
        e1.record(stream)
        sol.sync()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
        if dist is not None:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        barrier()
        return float(ms.item())

    # ---- device-resident throughput ------------------------------------------------------------
    advance = sol.step_ssprk2 if rk2 else sol.step
    kstep = lambda: advance(OUTPUT_INTERVAL, dt, sync=False)
    for _ in range(max(3, args.warmup)):
        kstep()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = sol.launch_count
    ms = timed(kstep, args.steps)
    launches = sol.launch_count - l0
    clocks = sampler.stop() if rank == 0 else None
    eta_max = sol.check()
    stages = nst * OUTPUT_INTERVAL * args.steps
    dsu_total = 3.0 * ctx.Np * K_global * stages
    value = dsu_total / (ms * 1e-3)
    lt = torch.tensor([launches], device="cuda", dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(lt)
    peak, peak_src = peaks()
    alg_bytes = sol.algorithmic_bytes_per_stage  # this rank, one stage == one launch of the dominant kernel
    if rk2:  # no residual planes: stage 1 reads q, writes q1 (48 B/node), stage 2 reads q1 and q, writes q (72 B/node); sponge 8 B/node
        alg_bytes += (60.0 - 96.0 + (8.0 if has_sponge else 0.0)) * ctx.Np * ctx.K
    stage_ms = ms / stages
    achieved = alg_bytes / (stage_ms * 1e-3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        try:
            traffic = json.load(open(tp)).get("%s_N%d_v%d" % (args.shape, args.order, sol.variant))
        except Exception:
            traffic = None

    # ---- end to end through the C ABI with host buffers -------------------------------------------
    e2e = None
    if not args.no_e2e:
        pin = [torch.empty(h.shape, dtype=torch.float64).pin_memory() for _ in range(6)]
        src = [p.numpy() for p in pin[:3]]
        dst = [p.numpy() for p in pin[3:]]
        for a, b in zip(src, (h, hu, hv)):
            a[...] = b

        def e2e_step():
            sol.set_state(src[0], src[1], src[2])
            advance(OUTPUT_INTERVAL, dt, sync=False)
            sol.get_state(out=dst)

        e2e_step()
        n_e2e = max(1, min(args.steps, 5))
        ms2 = timed(e2e_step, n_e2e)
        bytes_dir = 3 * ctx.Np * ctx.K * 8
        stepper = "bdg_sw2d_step_ssprk2" if rk2 else "bdg_sw2d_step"
        serial = 3.0 * ctx.Np * K_global * nst * OUTPUT_INTERVAL * n_e2e / (ms2 * 1e-3)
        # The same batches through the streaming calls: every step still uploads its own input from pinned host memory
        # and downloads its own result, but the upload of step i+1 and the download of step i-1 overlap the launches
        # of step i (both PCIe directions + the SMs busy at once).  The first upload and the last download are exposed.
        dst2 = [dst, [torch.empty(h.shape, dtype=torch.float64).pin_memory().numpy() for _ in range(3)]]
        n_pipe = max(2, min(args.steps, 10))
        it = {"i": 0, "n": 0}

        def pipe_step():
            i = it["i"]
            it["i"] += 1
            if i == 0:
                sol.stage_state(src[0], src[1], src[2])
            sol.commit_state()
            if i + 1 < it["n"]:
                sol.stage_state(src[0], src[1], src[2])
            advance(OUTPUT_INTERVAL, dt, sync=False)
            sol.get_state_async(dst2[i & 1])

        it.update(i=0, n=2)
        timed(pipe_step, 2, tail=sol.get_state_wait)  # warm-up: allocates the streaming buffers
        it.update(i=0, n=n_pipe)
        ms3 = timed(pipe_step, n_pipe, tail=sol.get_state_wait)
        piped = 3.0 * ctx.Np * K_global * nst * OUTPUT_INTERVAL * n_pipe / (ms3 * 1e-3)
        same = all(np.array_equal(a, b) for a, b in zip(dst2[0], dst2[1])) and np.array_equal(dst2[0][0], dst[0])
        e2e = {"value": piped, "unit": UNIT, "h2d_bytes_per_step": bytes_dir, "d2h_bytes_per_step": bytes_dir, "steps": n_pipe,
               "api": "per step: bdg_sw2d_commit_state (input staged by bdg_sw2d_stage_state from pinned host) -> %s -> bdg_sw2d_get_state_async; "
                      "copies of neighbouring steps overlap the launches; joined by bdg_sw2d_get_state_wait inside the timed region" % stepper,
               "serial": {"value": serial, "steps": n_e2e, "api": "bdg_sw2d_set_state (pinned host) -> %s -> bdg_sw2d_get_state, nothing overlapped" % stepper},
               "results_identical_to_serial": bool(same)}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        v, sample, _ = cpu_oracle_rate(args, 1, target_seconds=12.0)
        cpu = {"value": v, "unit": UNIT, "cores": 1, "kind": "port", "sample": sample}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": workload_config(args, world, sol.variant),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "peak_source": peak_src, "algorithmic_bytes_per_launch": alg_bytes,
                         "launch_ms": stage_ms,
                         "per": "rank 0, one RHS+stage: " + ("one fused launch" if sol.variant == 2 and not rk2 else "the launches of the two-kernel path (face, volume)")},
            "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(lt.item()), "clocks": clocks,
            "eta_max_after": eta_max, "dt": dt,
        }
        print(json.dumps(line), flush=True)
    sol.close()
    if dist is not None:
        dist.destroy_process_group()


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
