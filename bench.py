#!/usr/bin/env python
"""bench.py -- sw2d fp64 DOF-stage-updates/s on B200 (BASELINE.json metric), one JSON line.

  python bench.py --gpus N --steps K --warmup W            # the CUDA path (this repo)
  python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (oracle port) on host cores

A "step" of the CUDA arm is one output interval of the reference driver: OUTPUT_INTERVAL LSERK4 time steps (5 fused
RHS+stage launches each) over the synthetic mesh.  DSU/s = 3*Np*K*5*lserk_steps / seconds.

Workloads (--workload; "auto" = cfg2 at N=1, cfg5 at N>1):
  cfg2  BASELINE configs[1]: 1 M structured triangles, N=4, per GPU (N>1: weak scaling, 1 M elements per GPU)
  cfg3  configs[2]: 4 M triangles, N=8           cfg4  configs[3]: 2 M quads, N=6
  cfg5  configs[4]: the 16 M-triangle mesh, N=4, split over the N GPUs by MeshManager.partitionMesh (METIS), face-trace
        halo over NVLink; total work fixed ("scaling": "strong" among N = 2, 4, 8; N=1 runs cfg2, the config the metric
        is quoted on, and --workload cfg5 --gpus 1 gives the single-GPU figure of the same mesh)
  custom  --shape/--order/--nx/--ny (cells in y PER GPU)
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

OUTPUT_INTERVAL = 20  # reference numParams.outputInterval (src/sw2d/main.cpp:60)
METRIC = "sw2d fp64 DOF-stage-updates/s"
UNIT = "DSU/s"
FP64_PEAK_TFLOPS = 37.1  # measured on this pool: DFMA and DMMA m8n8k4 share one pipe (profiles/r01_fp64_microbench.txt)

WORKLOADS = {  # shape, order, nx, ny, ny is per GPU?
    "cfg2": ("tri", 4, 1000, 500, True),
    "cfg3": ("tri", 8, 2000, 1000, False),
    "cfg4": ("quad", 6, 2000, 1000, False),
    "cfg5": ("tri", 4, 4000, 2000, False),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="auto", choices=["auto", "custom"] + sorted(WORKLOADS))
    ap.add_argument("--order", type=int, default=None)
    ap.add_argument("--shape", default=None, choices=["tri", "quad"])
    ap.add_argument("--nx", type=int, default=None)
    ap.add_argument("--ny", type=int, default=None, help="cells in y PER GPU (custom workload)")
    ap.add_argument("--partition", default="metis", choices=["metis", "strips", "rcb"])
    ap.add_argument("--variant", type=int, default=0, help="0 auto, 1 generic, 2 fused")
    ap.add_argument("--mode", default="simple", choices=["simple", "full"])
    ap.add_argument("--stepper", default="lserk4", choices=["lserk4", "ssprk2"],
                    help="ssprk2 = the src/sw2d driver loop (SURVEY.md 8f rank 1): with --mode full the top edge becomes an open "
                         "boundary with tide, sloping 150-200 m bathymetry and a sponge behind the open boundary")
    ap.add_argument("--block", default="", help="element order: '' = cell-major, 'BXxBY' = blocks of BX x BY cells (locality study)")
    ap.add_argument("--shuffle", type=int, default=None, help="seed: random element / vertex numbering (unstructured-numbering study)")
    ap.add_argument("--jitter", type=float, default=0.0, help="interior vertex jitter in cell widths")
    ap.add_argument("--reps", type=int, default=3, help="repetitions of the K-step timed region (value = median, best also reported)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="N>1, workload auto: skip the 1 M-elements-per-GPU weak-scaling line")
    args = ap.parse_args()
    custom = any(v is not None for v in (args.order, args.shape, args.nx, args.ny))
    if args.workload == "auto":
        args.workload = "custom" if custom else ("cfg2" if args.gpus == 1 else "cfg5")
        args.auto = not custom
    else:
        args.auto = False
    if args.workload == "custom":
        args.shape, args.order = args.shape or "tri", args.order or 4
        args.nx, args.ny, args.ny_per_gpu = args.nx or 1000, args.ny or 500, True
    else:
        args.shape, args.order, args.nx, args.ny, args.ny_per_gpu = WORKLOADS[args.workload]
    return args


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured copy bandwidth (MEASURED_PEAKS.json hbm_gbs)"
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


def global_mesh(args, world):
    from blitzdg_b200.synthetic import structured_mesh

    block = tuple(int(x) for x in args.block.split("x")) if args.block else None
    ny = args.ny * world if args.ny_per_gpu else args.ny
    return structured_mesh(args.shape, args.nx, ny, jitter=args.jitter, block=block, shuffle=args.shuffle)


def partition_on_rank0(mesh, world, how, dist, rank):
    """The element partition is computed ONCE (rank 0: MeshManager.partitionMesh = METIS_PartMeshNodal with the reference's
    options) and broadcast, so that every rank cuts the mesh identically."""
    import torch

    from blitzdg_b200.partition import make_partition, partition_stats

    K = mesh.get_NumElements()
    buf = torch.zeros(K, dtype=torch.int32, device="cuda")
    desc = [None]
    if rank == 0:
        t0 = time.perf_counter()
        epart, d = make_partition(mesh, world, how)
        desc[0] = (d, time.perf_counter() - t0)
        buf.copy_(torch.from_numpy(np.ascontiguousarray(epart, dtype=np.int32)))
    dist.broadcast(buf, 0)
    dist.broadcast_object_list(desc, 0)
    epart = buf.cpu().numpy()
    st = partition_stats(mesh, epart)
    st["method"], st["seconds"] = desc[0]
    return epart, st


def build_problem(args, rank, world, dist=None):
    """Global synthetic mesh -> this rank's DGContext (+ halo) and initial state, all through the host setup mirror."""
    from blitzdg_b200.initial_conditions import gaussian_hump
    from blitzdg_b200.partition import build_local
    from blitzdg_b200.synthetic import provision

    mesh = global_mesh(args, world)
    pstats = None
    if world == 1:
        ctx = provision(mesh, args.order).get_DGContext()
        halo = None
    else:
        epart, pstats = partition_on_rank0(mesh, world, args.partition, dist, rank)
        ctx, halo, _ = build_local(mesh, epart, rank, args.order)
    H, h, hu, hv = gaussian_hump(ctx.x, ctx.y)
    if args.stepper == "ssprk2" and args.mode == "full":
        H, h = coastal_setup(ctx, h - H)
    return ctx, halo, (H, h, hu, hv), mesh.get_NumElements(), pstats


def coastal_setup(ctx, eta):
    """The coastal workload of --stepper ssprk2 --mode full: boundary faces on the top edge (y = max) become Out faces
    (tidal forcing, reference src/sw2d/main.cpp:348-353), depth 150-200 m sloping in x.  Returns H, h."""
    wall = np.asarray(ctx.BCmap[3]).reshape(-1, ctx.Nfp)
    yv = ctx.y.T.reshape(-1)[np.asarray(ctx.vmapM)[wall]]
    top = (yv > ctx.y.max() - 1e-9).all(axis=1)
    ctx.BCmap = {3: wall[~top].reshape(-1), 2: wall[top].reshape(-1)}
    H = 175.0 + 25.0 * ctx.x
    return H, H + eta


# ------------------------------------------------------------------------------------------------ CPU reference legs
def cpu_reference(args, world, threads, budget_s, steps=1, warmup=0, max_elements=None):
    """The oracle (port of the reference arithmetic + sweep structure) on the ARM'S OWN MESH.

    threads == 1: one Oracle instance over the whole mesh, exactly the single-threaded reference.  threads > 1: the same
    mesh cut into strips with a trace exchange per stage (oracle/threaded.py) on `threads` host threads -- the reference
    itself has no threading, this is the courtesy "all host cores" figure.  One timed "step" = n_stage RHS+stage
    evaluations over the whole mesh, n_stage chosen from a calibration stage so that steps+warmup fit `budget_s`.
    Returns (DSU/s, seconds per step, sample description, n_stage, setup seconds)."""
    from blitzdg_b200.initial_conditions import gaussian_hump
    from blitzdg_b200.synthetic import provision
    from oracle.binding import Oracle

    full, rk2 = args.mode == "full", args.stepper == "ssprk2"
    per_cell = 2 if args.shape == "tri" else 1
    ny_tot = args.ny * world if args.ny_per_gpu else args.ny
    slab = ""
    if max_elements and args.nx * ny_tot * per_cell > max_elements:  # bounded sample: the bottom rows of the same mesh
        from blitzdg_b200.synthetic import structured_mesh

        rows = max(4, int(max_elements // (args.nx * per_cell)))
        mesh = structured_mesh(args.shape, args.nx, rows, jitter=args.jitter)  # same cell size in x; y rescaled to [-1,1]
        slab = " [bounded sample: a %d x %d-cell slab of the %d x %d mesh]" % (args.nx, rows, args.nx, ny_tot)
    else:
        mesh = global_mesh(args, world)
    K = mesh.get_NumElements()
    t_setup = time.perf_counter()
    if full or rk2:
        threads = 1
        ctx = provision(mesh, args.order).get_DGContext()
        H, h, hu, hv = gaussian_hump(ctx.x, ctx.y)
        Hx = Hy = np.zeros_like(H)
        sponge = None
        if rk2 and full:
            from blitzdg_b200.sw2d import bedSlopes, buildSpongeCoeff

            H, h = coastal_setup(ctx, h - H)
            Hx, Hy = bedSlopes(ctx, H)
            sponge = buildSpongeCoeff(ctx, 10000.0, 0.1)
        orc = Oracle(ctx, CFL=0.65, lf_mode=1 if full else 0, sources=1 if full else 0)
        if full:
            orc.set_bathymetry(H, Hx, Hy)
        dt = 0.5 * orc.timestep(h, hu, hv, use_abs=True)[0]
        res = [np.zeros_like(h) for _ in range(3)]
        Np = ctx.Np
        if rk2:
            per_call, what = 2, "SSP-RK2 step(s)"
            advance = lambda n: orc.step_ssprk2(h, hu, hv, n, dt, sponge=sponge)
        else:
            per_call, what = 5, "LSERK4 step(s)"
            advance = lambda n: orc.step(h, hu, hv, res, n, dt)
    else:
        from oracle.threaded import StripOracle

        # threads == 1: ONE Oracle instance over the whole mesh, advanced stage by stage (the same C calls as oracle_sw2d_step)
        so = StripOracle(mesh, args.order, nparts=1 if threads == 1 else None, threads=threads, CFL=0.65)
        Np = so.Np
        xs = np.zeros((Np, K))
        ys = np.zeros((Np, K))
        for p in so.parts:
            xs[:, p["el"]] = p["ctx"].x
            ys[:, p["el"]] = p["ctx"].y
        H, h, hu, hv = gaussian_hump(xs, ys)
        del xs, ys
        so.set_state(h, hu, hv)
        dt = 0.5 * min(p["orc"].timestep(p["q"][0], p["q"][1], p["q"][2], use_abs=True)[0] for p in so.parts)
        per_call, what = 1, "LSERK4 stage(s)"
        advance = lambda n: so.stages(n, dt)
    t_setup = time.perf_counter() - t_setup
    t0 = time.perf_counter()
    advance(1)  # calibration (also the first warm-up)
    one = time.perf_counter() - t0
    n = max(1, int(budget_s / max(1, steps + warmup) / max(one, 1e-3)))
    if per_call == 1 and n < 5 and steps == 1 and one * 5 <= 35.0:
        n = 5  # the 1-thread baseline leg times at least one whole LSERK4 step where that stays within ~30 s
    if per_call == 1 and n >= 5:
        n -= n % 5  # whole LSERK4 steps
    for _ in range(max(0, warmup - 1)):
        advance(n)
    times = []
    for _ in range(steps):
        t0 = time.perf_counter()
        advance(n)
        times.append(time.perf_counter() - t0)
    sec = float(np.mean(times))
    n_stage = n * per_call
    dsu = 3.0 * Np * K * n_stage / sec
    fam = "triangles" if args.shape == "tri" else "quads"
    sample = "%d timed step(s), each %d %s = %d RHS+stage evaluations over the arm's own mesh (%d %s, N=%d)%s, %d host thread(s)%s, %.1f s per step" % (
        steps, n, what, n_stage, K, fam, args.order, slab, threads,
        "" if threads == 1 else " over %d strips with a trace exchange per stage" % so.nparts, sec)
    return dsu, sec, sample, n_stage, t_setup, threads


def verbatim_reference_rate(args):
    """Courtesy figure beside the reference arm: the reference's OWN computeRHS, compiled verbatim from /root/reference behind the
    blitz stand-in (oracle/_ref/libsw2d_ref.so, oracle/make_ref.py), against the oracle port on the same slab, one thread each,
    RHS evaluations only.  Says which of the two is the faster CPU code (the arm times the port); None where the library or the
    case (quads / full physics need more set-up than a courtesy figure is worth) is absent."""
    try:
        from oracle import ref_binding

        if args.shape != "tri" or args.mode == "full" or not ref_binding.available():
            return None
        from blitzdg_b200.initial_conditions import gaussian_hump
        from blitzdg_b200.synthetic import provision, structured_mesh
        from oracle.binding import Oracle

        rows = max(4, 50_000 // (2 * args.nx))
        ctx = provision(structured_mesh("tri", args.nx, rows), args.order).get_DGContext()
        H, h, hu, hv = gaussian_hump(ctx.x, ctx.y)
        ref, orc = ref_binding.CompiledReference(ctx), Oracle(ctx, CFL=0.65, lf_mode=0, sources=0)
        out = {}
        for name, f in (("verbatim", lambda: ref.rhs_simple(h, hu, hv)), ("port", lambda: orc.rhs(h, hu, hv))):
            f()
            t0 = time.perf_counter()
            for _ in range(3):
                f()
            out[name] = 3.0 * ctx.Np * ctx.K / ((time.perf_counter() - t0) / 3)
        return {"verbatim_rhs_per_s": out["verbatim"], "port_rhs_per_s": out["port"], "unit": "DOF RHS evaluations/s", "cores": 1,
                "sample": "3 RHS evaluations each on a %d x %d-cell slab (%d triangles, N=%d) of the arm's mesh; verbatim = the reference's "
                          "sw2d-simple computeRHS compiled from /root/reference behind oracle/blitz_shim (not blitz++ itself)" % (args.nx, rows, ctx.K, args.order)}
    except Exception as e:  # a courtesy figure never costs the arm its line
        return {"unavailable": repr(e)[:200]}


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return
    world = args.gpus
    cores = os.cpu_count() or 1
    value, sec, sample, n_stage, t_setup, used = cpu_reference(args, world, cores, budget_s=100.0, steps=max(1, args.steps), warmup=max(1, args.warmup))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * sec, "higher_is_better": True,
        "scaling": scaling_kind(args), "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, world, None, None),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": used, "kind": "port", "sample": sample, "stages_per_step": n_stage,
                         "host_setup_s": t_setup, "verbatim_reference": verbatim_reference_rate(args)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def scaling_kind(args):
    return "weak" if args.ny_per_gpu else "strong"


def workload_config(args, world, variant, pstats):
    fam = "triangles" if args.shape == "tri" else "quads"
    ny_tot = args.ny * world if args.ny_per_gpu else args.ny
    Ktot = args.nx * ny_tot * (2 if args.shape == "tri" else 1)
    rk2 = args.stepper == "ssprk2"
    if world == 1:
        part = "single GPU"
    else:
        part = "%d parts by %s; face-trace halo per stage by direct NVLink peer stores from the sender's kernel + epoch flags (NCCL send/recv fallback)" % (
            world, pstats["method"] if pstats else {"metis": "METIS_PartMeshNodal", "rcb": "recursive coordinate bisection", "strips": "horizontal strips"}[args.partition])
    cfg = {
        "workload": "%s: sw2d on synthetic structured %s mesh, %d elements in total (%dx%d cells), %s, N=%d, " % (
            args.workload, fam, Ktot, args.nx, ny_tot, "%d per GPU" % (Ktot // world), args.order) + (
            "src/sw2d driver loop (SSP-RK2 + sponge relaxation, reference src/sw2d/main.cpp:192-244), " if rk2 else "LSERK4, ") + (
            "full physics (global LF, star states, sources%s)" % (", open top edge with tide, 150-200 m bathymetry, sponge" if rk2 else "") if args.mode == "full"
            else "sw2d-simple physics (per-face LF, all-wall BC)"),
        "baseline_config": {"cfg2": "configs[1]", "cfg3": "configs[2]", "cfg4": "configs[3]", "cfg5": "configs[4]"}.get(args.workload),
        "elements_total": Ktot, "elements_per_gpu": Ktot // world, "order": args.order,
        "steps_per_step": OUTPUT_INTERVAL, "stages_per_step": (2 if rk2 else 5) * OUTPUT_INTERVAL,
        "partition": part,
        "l2": "working set (>=1 GB of state per GPU) exceeds the 126 MB L2; no explicit flush",
        "kernel_variant": variant,
    }
    if pstats:
        cfg["partition_stats"] = {k: pstats[k] for k in ("elements", "neighbours", "cut_faces", "cut_faces_total", "seconds")}
    return cfg


# ------------------------------------------------------------------------------------------------------ the CUDA arm
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

    def comm_init(sol):
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt = torch.frombuffer(bytearray(nccl_unique_id()), dtype=torch.uint8).cuda()
        dist.broadcast(idt, 0)
        sol.comm_init(bytes(idt.cpu().numpy().tobytes()))

    def barrier(sol=None):
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        if sol is not None:
            sol.sync()

    rk2 = args.stepper == "ssprk2"
    nst = 2 if rk2 else 5  # RHS evaluations per time step

    def measure(a, with_extras):
        """Builds the workload `a`, times it; returns the dict of everything measured on it."""
        t_host = time.perf_counter()
        ctx, halo, (H, h, hu, hv), K_global, pstats = build_problem(a, rank, world, dist)
        sol = SW2dSolver(ctx, mode=a.mode, halo=halo, device=local)
        if a.variant:
            sol.set_variant(a.variant)
        has_sponge = False
        if rk2 and a.mode == "full":
            from blitzdg_b200.sw2d import bedSlopes, buildSpongeCoeff

            Hx, Hy = bedSlopes(ctx, H)
            sol.set_bathymetry(H, Hx, Hy)
            sol.set_sponge(buildSpongeCoeff(ctx, 10000.0, 0.1))
            has_sponge = True
        else:
            sol.set_bathymetry(H, np.zeros_like(H), np.zeros_like(H))  # flat bottom; also feeds eta = h - H of the blow-up guard
        if world > 1:
            comm_init(sol)
        sol.set_state(h, hu, hv)
        dt = 0.5 * sol.computeTimeStep() * (0.65 / sol.num.CFL)  # reference formula (sw2d-simple CFL 0.65), halved for LSERK4 margin
        if rk2:
            dt = sol.computeTimeStep()  # the src/sw2d driver's own CFL 0.25 (main.cpp:56)
        t_host = time.perf_counter() - t_host
        stream = torch.cuda.ExternalStream(sol.stream, device=torch.device("cuda", local))

        def timed(fn, steps, tail=None):
            barrier(sol)
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
            barrier(sol)
            return float(ms.item())

        # ---- device-resident throughput: `reps` repetitions of exactly K steps, value = median ------------------------
        advance = sol.step_ssprk2 if rk2 else sol.step
        kstep = lambda: advance(OUTPUT_INTERVAL, dt, sync=False)
        for _ in range(max(3, a.warmup)):
            kstep()
        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()
        l0 = sol.launch_count
        times = [timed(kstep, a.steps) for _ in range(max(1, a.reps))]
        launches = (sol.launch_count - l0) // max(1, a.reps)
        clocks = sampler.stop() if rank == 0 else None
        eta_max = sol.check()
        stages = nst * OUTPUT_INTERVAL * a.steps
        dsu_total = 3.0 * ctx.Np * K_global * stages
        ms = float(np.median(times))
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
        Nfn = ctx.Nfaces * ctx.Nfp
        flops_stage = (20.0 * ctx.Np * ctx.Np + 6.0 * ctx.Np * Nfn + 100.0 * ctx.Np) * ctx.K  # SURVEY.md 8d, min-sharing count
        fp64_fraction = flops_stage / (stage_ms * 1e-3) / (FP64_PEAK_TFLOPS * 1e12)
        traffic, traffic_src = None, None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            try:
                tj = json.load(open(tp))
                key = "%s_N%d_v%d%s" % (a.shape, a.order, sol.variant, "_full" if a.mode == "full" else "")
                if key in tj and abs(ctx.K - 1_000_000) < 1000:
                    traffic = tj[key]
                    traffic_src = "profiles/traffic.json[%s]: dram__bytes_read.sum + dram__bytes_write.sum of one ncu --set full capture of this kernel on this workload (a stored constant, not re-measured by this run)" % key
            except Exception:
                pass
        out = {
            "value": value, "ms": ms, "times_ms": times, "stages": stages, "launches": int(lt.item()), "clocks": clocks, "eta_max": eta_max,
            "dt": dt, "variant": sol.variant, "pstats": pstats, "host_setup_s": t_host, "K_local": int(ctx.K),
            "repetitions": {"n": len(times), "steps_each": a.steps, "best": dsu_total / (min(times) * 1e-3), "median": value,
                            "values": [dsu_total / (t * 1e-3) for t in times]},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "algorithmic_bytes_per_launch": alg_bytes,
                         "launch_ms": stage_ms, "fp64_fraction": fp64_fraction,
                         "fp64": {"flops_per_launch": flops_stage, "peak_tflops": FP64_PEAK_TFLOPS, "achieved_tflops": flops_stage / (stage_ms * 1e-3) / 1e12,
                                  "note": "flops = (20 Np^2 + 6 Np Nfn + 100 Np) K (SURVEY.md 8d); peak = measured DFMA/DMMA pipe (profiles/r01_fp64_microbench.txt)"},
                         "per": "rank 0, one RHS+stage: " + ("one fused launch" if sol.variant == 2 and not rk2 else "the launches of the two-kernel path (face, volume)")},
        }
        if not with_extras:
            sol.close()
            return out

        # ---- end to end through the C ABI with host buffers -------------------------------------------------------
        if not a.no_e2e:
            pin = [torch.empty(h.shape, dtype=torch.float64).pin_memory() for _ in range(6)]
            src = [p.numpy() for p in pin[:3]]
            dst = [p.numpy() for p in pin[3:]]
            for x, y in zip(src, (h, hu, hv)):
                x[...] = y
            bytes_dir = 3 * ctx.Np * ctx.K * 8
            stepper = "bdg_sw2d_step_ssprk2" if rk2 else "bdg_sw2d_step"

            def e2e_serial(interval):
                def f():
                    sol.set_state(src[0], src[1], src[2])
                    advance(interval, dt, sync=False)
                    sol.get_state(out=dst)
                return f

            e2e_serial(OUTPUT_INTERVAL)()
            n_e2e = max(1, min(a.steps, 5))
            ms2 = timed(e2e_serial(OUTPUT_INTERVAL), n_e2e)
            serial = 3.0 * ctx.Np * K_global * nst * OUTPUT_INTERVAL * n_e2e / (ms2 * 1e-3)
            ms1 = timed(e2e_serial(1), n_e2e)  # outputInterval = 1: the PCIe-bound end of the range
            serial1 = 3.0 * ctx.Np * K_global * nst * 1 * n_e2e / (ms1 * 1e-3)
            # The same batches through the streaming calls: every step still uploads its own input from pinned host memory
            # and downloads its own result, but the upload of step i+1 and the download of step i-1 overlap the launches
            # of step i (both PCIe directions + the SMs busy at once).  The first upload and the last download are exposed.
            dst2 = [dst, [torch.empty(h.shape, dtype=torch.float64).pin_memory().numpy() for _ in range(3)]]
            n_pipe = max(2, min(a.steps, 10))
            it = {"i": 0, "n": 0, "k": OUTPUT_INTERVAL}

            def pipe_step():
                i = it["i"]
                it["i"] += 1
                if i == 0:
                    sol.stage_state(src[0], src[1], src[2])
                sol.commit_state()
                if i + 1 < it["n"]:
                    sol.stage_state(src[0], src[1], src[2])
                advance(it["k"], dt, sync=False)
                sol.get_state_async(dst2[i & 1])

            it.update(i=0, n=2)
            timed(pipe_step, 2, tail=sol.get_state_wait)  # warm-up: allocates the streaming buffers
            it.update(i=0, n=n_pipe)
            ms3 = timed(pipe_step, n_pipe, tail=sol.get_state_wait)
            piped = 3.0 * ctx.Np * K_global * nst * OUTPUT_INTERVAL * n_pipe / (ms3 * 1e-3)
            same = all(np.array_equal(x, y) for x, y in zip(dst2[0], dst2[1])) and np.array_equal(dst2[0][0], dst[0])
            it.update(i=0, n=n_pipe, k=1)
            ms4 = timed(pipe_step, n_pipe, tail=sol.get_state_wait)
            piped1 = 3.0 * ctx.Np * K_global * nst * 1 * n_pipe / (ms4 * 1e-3)
            out["e2e"] = {
                "value": piped, "unit": UNIT, "h2d_bytes_per_step": bytes_dir, "d2h_bytes_per_step": bytes_dir, "steps": n_pipe,
                "api": "per step: bdg_sw2d_commit_state (input staged by bdg_sw2d_stage_state from pinned host) -> %s(%d steps) -> bdg_sw2d_get_state_async; "
                       "copies of neighbouring steps overlap the launches; joined by bdg_sw2d_get_state_wait inside the timed region" % (stepper, OUTPUT_INTERVAL),
                "serial": {"value": serial, "steps": n_e2e, "api": "bdg_sw2d_set_state (pinned host) -> %s -> bdg_sw2d_get_state, nothing overlapped" % stepper},
                "output_interval_1": {"value": piped1, "serial": serial1, "host_gbs_per_direction_per_gpu": bytes_dir * n_pipe / (ms4 * 1e-3) / 1e9,
                                      "note": "the same calls with ONE time step between upload and download (the reference's outputInterval = 1): the PCIe-bound end of the range"},
                "host_gbs_per_direction_per_gpu": bytes_dir * n_pipe / (ms3 * 1e-3) / 1e9,
                "results_identical_to_serial": bool(same)}

        # ---- parity of the timed configuration against the oracle and the arbiter (rank 0, single GPU) -------------
        if rank == 0 and world == 1 and not a.no_parity and not rk2 and a.mode == "simple" and ctx.K <= 2_100_000:
            out["parity"] = parity_single(a, ctx, sol, (h, hu, hv), local)
        sol.close()
        return out

    res = measure(args, True)

    parity_multi = None
    if world > 1 and not args.no_parity:
        parity_multi = multi_gpu_parity(args, rank, world, local, dist, comm_init)

    secondary = None
    if world > 1 and args.auto and not args.no_secondary:
        try:
            a2 = argparse.Namespace(**vars(args))
            a2.workload = "cfg2"
            a2.shape, a2.order, a2.nx, a2.ny, a2.ny_per_gpu = WORKLOADS["cfg2"]
            a2.reps = 1
            r2 = measure(a2, False)
            secondary = {"metric": METRIC, "value": r2["value"], "unit": UNIT, "scaling": "weak", "ms_per_step": r2["ms"] / a2.steps,
                         "config": workload_config(a2, world, r2["variant"], r2["pstats"]), "roofline_frac": r2["roofline"]["frac"]}
        except Exception as exc:  # the headline line must survive
            secondary = {"error": repr(exc)}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        v, sec, sample, n_stage, t_setup, used = cpu_reference(args, 1, 1, budget_s=20.0, steps=1, warmup=0, max_elements=1_050_000)
        cpu = {"value": v, "unit": UNIT, "cores": used, "kind": "port", "sample": sample}

    if rank == 0:
        par = res.get("parity") or {}
        line = {
            "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": res["ms"] / args.steps, "higher_is_better": True, "scaling": scaling_kind(args), "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": workload_config(args, world, res["variant"], res["pstats"]),
            "roofline": res["roofline"], "fp64_fraction": res["roofline"]["fp64_fraction"], "repetitions": res["repetitions"],
            "cpu_baseline": cpu, "e2e": res.get("e2e"), "gpu_launches": res["launches"], "clocks": res["clocks"],
            "max_rel_err_rhs": par.get("max_rel_err_rhs"), "max_rel_err_1000": par.get("max_rel_err_1000"), "parity": par or None,
            "parity_multi": parity_multi, "secondary": secondary,
            "eta_max_after": res["eta_max"], "dt": res["dt"], "host_setup_s": res["host_setup_s"],
        }
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def parity_single(args, ctx, sol, state, device):
    """max_rel_err_rhs of the TIMED configuration: the kernel's RHS of the initial state over the whole mesh against the
    oracle (reference arithmetic, fp64) and against the extended-precision arbiter; max_rel_err_1000: BASELINE config 1
    (coarse_box.msh, N=4, 1000 LSERK4 steps) through the same library.  The oracle is the checker here, never the timed path."""
    from oracle.binding import Arbiter, Oracle
    from oracle.threaded import StripOracle

    rel = lambda x, y: float(np.abs(x - y).max() / max(np.abs(y).max(), 1e-300))
    h, hu, hv = state
    hu = np.ascontiguousarray(0.5 * h * np.sin(np.pi * ctx.x) * np.cos(2.0 * ctx.y))  # a moving state without normal flow at the walls
    hv = np.ascontiguousarray(0.4 * h * np.cos(2.5 * ctx.x) * np.sin(np.pi * ctx.y))
    sol.set_state(h, hu, hv)
    R = sol.computeRHS()
    Ra = Arbiter(ctx).rhs(h, hu, hv)
    mesh = global_mesh(args, 1)
    so = StripOracle(mesh, args.order, threads=min(8, os.cpu_count() or 1))
    so.set_state(h, hu, hv)
    Ro = so.rhs()
    so.close()
    geo = None
    if sol.variant == 2:  # the fused kernel folds the per-node copies of the affine geometry into element / face means
        import copy

        cm = copy.copy(ctx)  # the geometry the fused kernel reads: element / face means, per-node normals on wall face nodes
        for name in ("rx", "sx", "ry", "sy"):
            a = getattr(ctx, name)
            setattr(cm, name, np.ascontiguousarray(np.broadcast_to(a.mean(axis=0, keepdims=True), a.shape)))
        Nfn = ctx.Nfaces * ctx.Nfp
        wall = np.asarray(ctx.BCmap.get(3, ()), dtype=np.int64)
        for name in ("nx", "ny", "Fscale"):
            a = getattr(ctx, name).reshape(ctx.Nfaces, ctx.Nfp, -1)
            m = np.ascontiguousarray(np.broadcast_to(a.mean(axis=1, keepdims=True), a.shape).reshape(Nfn, -1))
            if name != "Fscale" and wall.size:
                m[wall % Nfn, wall // Nfn] = getattr(ctx, name)[wall % Nfn, wall // Nfn]
            setattr(cm, name, m)
        Rm = Arbiter(cm).rhs(h, hu, hv)
        geo = {"cuda_vs_exact_on_averaged_geometry": max(rel(R[c], Rm[c]) for c in range(3)),
               "geometry_effect": max(rel(Rm[c], Ra[c]) for c in range(3)),
               "note": "geometry_effect = arbiter(per-node rx..Fscale as the reference stores them) vs arbiter(their element/face means + per-node "
                       "wall normals, what the 104 B/element affine kernel reads): the scatter of the reference's own geometry arrays, no kernel involved"}
    out = {"rhs_state": "Gaussian hump + smooth velocity field without normal flow at the walls, whole mesh (%d elements), kernel variant %d" % (ctx.K, sol.variant),
           "geometry": geo,
           "max_rel_err_rhs": max(rel(R[c], Ro[c]) for c in range(3)),
           "cuda_vs_exact": max(rel(R[c], Ra[c]) for c in range(3)),
           "oracle_vs_exact": max(rel(Ro[c], Ra[c]) for c in range(3)),
           "note": "max|d|/max|ref| per field, worst field; 'exact' = the long-double arbiter (oracle/sw2d_arbiter.c). At this mesh size the "
                   "reference arithmetic's own fp64 rounding error (oracle_vs_exact) exceeds 1e-12, see DESIGN.md section 2"}
    try:
        from blitzdg_b200 import MeshManager, TriangleNodesProvisioner
        from blitzdg_b200.initial_conditions import gaussian_hump
        from blitzdg_b200.sw2d import SW2dSolver

        m = MeshManager()
        m.readMesh(os.path.join(ROOT, "tests", "golden", "coarse_box.msh"))
        c1 = TriangleNodesProvisioner(4, m).get_DGContext()
        H1, h1, hu1, hv1 = gaussian_hump(c1.x, c1.y)
        o1 = Oracle(c1, CFL=0.65)
        dt1 = 0.5 * o1.timestep(h1, hu1, hv1, use_abs=True)[0]
        s1 = SW2dSolver(c1, device=device)
        s1.set_state(h1, hu1, hv1)
        s1.step(1000, dt1)
        got = s1.get_state()
        s1.close()
        ho, huo, hvo = h1.copy(), hu1.copy(), hv1.copy()
        o1.step(ho, huo, hvo, [np.zeros_like(h1) for _ in range(3)], 1000, dt1)
        out["max_rel_err_1000"] = max(rel(a, b) for a, b in zip(got[:3], (ho, huo, hvo)))
        out["steps_1000_config"] = "BASELINE configs[0]: coarse_box.msh, N=4, 1000 LSERK4 steps, dt = %.3e" % dt1
    except Exception as exc:
        out["max_rel_err_1000_error"] = repr(exc)
    return out


def multi_gpu_parity(args, rank, world, local, dist, comm_init):
    """Before a multi-GPU number is reported: 5 LSERK4 steps of a 10 k-element mesh partitioned `world` ways with the SAME
    partitioner and halo path, against rank 0's single-rank run through the same C ABI."""
    import torch

    from blitzdg_b200.initial_conditions import gaussian_hump
    from blitzdg_b200.partition import build_local
    from blitzdg_b200.sw2d import SW2dSolver
    from blitzdg_b200.synthetic import provision, structured_mesh

    try:
        N = args.order
        mesh = structured_mesh(args.shape, 80, 64)
        K = mesh.get_NumElements()
        epart, pstats = partition_on_rank0(mesh, world, args.partition, dist, rank)
        lctx, halo, info = build_local(mesh, epart, rank, N)
        el = info["elements"]
        H, h, hu, hv = gaussian_hump(lctx.x, lctx.y)
        hu = np.ascontiguousarray(0.5 * h * np.sin(3.0 * lctx.x + 1.0))
        hv = np.ascontiguousarray(0.4 * h * np.cos(2.5 * lctx.y))
        sol = SW2dSolver(lctx, mode=args.mode, halo=halo, device=local)
        if args.variant:
            sol.set_variant(args.variant)
        sol.set_bathymetry(H, np.zeros_like(H), np.zeros_like(H))
        comm_init(sol)
        sol.set_state(h, hu, hv)
        dt = 0.5 * sol.computeTimeStep() * (0.65 / sol.num.CFL)
        R = sol.computeRHS()
        sol.step(5, dt)
        q = sol.get_state()[:3]
        sol.check()
        var = sol.variant
        sol.close()
        full = torch.zeros((6, lctx.Np, K), dtype=torch.float64)
        for c in range(3):
            full[c][:, el] = torch.from_numpy(q[c])
            full[3 + c][:, el] = torch.from_numpy(R[c])
        full = full.cuda()
        dist.all_reduce(full)
        out = None
        if rank == 0:
            try:
                got = full.cpu().numpy()
                ctx = provision(mesh, N).get_DGContext()
                Hg, hg, _, _ = gaussian_hump(ctx.x, ctx.y)
                hug = np.ascontiguousarray(0.5 * hg * np.sin(3.0 * ctx.x + 1.0))
                hvg = np.ascontiguousarray(0.4 * hg * np.cos(2.5 * ctx.y))
                one = SW2dSolver(ctx, mode=args.mode, device=local)
                if args.variant:
                    one.set_variant(args.variant)
                one.set_bathymetry(Hg, np.zeros_like(Hg), np.zeros_like(Hg))
                one.set_state(hg, hug, hvg)
                R1 = one.computeRHS()
                one.step(5, dt)
                q1 = one.get_state()[:3]
                one.close()
                rel = lambda x, y: float(np.abs(x - y).max() / max(np.abs(y).max(), 1e-300))
                out = {"max_rel_err": max(rel(got[c], q1[c]) for c in range(3)), "max_rel_err_rhs": max(rel(got[3 + c], R1[c]) for c in range(3)),
                       "ranks": world, "elements": K, "steps": 5, "kernel_variant": var, "partition": pstats["method"], "neighbours": pstats["neighbours"],
                       "against": "the same 5 LSERK4 steps on ONE GPU through the same C ABI (rank 0)"}
            except Exception as exc:  # rank 0 alone is here: it must still reach the barrier below
                out = {"error": repr(exc), "ranks": world}
        dist.barrier()
        return out
    except Exception as exc:
        return {"error": repr(exc), "ranks": world}


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
