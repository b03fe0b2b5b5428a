"""Host enqueue time vs device time of bdg_sw2d_step under torchrun (is the multi-rank path launch-bound?)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
import bench
from blitzdg_b200.sw2d import SW2dSolver, nccl_unique_id

rank, world, local = bench.dist_env()
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
class A: pass
a = A(); a.shape = "tri"; a.nx = 1000; a.ny = 500; a.order = 4; a.mode = "simple"
ctx, halo, (H, h, hu, hv), Kg = bench.build_problem(a, rank, world)
sol = SW2dSolver(ctx, halo=halo, device=local)
if world > 1:
    idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        idt = torch.frombuffer(bytearray(nccl_unique_id()), dtype=torch.uint8).cuda()
    dist.broadcast(idt, 0)
    sol.comm_init(bytes(idt.cpu().numpy().tobytes()))
sol.set_state(h, hu, hv)
dt = 0.5 * sol.computeTimeStep() * (0.65 / 0.25)
for _ in range(3):
    sol.step(20, dt)
for rep in range(3):
    torch.cuda.synchronize(); sol.sync()
    if world > 1: dist.barrier()
    t0 = time.perf_counter(); sol.step(20, dt, sync=False); t1 = time.perf_counter(); sol.sync(); t2 = time.perf_counter()
    print("rank %d: enqueue %.2f ms, total %.2f ms for 100 stages; n_interior=%s K=%d launches=%d" % (
        rank, 1e3 * (t1 - t0), 1e3 * (t2 - t0), halo["n_interior"] if halo else None, ctx.K, sol.launch_count), flush=True)
sol.close()
if world > 1: dist.destroy_process_group()
