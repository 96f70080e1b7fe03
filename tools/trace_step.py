"""Kernel timeline of a few steps on rank 0 (torch.profiler / CUPTI).  Run under torchrun for N > 1."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, torch.distributed as dist
from torch.profiler import profile, ProfilerActivity
from minicfd_b200 import capi, fields, halo, mesh as meshmod
from minicfd_b200.part import Part
from minicfd_b200.spatial import RungeKutta, Spatial, time_step_device

world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if world > 1:
    os.environ.setdefault("TORCH_NCCL_HIGH_PRIORITY", "1"); dist.init_process_group("nccl", device_id=torch.device("cuda", local))
px, py, pz = meshmod.factor3(world)
n = 42
m = meshmod.box(n * px, n * py, n * pz, hi=(10.0 * px, 10.0 * py, 10.0 * pz), periodic=True)
parts = meshmod.block_partition(m, px, py, pz) if world > 1 else None
part = Part(m, parts, rank, world)
comm = halo.for_part(part, 135) if world > 1 else None
sp = Spatial(part, degree=2, riemann=capi.HLLC, device=local, comm=comm)
sp.ctx.set_stream(torch.cuda.current_stream().cuda_stream)
sp.Approximate(lambda x: fields.vortex()(x, 0.0))
if comm: comm.exchange()
rk = RungeKutta(3); dt_dev = torch.zeros(1, dtype=torch.float64, device="cuda")
use_peer = world > 1 and "--nccl" not in sys.argv
if use_peer:
    from minicfd_b200 import peer
    peer.connect(sp, part)
def step():
    if use_peer:
        sp.ctx.peer_step(3, 0.0, 1.0)
    else:
        time_step_device(sp, 1.0, 3, dt_dev); rk.UpdateDevice(sp, 0.0, dt_dev.data_ptr())
for _ in range(5): step()
torch.cuda.synchronize()
if world > 1: dist.barrier()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3): step()
    torch.cuda.synchronize()
if rank == 0:
    ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    ev.sort(key=lambda e: e.time_range.start)
    t0 = ev[0].time_range.start
    for e in ev:
        print(f"{e.time_range.start - t0:9.1f} us  +{e.time_range.elapsed_us():8.1f}  stream {getattr(e, 'stream', '?')}  {e.name[:70]}")
if world > 1:
    comm.drain(); torch.cuda.synchronize(); dist.destroy_process_group()
