"""Device time of one weno::Eigen Reconstruct on BASELINE config 2's mesh (run under gpurun)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from minicfd_b200 import capi, fields, mesh as meshmod
from minicfd_b200.part import Part
from minicfd_b200.spatial import Spatial

m = meshmod.box(42, 42, 42, hi=(10.0, 10.0, 10.0))
sp = Spatial(Part(m), degree=2, riemann=capi.HLLC, device=0)
for tag in range(6):
    sp.SetInviscidWall(tag)
sp.ctx.set_stream(torch.cuda.current_stream().cuda_stream)
sp.SetLimiter(0.001, 1e-6, eigen=True)
f = fields.vortex()
def shocked(x):
    u = f(x, 0.0)
    u[x[..., 0] > 5.0] *= 1.3          # a jump: the detector marks cells, so the characteristic path runs
    return u
sp.Approximate(shocked)
for _ in range(3):
    sp.Reconstruct()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); e0.record()
for _ in range(10):
    sp.Reconstruct()
e1.record(); torch.cuda.synchronize()
print("weno::Eigen Reconstruct: %.3f ms" % (e0.elapsed_time(e1) / 10))
