"""Ghost exchange and time-step minimum through peer-mapped memory (mcfd_peer_*, csrc/peer.cuh): several parts held by
several contexts of this process on one GPU, wired with plain pointers — the same kernels, flags and stream choreography
as one process per GPU over CUDA IPC (tests/mgpu_worker.py `peer`, run at 2 and 8 GPUs: profiles/)."""
import numpy as np
import pytest

import cases
import oracle_check as oc
from minicfd_b200 import capi, mesh as meshmod, peer
from minicfd_b200.part import Part, cell_points
from minicfd_b200.spatial import Spatial

pytestmark = pytest.mark.gpu


def relerr(a, b):
    return float(np.abs(np.asarray(a) - np.asarray(b)).max() / np.abs(np.asarray(b)).max())


@pytest.mark.parametrize("degree,world,riemann", [(2, 2, capi.HLLC), (2, 4, capi.AUSM), (3, 2, capi.HLLC), (2, 3, capi.HLLC)])
def test_peer_step_matches_single_part_oracle(degree, world, riemann):
    """mcfd_peer_step on N parts = GetTimeStep + min over the parts + RK3 on the whole mesh (oracle), several steps:
    covers the push / wait epochs, buffer rotation over steps, and the time-step slots' parity."""
    import torch
    m = meshmod.scramble(meshmod.box(6, 4, 3, warp=0.1, seed=2), seed=3)
    parts = (np.arange(m.n_cells) * world // m.n_cells).astype(np.int32)
    kf = 5 * (degree + 1) ** 3
    u = cases.noisy_state(cell_points(m.nodes, m.cells, degree), seed=9)
    o = oc.Oracle(m, degree, 0, riemann)
    for tag in range(6):
        o.set_bc(tag, 0)
    P = [Part(m, parts, r, world) for r in range(world)]
    S = [Spatial(p, degree=degree, riemann=riemann, device=0) for p in P]
    streams = [torch.cuda.Stream() for _ in S]
    for sp, p, st in zip(S, P, streams):
        for tag in range(6):
            sp.SetInviscidWall(tag)
        sp.ctx.set_stream(st.cuda_stream)
        sp.SetSolutionColumn(p.gather_from_global(u, kf))
    peer.connect_local(S, P)
    ref, t = u.copy(), 0.0
    for step in range(4):
        capi.peer_step_group([sp.ctx for sp in S], 3, t, 2e-3)     # nothing waits for the host
        dt = o.time_step(ref, 2e-3, 3)
        ref = o.rk_update(ref, 3, t, dt)
        t += dt
    for sp in S:
        sp.ctx.peer_check()
    got = np.zeros(m.n_cells * kf)
    for sp, p in zip(S, P):
        p.scatter_to_global(sp.GetSolutionColumn(), kf, got)
    assert relerr(got, ref) < 1e-12


def test_peer_calls_need_a_connection():
    m = cases.small_mesh(seed=4)
    sp = Spatial(Part(m), degree=2, riemann=capi.HLLC, device=0)
    with pytest.raises(capi.McfdError):
        sp.ctx.peer_push()
    with pytest.raises(capi.McfdError):
        sp.ctx.peer_step(3, 0.0, 1e-3)
