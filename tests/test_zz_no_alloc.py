"""mcfd_time_step_static: the time step for hosts that must not allocate device memory while running
(the whole GPU handed to their own pool, a fixed memory budget, or a loop that cannot afford the implicit
synchronisation of cudaMalloc).  The test plays such a host: every buffer is allocated up front, the entry
point is called through the C ABI with raw device pointers, and its answer is checked against the existing
mcfd_time_step on two calls in a row with different states (so nothing left in the library's fixed scratch
array by the first call can leak into the second).  Collected last on purpose."""
import gc

import numpy as np
import pytest

import cases
from minicfd_b200 import capi, mesh as meshmod
from minicfd_b200.part import Part, cell_points
from minicfd_b200.spatial import Spatial

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("degree,scheme", [(2, capi.FR_LOBATTO), (3, capi.DG_LOBATTO)])
def test_static_time_step_matches_allocating_entry_point(degree, scheme):
    import torch
    m = meshmod.scramble(meshmod.box(7, 5, 4, warp=0.15, seed=2), seed=6)
    sp = Spatial(Part(m), scheme=scheme, degree=degree, riemann=capi.HLLC, device=0)
    n_cells = m.n_cells
    length, _ = sp.ctx.cell_geometry()
    pts = cell_points(m.nodes, m.cells, degree)
    # the caller's own device memory, allocated once before the "run"
    column_dev = torch.empty(n_cells * sp.kfields, dtype=torch.float64, device="cuda")
    length_dev = torch.from_numpy(length).cuda()
    dt_dev = torch.zeros(1, dtype=torch.float64, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    for call, (seed, amp, dt_guess) in enumerate([(1, 1e-3, 1.0), (2, 0.3, 1.0), (3, 1e-3, 1e-4)]):
        u = cases.noisy_state(pts, seed=seed, amp=amp)
        sp.SetSolutionColumn(u)
        expect = sp.ctx.time_step(dt_guess, 3)                 # existing entry point
        column_dev.copy_(torch.from_numpy(u))
        if call == 0:     # the first call on a device looks the scratch array's address up (module load included)
            capi.time_step_static(scheme, degree, n_cells, column_dev.data_ptr(), length_dev.data_ptr(), dt_guess, 3,
                                  dt_dev.data_ptr(), stream)
            torch.cuda.synchronize()
        # what the driver itself reports as free: sees a cudaMalloc / cudaMallocAsync inside the library, which torch's
        # own counters would not
        gc.collect()      # contexts of earlier tests free their device memory when collected: not during the measurement
        torch.cuda.synchronize()
        free_before, allocated_before = torch.cuda.mem_get_info()[0], torch.cuda.memory_allocated()
        capi.time_step_static(scheme, degree, n_cells, column_dev.data_ptr(), length_dev.data_ptr(), dt_guess, 3,
                              dt_dev.data_ptr(), stream)
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] >= free_before and torch.cuda.memory_allocated() == allocated_before
        got = float(dt_dev.item())
        assert got == pytest.approx(expect, rel=1e-13), (call, got, expect)


def test_static_time_step_rejects_oversized_parts():
    import torch
    dummy = torch.zeros(8, dtype=torch.float64, device="cuda")
    with pytest.raises(capi.McfdError, match="exceeds the fixed scratch"):
        capi.time_step_static(capi.FR_LOBATTO, 2, (1 << 20) + 1, dummy.data_ptr(), dummy.data_ptr(), 1.0, 3,
                              dummy.data_ptr(), 0)
