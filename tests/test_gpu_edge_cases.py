"""Edge cases of the path through the C ABI against the oracle: a single cell whose six faces are all boundaries,
cell counts around the 32-cell tile (ragged last tile, exactly one tile, one cell more), parts with fewer local cells
than a tile and with no inner cell at all, and what the library answers to calls it cannot serve.
Tolerance 1e-12 relative."""
import numpy as np
import pytest

import cases
import oracle_check as oc
from minicfd_b200 import capi, mesh as meshmod
from minicfd_b200.part import Part, cell_points
from minicfd_b200.spatial import RungeKutta, Spatial

pytestmark = pytest.mark.gpu
TOL = 1e-12


def relerr(a, b):
    return np.abs(a - b).max() / np.abs(b).max()


def check_against_oracle(m, degree, scheme, oracle_scheme, riemann, viscous=False, limiter=False, **kw):
    o = oc.Oracle(m, degree, oracle_scheme, riemann, **({"c_next": kw["fr_c_next"]} if "fr_c_next" in kw else {}))
    sp = Spatial(Part(m), scheme=scheme, degree=degree, riemann=riemann, device=0,
                 diffusion=capi.NAVIER_STOKES_DDG if viscous else capi.INVISCID, **kw)
    if viscous:
        cases.apply_viscous_oracle(o)
        cases.apply_viscous_spatial(sp)
    else:
        cases.apply_bcs_oracle(o)
        cases.apply_bcs_spatial(sp, cases.MIXED_BCS)
    u = cases.noisy_state(cell_points(m.nodes, m.cells, degree), seed=3, amp=1e-2 if limiter else 1e-4)
    if limiter:
        o.set_limiter(*cases.WENO)
        sp.SetLimiter(*cases.WENO)
        expect = o.weno_lazy(u, *cases.WENO)
    else:
        expect = u
    sp.SetSolutionColumn(u)
    assert relerr(sp.GetSolutionColumn(), expect) < TOL
    sp.SetTime(0.2)
    assert relerr(sp.GetResidualColumn(), o.residual(expect, 0.2)) < TOL
    assert sp.GetTimeStep(1.0, 3) == pytest.approx(o.time_step(expect, 1.0, 3), rel=1e-13)
    RungeKutta(3).Update(sp, 0.0, 1e-3)
    assert relerr(sp.GetSolutionColumn(), o.rk_update(expect, 3, 0.0, 1e-3)) < TOL
    return sp


@pytest.mark.parametrize("degree", [2, 3])
@pytest.mark.parametrize("scheme,oracle_scheme,kw", [(capi.FR_LOBATTO, 0, {}), (capi.DG_LOBATTO, 1, {}),
                                                      (capi.FR_GENERAL, 2, {"fr_c_next": 0.4})])
def test_single_cell_all_boundary_kinds(degree, scheme, oracle_scheme, kw):
    m = meshmod.box(1, 1, 1, warp=0.0)
    check_against_oracle(m, degree, scheme, oracle_scheme, capi.HLLC, **kw)


def test_single_cell_viscous_limited_and_sampled():
    m = meshmod.box(1, 1, 1)
    check_against_oracle(m, 2, capi.FR_LOBATTO, 3, capi.AUSM, viscous=True)
    sp = check_against_oracle(m, 2, capi.FR_LOBATTO, 0, capi.HLLC, limiter=True)    # weno::Lazy without any neighbour
    points, values = sp.ctx.sample_vtk(True)
    assert points.shape == (1, 27, 3) and np.isfinite(values).all()
    assert bytes(sp.ctx.vtk_component_text(0)).count(b" ") == 27


@pytest.mark.parametrize("dims", [(31, 1, 1), (32, 1, 1), (33, 1, 1), (4, 4, 4), (13, 5, 1)])
def test_cell_counts_around_the_tile_size(dims):
    m = meshmod.scramble(meshmod.box(*dims, warp=0.1, seed=1), seed=2)
    check_against_oracle(m, 2, capi.FR_LOBATTO, 0, capi.HLLC)


def test_parts_smaller_than_a_tile_and_without_inner_cells():
    """A 6 x 1 x 1 row cut into three parts of two cells: the middle part has no inner cell, every part fewer local
    cells than a tile; driven through the halo buffers like the other two-part tests."""
    import torch
    from minicfd_b200.halo import _DeviceArray
    m = meshmod.box(6, 1, 1, warp=0.0)
    parts = (np.arange(6) // 2).astype(np.int32)
    o = oc.Oracle(m, 2, 0, 1)
    cases.apply_bcs_oracle(o)
    u = cases.noisy_state(cell_points(m.nodes, m.cells, 2), seed=5, amp=1e-4)
    expect = o.rk_update(u, 3, 0.0, 1e-3).reshape(-1, 135)
    P = [Part(m, parts, r, 3) for r in range(3)]
    assert P[1].n_inner == 0 and all(p.n_local == 2 for p in P)
    S = [Spatial(p, degree=2, riemann=capi.HLLC, device=0) for p in P]
    for sp, p in zip(S, P):
        cases.apply_bcs_spatial(sp, cases.MIXED_BCS)
        sp.SetSolutionColumn(p.gather_from_global(u, 135))
        sp.ctx.rk_begin(3, 0.0, 1e-3)
    bufs = [sum(sp.ctx.halo_buffers(), ()) for sp in S]
    for stage in range(3):
        for sp in S:
            sp.ctx.halo_pack()
        torch.cuda.synchronize()
        for a, p in enumerate(P):        # deliver every neighbour's slice of a's send buffer
            so = np.concatenate([[0], np.cumsum([len(c) for c in p.send_cells])]) * 135
            for i, b in enumerate(p.neighbours):
                q = P[b]
                j = list(q.neighbours).index(a)
                ro = np.concatenate([[0], np.cumsum([len(c) for c in q.recv_cells])]) * 135
                src = torch.as_tensor(_DeviceArray(bufs[a][0], bufs[a][1]), device="cuda")[so[i]:so[i + 1]]
                dst = torch.as_tensor(_DeviceArray(bufs[b][2], bufs[b][3]), device="cuda")[ro[j]:ro[j + 1]]
                dst.copy_(src)
        torch.cuda.synchronize()
        for sp in S:
            sp.ctx.rk_stage_inner(stage)
            sp.ctx.halo_unpack()
            sp.ctx.rk_stage_inter(stage)
    for sp, p in zip(S, P):
        assert relerr(sp.GetSolutionColumn().reshape(-1, 135), expect[p.local_global]) < TOL


def test_calls_the_library_cannot_serve_fail_loudly():
    m = cases.small_mesh()
    with pytest.raises(capi.McfdError):
        capi.Context(scheme=capi.FR_LOBATTO, degree=5, riemann=capi.HLLC, device=0)        # no such degree
    with pytest.raises(capi.McfdError):
        capi.Context(scheme=capi.FR_LOBATTO, degree=2, riemann=7, device=0)                # no such solver
    ctx = capi.Context(scheme=capi.FR_LOBATTO, degree=2, riemann=capi.HLLC, device=0)
    with pytest.raises(capi.McfdError):
        ctx.time_step(1.0, 3)                                                             # no mesh yet
    sp = Spatial(Part(m), degree=2, riemann=capi.HLLC, device=0)
    with pytest.raises(capi.McfdError):
        sp.ctx.rk_stage_inner(0)                                                          # no RK step open
    with pytest.raises(capi.McfdError):
        sp.ctx.set_diffusion(2.0, 1 / 12, 0.01, 0.7)                                      # context created inviscid
    with pytest.raises(capi.McfdError):
        sp.ctx.build_damping_matrices()
    with pytest.raises(capi.McfdError):
        sp.ctx.vtk_component_text(0)                                                      # nothing sampled yet
    bad = meshmod.HexMesh(nodes=m.nodes, cells=m.cells[:, [0, 3, 2, 1, 4, 7, 6, 5]].copy(), bfaces=m.bfaces, btags=m.btags)
    with pytest.raises(capi.McfdError):
        Spatial(Part(bad), degree=2, riemann=capi.HLLC, device=0)                          # inverted cells
