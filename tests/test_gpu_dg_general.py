"""GPU parity tests of dg::General on the modal Projection polynomial with Gauss-Legendre rules,
through the C ABI, against the golden vectors of the reference and the oracle.  Tolerance 1e-12 relative."""
import os

import numpy as np
import pytest

import cases
import oracle_check as oc
from minicfd_b200 import capi, fields, mesh as meshmod
from minicfd_b200.part import Part
from minicfd_b200.spatial import RungeKutta, Spatial

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL = 1e-12


def relerr(a, b):
    return np.abs(a - b).max() / np.abs(b).max()


@pytest.mark.parametrize("name,degree,riemann", cases.DGM_GOLDEN_CASES)
def test_dg_general_golden(name, degree, riemann):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    m = meshmod.HexMesh(nodes=g["nodes"], cells=g["cells"], bfaces=g["bfaces"], btags=g["btags"])
    sp = Spatial(Part(m), scheme=capi.DG_GENERAL, degree=degree, riemann=riemann, device=0)
    assert sp.kfields == 5 * {2: 10, 3: 20}[degree]
    cases.apply_bcs_spatial(sp, cases.MIXED_BCS)
    sp.Approximate(lambda x: fields.smooth(1.0)(x, 0.0))         # Projection::Approximate
    assert relerr(sp.GetSolutionColumn(), g["projected"]) < TOL
    sp.SetSolutionColumn(g["u"])
    sp.SetTime(cases.GOLDEN_T)
    assert relerr(sp.GetResidualColumn(), g["residual"]) < TOL
    assert sp.GetTimeStep(1.0, 3) == pytest.approx(float(g["dt"]), rel=1e-13)
    RungeKutta(3).Update(sp, 0.0, cases.GOLDEN_DT)
    assert relerr(sp.GetSolutionColumn(), g["u_rk3"]) < TOL


@pytest.mark.parametrize("degree", [2, 3])
@pytest.mark.parametrize("riemann", [capi.EXACT, capi.HLLC, capi.AUSM])
def test_dg_general_vs_oracle(degree, riemann):
    m = cases.small_mesh(seed=21)
    o = oc.Oracle(m, degree, 9, riemann)
    cases.apply_bcs_oracle(o)
    u = o.approximate(oc.F_SMOOTH, [1.0])
    u = u * (1 + 1e-3 * np.random.default_rng(1).uniform(-1, 1, u.shape))
    sp = Spatial(Part(m), scheme=capi.DG_GENERAL, degree=degree, riemann=riemann, device=0)
    cases.apply_bcs_spatial(sp, cases.MIXED_BCS)
    sp.SetSolutionColumn(u)
    sp.SetTime(0.2)
    assert relerr(sp.GetResidualColumn(), o.residual(u, 0.2)) < TOL
    for order in (1, 2, 3):
        sp.SetSolutionColumn(u)
        RungeKutta(order).Update(sp, 0.0, 2e-3)
        assert relerr(sp.GetSolutionColumn(), o.rk_update(u, order, 0.0, 2e-3)) < TOL
    sp.SetSolutionColumn(u)
    RungeKutta(3).UpdateGeneric(sp, 0.0, 2e-3)
    assert relerr(sp.GetSolutionColumn(), o.rk_update(u, 3, 0.0, 2e-3)) < TOL


def test_dg_general_conserves_on_a_periodic_box_and_keeps_free_stream():
    m = meshmod.box(6, 5, 4, hi=(6.0, 5.0, 4.0), periodic=True)
    sp = Spatial(Part(m), scheme=capi.DG_GENERAL, degree=2, riemann=capi.HLLC, device=0)
    sp.Approximate(lambda x: fields.vortex(x0=3.0, y0=2.5)(x, 0.0))
    res = sp.GetResidualColumn().reshape(m.n_cells, 10, 5)
    # the first mode is constant (phi_0 = 1/sqrt(V)): sum_cells sqrt(V) R_0 is d/dt of the integral of u
    _, vol = sp.ctx.cell_geometry()
    total = (res[:, 0, :] * np.sqrt(vol)[:, None]).sum(0)
    assert np.abs(total).max() < 1e-12 * np.abs(res).max() * m.n_cells
    mw = meshmod.scramble(meshmod.box(4, 4, 3, warp=0.12, seed=5), seed=5)
    spw = Spatial(Part(mw), scheme=capi.DG_GENERAL, degree=3, riemann=capi.AUSM, device=0)
    state = np.array([1.2, 0.3, -0.2, 0.1, 2.5])
    for tag in range(6):
        spw.SetSmartBoundary(tag, fields.constant(state))
    spw.Approximate(lambda x: fields.constant(state)(x, 0.0))
    assert np.abs(spw.GetResidualColumn()).max() < 1e-11


def test_dg_general_two_parts_with_halo_buffers():
    import torch
    from minicfd_b200.halo import _DeviceArray
    m = meshmod.scramble(meshmod.box(6, 4, 2, warp=0.1, seed=2), seed=3)
    parts = (np.arange(m.n_cells) * 2 // m.n_cells).astype(np.int32)
    o = oc.Oracle(m, 2, 9, 1)
    for tag in range(6):
        o.set_bc(tag, 0)
    u = o.approximate(oc.F_SMOOTH, [1.0])
    u = u * (1 + 1e-3 * np.random.default_rng(2).uniform(-1, 1, u.shape))
    expect_res = o.residual(u, 0.0).reshape(-1, 50)
    expect_u = o.rk_update(u, 3, 0.0, 1e-3).reshape(-1, 50)
    P = [Part(m, parts, r, 2) for r in range(2)]
    S = [Spatial(p, scheme=capi.DG_GENERAL, degree=2, riemann=capi.HLLC, device=0) for p in P]
    for sp in S:
        for tag in range(6):
            sp.SetInviscidWall(tag)
    bufs = [sum(sp.ctx.halo_buffers(), ()) for sp in S]

    def exchange():
        for sp in S:
            sp.ctx.halo_pack()
        torch.cuda.synchronize()
        for a, b in ((0, 1), (1, 0)):
            src = torch.as_tensor(_DeviceArray(bufs[a][0], bufs[a][1]), device="cuda")
            dst = torch.as_tensor(_DeviceArray(bufs[b][2], bufs[b][3]), device="cuda")
            dst.copy_(src)
        torch.cuda.synchronize()

    for sp, p in zip(S, P):
        sp.SetSolutionColumn(p.gather_from_global(u, 50))
    exchange()
    for sp, p in zip(S, P):
        sp.ctx.halo_unpack()
        assert relerr(sp.ctx.residual(0.0).reshape(-1, 50), expect_res[p.local_global]) < TOL
    for sp in S:
        sp.ctx.rk_begin(3, 0.0, 1e-3)
    for stage in range(3):
        exchange()
        for sp in S:
            sp.ctx.rk_stage_inner(stage)
            sp.ctx.halo_unpack()
            sp.ctx.rk_stage_inter(stage)
    for sp, p in zip(S, P):
        assert relerr(sp.GetSolutionColumn().reshape(-1, 50), expect_u[p.local_global]) < TOL


@pytest.mark.parametrize("name,degree,riemann", cases.DGM_WENO_GOLDEN_CASES)
def test_dg_general_with_weno_lazy_golden(name, degree, riemann):
    """WithLimiter<dg::General<Part<Projection>>, weno::Lazy>: the reference's DGFEM + LIMITER build."""
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    m = meshmod.HexMesh(nodes=g["nodes"], cells=g["cells"], bfaces=g["bfaces"], btags=g["btags"])
    sp = Spatial(Part(m), scheme=capi.DG_GENERAL, degree=degree, riemann=riemann, device=0)
    cases.apply_bcs_spatial(sp, cases.MIXED_BCS)
    sp.SetLimiter(*cases.WENO)
    sp.SetSolutionColumn(g["u"])
    assert relerr(sp.GetSolutionColumn(), g["limited"]) < TOL
    RungeKutta(3).Update(sp, 0.0, cases.GOLDEN_DT)
    assert relerr(sp.GetSolutionColumn(), g["u_rk3"]) < TOL


def test_dg_general_weno_two_parts_match_the_reference_parts():
    import refapi
    import torch
    from minicfd_b200.halo import _DeviceArray
    if not refapi.available(2):
        pytest.skip("oracle/_ref not shipped")
    m = meshmod.scramble(meshmod.box(6, 4, 2, warp=0.1, seed=2), seed=3)
    parts = (np.arange(m.n_cells) * 2 // m.n_cells).astype(np.int32)
    o = oc.Oracle(m, 2, 9, 1)
    u = o.approximate(oc.F_SMOOTH, [1.0])
    u = u * (1 + 5e-2 * np.random.default_rng(2).uniform(-1, 1, u.shape))
    r = refapi.Ref(m, degree=2, scheme=refapi.DG_GENERAL_WENO_LAZY, riemann=1, cell_parts=parts, n_parts=2, weno=cases.WENO)
    P = [Part(m, parts, k, 2) for k in range(2)]
    r.set_solution_all([p.gather_from_global(u, 50) for p in P])
    S = [Spatial(p, scheme=capi.DG_GENERAL, degree=2, riemann=capi.HLLC, device=0) for p in P]
    bufs = [sum(sp.ctx.halo_buffers(), ()) for sp in S]
    for sp, p in zip(S, P):
        sp.ctx.set_solution(p.gather_from_global(u, 50))
    for sp in S:
        sp.ctx.halo_pack()
    torch.cuda.synchronize()
    for a, b in ((0, 1), (1, 0)):
        src = torch.as_tensor(_DeviceArray(bufs[a][0], bufs[a][1]), device="cuda")
        dst = torch.as_tensor(_DeviceArray(bufs[b][2], bufs[b][3]), device="cuda")
        dst.copy_(src)
    torch.cuda.synchronize()
    for sp in S:
        sp.ctx.limiter_weno_lazy(*cases.WENO, 0)
        sp.ctx.halo_unpack()
        sp.ctx.limiter_weno_lazy(*cases.WENO, 1)
    for k, sp in enumerate(S):
        assert relerr(sp.GetSolutionColumn(), r.get_solution(k)) < TOL
