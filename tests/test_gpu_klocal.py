"""GPU parity tests of the kLocal == true expansions (mini::polynomial::Hexahedron<Gx, Gy, Gz, K, true>): columns hold
det(J) u, as the reference's laminar drivers use them for FR (demo/laminar/shockless.hpp:66-69, BASELINE config 4).
Through the C ABI, against the reference's golden vectors and the oracle.  Tolerance 1e-12 relative."""
import os

import numpy as np
import pytest

import cases
import oracle_check as oc
from minicfd_b200 import capi, fields, mesh as meshmod
from minicfd_b200.part import Part, cell_points
from minicfd_b200.spatial import RungeKutta, Spatial

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL = 1e-12


def relerr(a, b):
    return np.abs(a - b).max() / np.abs(b).max()


def local_spatial(m, degree, riemann, viscous, **kw):
    sp = Spatial(Part(m), scheme=capi.FR_LOBATTO, degree=degree, riemann=riemann, device=0, k_local=True,
                 diffusion=capi.NAVIER_STOKES_DDG if viscous else capi.INVISCID, **kw)
    if viscous:
        cases.apply_viscous_spatial(sp)
    else:
        cases.apply_bcs_spatial(sp, cases.MIXED_BCS)
    return sp


@pytest.mark.parametrize("name,degree,riemann,viscous", cases.LOCAL_GOLDEN_CASES)
def test_klocal_golden(name, degree, riemann, viscous):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    m = meshmod.HexMesh(nodes=g["nodes"], cells=g["cells"], bfaces=g["bfaces"], btags=g["btags"])
    sp = local_spatial(m, degree, riemann, viscous)
    assert relerr(sp.ctx.jacobian_determinants(), g["det"]) < 1e-14
    sp.SetSolutionColumn(g["u"])                        # det(J) u in ...
    assert relerr(sp.GetSolutionColumn(), g["u"]) < 1e-15   # ... and out again
    sp.SetTime(cases.GOLDEN_T)
    assert relerr(sp.GetResidualColumn(), g["residual"]) < TOL
    assert sp.GetTimeStep(1.0, 3) == pytest.approx(float(g["dt"]), rel=1e-13)
    RungeKutta(3).Update(sp, 0.0, cases.GOLDEN_DT)
    assert relerr(sp.GetSolutionColumn(), g["u_rk3"]) < TOL


@pytest.mark.parametrize("degree,riemann", [(2, capi.AUSM), (3, capi.HLLC)])
def test_klocal_navier_stokes_vs_oracle(degree, riemann):
    m = cases.small_mesh(seed=17)
    n = (degree + 1) ** 3
    o = oc.Oracle(m, degree, 3, riemann)
    cases.apply_viscous_oracle(o)
    det = o.jacobian_determinants()
    o.set_local(True)
    u = cases.noisy_state(cell_points(m.nodes, m.cells, degree), seed=6, amp=1e-4).reshape(-1, n, 5)
    column = (u * det[:, :, None]).reshape(-1)
    sp = local_spatial(m, degree, riemann, True)
    sp.SetSolutionColumn(column)
    sp.SetTime(0.2)
    assert relerr(sp.GetResidualColumn(), o.residual(column, 0.2)) < TOL
    for order in (1, 2, 3):
        sp.SetSolutionColumn(column)
        RungeKutta(order).Update(sp, 0.0, 1e-3)
        assert relerr(sp.GetSolutionColumn(), o.rk_update(column, order, 0.0, 1e-3)) < TOL
    sp.SetSolutionColumn(column)
    RungeKutta(3).UpdateGeneric(sp, 0.0, 1e-3)          # rk.hpp's column algebra on det(J) u columns
    assert relerr(sp.GetSolutionColumn(), o.rk_update(column, 3, 0.0, 1e-3)) < TOL


def test_klocal_approximate_and_euler_equivalence():
    """Approximate stores value * det J (Hexahedron::SetValue); without viscosity the two expansions are the same
    scheme, so the kLocal column is det J times the other one's, residuals included."""
    m = meshmod.scramble(meshmod.box(4, 3, 3, warp=0.15, seed=2), seed=2)
    a = local_spatial(m, 2, capi.HLLC, False)
    b = Spatial(Part(m), degree=2, riemann=capi.HLLC, device=0)
    cases.apply_bcs_spatial(b, cases.MIXED_BCS)
    f = lambda x: fields.smooth(1.0)(x, 0.0)
    a.Approximate(f)
    b.Approximate(f)
    det = a.ctx.jacobian_determinants()
    ua, ub = a.GetSolutionColumn().reshape(-1, 27, 5), b.GetSolutionColumn().reshape(-1, 27, 5)
    assert relerr(ua, ub * det[:, :, None]) < 1e-15
    ra, rb = a.GetResidualColumn().reshape(-1, 27, 5), b.GetResidualColumn().reshape(-1, 27, 5)
    assert relerr(ra, rb * det[:, :, None]) < 1e-14


def test_klocal_navier_stokes_two_parts_with_halo_buffers():
    import torch
    from minicfd_b200.halo import _DeviceArray
    m = meshmod.scramble(meshmod.box(6, 4, 2, warp=0.1, seed=2), seed=3)
    parts = (np.arange(m.n_cells) * 2 // m.n_cells).astype(np.int32)
    o = oc.Oracle(m, 2, 3, 1)
    cases.apply_viscous_oracle(o)
    det = o.jacobian_determinants()
    o.set_local(True)
    u = cases.noisy_state(cell_points(m.nodes, m.cells, 2), seed=9, amp=1e-4).reshape(-1, 27, 5)
    column = (u * det[:, :, None]).reshape(-1)
    expect = o.rk_update(column, 3, 0.0, 1e-3).reshape(-1, 135)
    P = [Part(m, parts, r, 2) for r in range(2)]
    S = [Spatial(p, degree=2, riemann=capi.HLLC, device=0, diffusion=capi.NAVIER_STOKES_DDG, k_local=True) for p in P]
    for sp in S:
        cases.apply_viscous_spatial(sp)
    bufs = [sum(sp.ctx.halo_buffers(), ()) for sp in S]
    for sp, p in zip(S, P):
        sp.SetSolutionColumn(p.gather_from_global(column, 135))
    for sp in S:
        sp.ctx.rk_begin(3, 0.0, 1e-3)
    for stage in range(3):
        for sp in S:
            sp.ctx.halo_pack()
        torch.cuda.synchronize()
        for a, b in ((0, 1), (1, 0)):
            src = torch.as_tensor(_DeviceArray(bufs[a][0], bufs[a][1]), device="cuda")
            dst = torch.as_tensor(_DeviceArray(bufs[b][2], bufs[b][3]), device="cuda")
            dst.copy_(src)
        torch.cuda.synchronize()
        for sp in S:
            sp.ctx.rk_stage_inner(stage)
            sp.ctx.halo_unpack()
            sp.ctx.rk_stage_inter(stage)
    for sp, p in zip(S, P):
        assert relerr(sp.GetSolutionColumn().reshape(-1, 135), expect[p.local_global]) < TOL
