"""GPU parity tests of the viscous path: fr::Lobatto on
ConvectionDiffusion<rotated::Euler<..>, Direct<NavierStokes<Gas>>> through the C ABI, against the oracle.
Tolerance 1e-12 relative (max norm)."""
import os

import numpy as np
import pytest

import cases
import oracle_check as oc
import refapi
from minicfd_b200 import capi, fields, mesh as meshmod
from minicfd_b200.part import Part, cell_points
from minicfd_b200.spatial import RungeKutta, Spatial

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL = 1e-12


def relerr(a, b):
    return np.abs(a - b).max() / np.abs(b).max()


def viscous_spatial(m, degree, riemann, **kw):
    return Spatial(Part(m), scheme=capi.FR_LOBATTO, degree=degree, riemann=riemann, device=0,
                   diffusion=capi.NAVIER_STOKES_DDG, **kw)


@pytest.mark.parametrize("name,degree,riemann", cases.VISCOUS_GOLDEN_CASES)
def test_navier_stokes_golden(name, degree, riemann):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    m = meshmod.HexMesh(nodes=g["nodes"], cells=g["cells"], bfaces=g["bfaces"], btags=g["btags"])
    sp = viscous_spatial(m, degree, riemann)
    cases.apply_viscous_spatial(sp)
    sp.SetSolutionColumn(g["u"])
    sp.SetTime(cases.GOLDEN_T)
    assert relerr(sp.GetResidualColumn(), g["residual"]) < TOL
    assert sp.GetTimeStep(1.0, 3) == pytest.approx(float(g["dt"]), rel=1e-13)
    RungeKutta(3).Update(sp, 0.0, cases.GOLDEN_DT)
    assert relerr(sp.GetSolutionColumn(), g["u_rk3"]) < TOL


@pytest.mark.parametrize("degree", [2, 3])
@pytest.mark.parametrize("riemann", [capi.EXACT, capi.HLLC, capi.AUSM])
def test_navier_stokes_vs_oracle(degree, riemann):
    m = cases.small_mesh(seed=12)
    vis = (1.5, 0.1, 0.03, 0.72)
    u = cases.noisy_state(cell_points(m.nodes, m.cells, degree), seed=4, amp=1e-4)
    o = oc.Oracle(m, degree, 3, riemann)
    cases.apply_viscous_oracle(o, vis=vis)
    sp = viscous_spatial(m, degree, riemann)
    cases.apply_viscous_spatial(sp, vis=vis)
    sp.SetSolutionColumn(u)
    sp.SetTime(0.2)
    assert relerr(sp.GetResidualColumn(), o.residual(u, 0.2)) < TOL
    for order in (1, 2, 3):
        sp.SetSolutionColumn(u)
        RungeKutta(order).Update(sp, 0.0, 1e-3)
        assert relerr(sp.GetSolutionColumn(), o.rk_update(u, order, 0.0, 1e-3)) < TOL
    # the column algebra of rk.hpp through GetResidualColumn gives the same state as the fused stages
    sp.SetSolutionColumn(u)
    RungeKutta(3).UpdateGeneric(sp, 0.0, 1e-3)
    assert relerr(sp.GetSolutionColumn(), o.rk_update(u, 3, 0.0, 1e-3)) < TOL


@pytest.mark.skipif(not refapi.available(2), reason="oracle/_ref not shipped")
def test_navier_stokes_against_reference_binary_many_steps():
    m = cases.small_mesh(seed=5)
    u = cases.noisy_state(cell_points(m.nodes, m.cells, 2), seed=1, amp=1e-4)
    r = refapi.Ref(m, degree=2, scheme=refapi.FR_VISCOUS, riemann=refapi.HLLC)
    cases.apply_viscous_ref(r)
    r.set_solution(u)
    sp = viscous_spatial(m, 2, capi.HLLC)
    cases.apply_viscous_spatial(sp)
    sp.SetSolutionColumn(u)
    t, dt = 0.0, 1e-3
    for _ in range(10):
        RungeKutta(3).Update(sp, t, dt)
        t += dt
    r.rk_update(3, 0.0, dt, 10)
    assert relerr(sp.GetSolutionColumn(), r.get_solution()) < 1e-12


def test_zero_viscosity_is_the_euler_path():
    """With nu = 0 and no no-slip wall every viscous term is an exact zero; the two kernel instantiations may
    still contract their FMAs differently, hence a rounding-level bound instead of bit equality."""
    m = meshmod.scramble(meshmod.box(5, 4, 3, warp=0.1, seed=3), seed=3)
    u = cases.noisy_state(cell_points(m.nodes, m.cells, 2), seed=2)
    a = Spatial(Part(m), degree=2, riemann=capi.HLLC, device=0)
    b = viscous_spatial(m, 2, capi.HLLC)
    b.SetDiffusion(2.0, 1.0 / 12, 0.0, 0.708)
    for sp in (a, b):
        cases.apply_bcs_spatial(sp, cases.MIXED_BCS)
        sp.SetSolutionColumn(u)
        sp.SetTime(0.1)
    ra, rb = a.GetResidualColumn(), b.GetResidualColumn()
    assert np.abs(ra - rb).max() <= 1e-14 * np.abs(ra).max()


def test_viscous_free_stream_and_couette_like_shear():
    """Uniform flow has zero residual on a warped mesh (gradients and Hessians of a constant vanish up to
    rounding); a linear shear profile u(y) produces no net viscous force in the interior (tau is constant)."""
    m = meshmod.scramble(meshmod.box(4, 4, 4, warp=0.12, seed=7, periodic=False), seed=7)
    sp = viscous_spatial(m, 2, capi.HLLC)
    sp.SetDiffusion(2.0, 1.0 / 12, 0.05, 0.708)
    state = np.array([1.2, 0.3, -0.2, 0.1, 2.5])
    for tag in range(6):
        sp.SetSmartBoundary(tag, fields.constant(state))
    sp.Approximate(lambda x: fields.constant(state)(x, 0.0))
    res = sp.GetResidualColumn()
    assert np.abs(res).max() < 1e-12


def test_navier_stokes_two_parts_with_halo_buffers():
    """Two parts on one GPU: the ghost cells' traces (gradient, Hessian . n) are rebuilt from the unpacked
    coefficients, as the reference rebuilds them from the ghost cells' polynomials."""
    import torch
    from minicfd_b200.halo import _DeviceArray
    m = meshmod.scramble(meshmod.box(6, 4, 2, warp=0.1, seed=2), seed=3)
    parts = (np.arange(m.n_cells) * 2 // m.n_cells).astype(np.int32)
    u = cases.noisy_state(cell_points(m.nodes, m.cells, 2), seed=9, amp=1e-4)
    o = oc.Oracle(m, 2, 3, 1)
    cases.apply_viscous_oracle(o)
    expect_res = o.residual(u, 0.0).reshape(-1, 135)
    expect_u = o.rk_update(u, 3, 0.0, 1e-3).reshape(-1, 135)
    P = [Part(m, parts, r, 2) for r in range(2)]
    S = [Spatial(p, degree=2, riemann=capi.HLLC, device=0, diffusion=capi.NAVIER_STOKES_DDG) for p in P]
    for sp in S:
        cases.apply_viscous_spatial(sp)
    bufs = []
    for sp in S:
        (s_ptr, s_n), (r_ptr, r_n) = sp.ctx.halo_buffers()
        bufs.append((s_ptr, s_n, r_ptr, r_n))

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
        sp.SetSolutionColumn(p.gather_from_global(u, 135))
    exchange()
    for sp, p in zip(S, P):
        sp.ctx.halo_unpack()
        assert relerr(sp.ctx.residual(0.0).reshape(-1, 135), expect_res[p.local_global]) < TOL
    for sp in S:
        sp.ctx.rk_begin(3, 0.0, 1e-3)
    for stage in range(3):
        exchange()
        for sp in S:
            sp.ctx.rk_stage_inner(stage)
            sp.ctx.halo_unpack()
            sp.ctx.rk_stage_inter(stage)
    for sp, p in zip(S, P):
        assert relerr(sp.GetSolutionColumn().reshape(-1, 135), expect_u[p.local_global]) < TOL


# ---- component-wise isotropic diffusion, one viscosity per cell and component (flux path of SURVEY row a24) ----
@pytest.mark.parametrize("name,degree,riemann", cases.ISO_GOLDEN_CASES)
def test_cell_viscosity_golden_and_table(name, degree, riemann):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    m = meshmod.HexMesh(nodes=g["nodes"], cells=g["cells"], bfaces=g["bfaces"], btags=g["btags"])
    sp = Spatial(Part(m), degree=degree, riemann=riemann, device=0, diffusion=capi.CELL_VISCOSITY_DDG)
    cases.apply_viscous_spatial(sp, vis=cases.ISO_VISCOUS)      # Isotropic::SetProperty(nu): uniform
    sp.SetSolutionColumn(g["u"])
    sp.SetTime(cases.GOLDEN_T)
    assert relerr(sp.GetResidualColumn(), g["residual"]) < TOL
    RungeKutta(3).Update(sp, 0.0, cases.GOLDEN_DT)
    assert relerr(sp.GetSolutionColumn(), g["u_rk3"]) < TOL
    # EnergyBasedViscosity::properties_: a table with zeros in it
    nu = cases.cell_viscosities(m.n_cells)
    o = oc.Oracle(m, degree, 6, riemann)
    cases.apply_viscous_oracle(o, vis=cases.ISO_VISCOUS)
    o.set_cell_viscosity(nu)
    sp.SetCellViscosity(nu)
    sp.SetSolutionColumn(g["u"])
    sp.SetTime(cases.GOLDEN_T)
    assert relerr(sp.GetResidualColumn(), o.residual(g["u"], cases.GOLDEN_T)) < TOL
    for order in (1, 2, 3):
        sp.SetSolutionColumn(g["u"])
        RungeKutta(order).Update(sp, 0.0, 1e-3)
        assert relerr(sp.GetSolutionColumn(), o.rk_update(g["u"], order, 0.0, 1e-3)) < TOL


# ---- EnergyBasedViscosity + WithViscosity -------------------------------------------------------
@pytest.mark.parametrize("name,degree,riemann", cases.ENERGY_GOLDEN_CASES)
def test_energy_based_viscosity_golden(name, degree, riemann):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    m = meshmod.HexMesh(nodes=g["nodes"], cells=g["cells"], bfaces=g["bfaces"], btags=g["btags"])
    sp = Spatial(Part(m), degree=degree, riemann=riemann, device=0, diffusion=capi.CELL_VISCOSITY_DDG)
    cases.apply_bcs_spatial(sp, cases.MIXED_BCS)
    sp.SetDiffusion(*cases.ENERGY_BETAS, 0.0, 0.708)           # Direct::SetBetaValues before the spatial object exists
    sp.EnableEnergyBasedViscosity()                            # WithViscosity(...): Viscosity::Initialize
    assert relerr(sp.ctx.damping_matrices()[:4], g["damping"]) < TOL
    sp.SetSolutionColumn(g["u"])                               # -> UpdateProperties
    nu = sp.ctx.cell_viscosity()
    assert np.array_equal(nu > 0, g["nu"] > 0)
    assert relerr(nu, g["nu"]) < TOL
    assert sp.GetTimeStep(1.0, 3) == pytest.approx(float(g["dt"]), rel=1e-13)
    sp.SetTime(cases.GOLDEN_T)
    assert relerr(sp.GetResidualColumn(), g["residual"]) < TOL
    RungeKutta(3).Update(sp, 0.0, cases.GOLDEN_DT)
    assert relerr(sp.GetSolutionColumn(), g["u_rk3"]) < 1e-12
    assert relerr(sp.ctx.cell_viscosity(), g["nu_rk3"]) < 1e-9


def test_energy_based_viscosity_properties_at_size():
    """Smooth data gets no viscosity at all; a jump switches it on only next to the jump; it never exceeds
    lambda_max h / P; the damping matrices are negative semi-definite in the quadrature-weighted sense
    (u^T D u <= 0, the assert of viscosity.hpp:520)."""
    m = meshmod.box(12, 10, 8, hi=(1.2, 1.0, 0.8))
    sp = Spatial(Part(m), degree=2, riemann=capi.HLLC, device=0, diffusion=capi.CELL_VISCOSITY_DDG)
    for tag in range(6):
        sp.SetInviscidWall(tag)
    sp.SetDiffusion(*cases.ENERGY_BETAS, 0.0, 0.708)
    sp.EnableEnergyBasedViscosity()
    pts = cell_points(m.nodes, m.cells, 2)
    smooth = fields.smooth(0.5)(pts, 0.0)
    sp.SetSolutionColumn(smooth.reshape(-1))
    assert sp.ctx.cell_viscosity().max() == 0.0
    xc = pts[..., 0].mean(1)
    jumped = smooth.copy()
    jumped[xc > 0.6] *= 1.5                                     # whole cells: x = 0.6 is a plane of faces
    sp.SetSolutionColumn(jumped.reshape(-1))
    nu = sp.ctx.cell_viscosity()
    near = np.abs(xc - 0.6) < 0.11
    assert nu[near].max() > 0 and nu[~near].max() == 0.0
    D = sp.ctx.damping_matrices()
    rng = np.random.default_rng(0)
    v = rng.normal(size=(len(D), 27))
    assert (np.einsum("ci,cij,cj->c", v, D, v) <= 1e-12).all()


@pytest.mark.parametrize("degree,c_next,riemann", [(2, 0.3, capi.HLLC), (3, 0.8, capi.EXACT)])
def test_fr_general_with_navier_stokes(degree, c_next, riemann):
    """fr::General (Vincent correction over the whole node line) on the Navier-Stokes Riemann type; no-slip walls
    contribute nothing there (fr::General does not override AddFluxOnNoSlipWalls, fem.hpp:382-383)."""
    m = cases.small_mesh(seed=6)
    u = cases.noisy_state(cell_points(m.nodes, m.cells, degree), seed=8, amp=1e-4)
    o = oc.Oracle(m, degree, 8, riemann, c_next=c_next)
    cases.apply_viscous_oracle(o)
    sp = Spatial(Part(m), scheme=capi.FR_GENERAL, degree=degree, riemann=riemann, device=0, fr_c_next=c_next,
                 diffusion=capi.NAVIER_STOKES_DDG)
    cases.apply_viscous_spatial(sp)
    sp.SetSolutionColumn(u)
    sp.SetTime(0.2)
    assert relerr(sp.GetResidualColumn(), o.residual(u, 0.2)) < TOL
    assert sp.GetTimeStep(1.0, 3) == pytest.approx(o.time_step(u, 1.0, 3), rel=1e-13)
    RungeKutta(3).Update(sp, 0.0, 1e-3)
    assert relerr(sp.GetSolutionColumn(), o.rk_update(u, 3, 0.0, 1e-3)) < TOL


def test_energy_based_viscosity_two_parts_with_halo_buffers():
    """Two parts on one GPU: damping matrices with ghost cells present, UpdateProperties over ghost faces, and the ghost
    cells' viscosities through mcfd_viscosity_halo_*.  With the new state exchanged BEFORE every UpdateProperties the
    result does not depend on the partition and equals the single-part oracle."""
    import torch
    from minicfd_b200.halo import _DeviceArray
    m = meshmod.scramble(meshmod.box(6, 4, 2, warp=0.1, seed=2), seed=3)
    parts = (np.arange(m.n_cells) * 2 // m.n_cells).astype(np.int32)
    u = cases.noisy_state(cell_points(m.nodes, m.cells, 2), seed=9, amp=cases.ENERGY_AMP)
    o = oc.Oracle(m, 2, 6, 1)
    for tag in range(6):
        o.set_bc(tag, 0)
    o.set_viscous(*cases.ENERGY_BETAS, 0.0, 0.708)
    damping = o.damping_matrices()
    dt = 1e-3
    def set_solution(uu):
        nu, _ = o.update_viscosity(uu, damping)
        o.set_cell_viscosity(nu)
        return uu, nu
    u0, nu0 = set_solution(u)
    assert 0 < (nu0 > 0).sum() < nu0.size
    u1, _ = set_solution(u0 + dt * o.residual(u0, 0.0))
    u2, _ = set_solution((u0 * 3 + u1 + dt * o.residual(u1, dt)) / 4)
    u3, nu3 = set_solution((u0 + 2 * (u2 + dt * o.residual(u2, dt / 2))) / 3)

    P = [Part(m, parts, r, 2) for r in range(2)]
    S = [Spatial(p, degree=2, riemann=capi.HLLC, device=0, diffusion=capi.CELL_VISCOSITY_DDG) for p in P]
    for sp in S:
        for tag in range(6):
            sp.SetInviscidWall(tag)
        sp.SetDiffusion(*cases.ENERGY_BETAS, 0.0, 0.708)
        sp.ctx.build_damping_matrices()
    for sp, p in zip(S, P):
        assert relerr(sp.ctx.damping_matrices(), damping[p.local_global]) < TOL
    ubuf = [sum(sp.ctx.halo_buffers(), ()) for sp in S]
    nbuf = [sum(sp.ctx.viscosity_halo_buffers(), ()) for sp in S]

    def copy(bufs):
        torch.cuda.synchronize()
        for a, b in ((0, 1), (1, 0)):
            src = torch.as_tensor(_DeviceArray(bufs[a][0], bufs[a][1]), device="cuda")
            dst = torch.as_tensor(_DeviceArray(bufs[b][2], bufs[b][3]), device="cuda")
            dst.copy_(src)
        torch.cuda.synchronize()

    def refresh():                      # new state to the ghosts, UpdateProperties, viscosities to the ghosts
        for sp in S:
            sp.ctx.halo_pack()
        copy(ubuf)
        for sp in S:
            sp.ctx.halo_unpack()
            sp.ctx.update_cell_viscosity(1.0)
            sp.ctx.viscosity_halo_pack()
        copy(nbuf)
        for sp in S:
            sp.ctx.viscosity_halo_unpack()

    for sp, p in zip(S, P):
        sp.ctx.set_solution(p.gather_from_global(u, 135))
    refresh()
    for sp, p in zip(S, P):
        nu = sp.ctx.cell_viscosity(p.n_local + p.n_ghost)
        assert relerr(nu, nu0[p.global_ids]) < TOL          # local AND ghost cells
    for sp in S:
        sp.ctx.rk_begin(3, 0.0, dt)
    for stage in range(3):
        for sp in S:
            sp.ctx.rk_stage_inner(stage)
            sp.ctx.rk_stage_inter(stage)
        refresh()
    for sp, p in zip(S, P):
        assert relerr(sp.GetSolutionColumn().reshape(-1, 135), u3.reshape(-1, 135)[p.local_global]) < 1e-12
        assert relerr(sp.ctx.cell_viscosity(), nu3[p.local_global]) < 1e-9


@pytest.mark.parametrize("name,degree,riemann", cases.ENERGY_NS_GOLDEN_CASES)
def test_energy_based_viscosity_over_navier_stokes_golden(name, degree, riemann):
    """WithViscosity<fr::Lobatto<Part, EnergyBasedViscosity<Part, ConvectionDiffusion<Euler, Direct<NavierStokes>>>>>
    (demo/laminar/shocked.hpp:100-111): MCFD_NAVIER_STOKES_CELL_VISCOSITY_DDG, against the reference's vectors —
    damping matrices, viscosities, time step, residual, one RK3 step with the update after every stage."""
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    m = meshmod.HexMesh(nodes=g["nodes"], cells=g["cells"], bfaces=g["bfaces"], btags=g["btags"])
    sp = Spatial(Part(m), degree=degree, riemann=riemann, device=0, diffusion=capi.NAVIER_STOKES_CELL_VISCOSITY_DDG)
    cases.apply_viscous_spatial(sp)                            # SetDiffusion(beta_0, beta_1, nu, Pr) + viscous boundaries
    sp.EnableEnergyBasedViscosity()
    assert relerr(sp.ctx.damping_matrices()[:4], g["damping"]) < TOL
    sp.SetSolutionColumn(g["u"])                               # -> UpdateProperties
    nu = sp.ctx.cell_viscosity()
    assert np.array_equal(nu > 0, g["nu"] > 0)
    assert relerr(nu, g["nu"]) < TOL
    assert sp.GetTimeStep(1.0, 3) == pytest.approx(float(g["dt"]), rel=1e-13)
    sp.SetTime(cases.GOLDEN_T)
    assert relerr(sp.GetResidualColumn(), g["residual"]) < TOL
    RungeKutta(3).Update(sp, 0.0, cases.GOLDEN_DT)
    assert relerr(sp.GetSolutionColumn(), g["u_rk3"]) < 1e-12
    assert relerr(sp.ctx.cell_viscosity(), g["nu_rk3"]) < 1e-9
