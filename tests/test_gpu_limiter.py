"""GPU parity tests of the WENO limiter: mini::limiter::Reconstruct(part, weno::Lazy) on
Extrapolation<Hexahedron> cells, alone and inside the Runge-Kutta stages (WithLimiter), through the C ABI."""
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


@pytest.mark.parametrize("name,degree,riemann", cases.WENO_GOLDEN_CASES)
def test_weno_lazy_golden(name, degree, riemann):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    m = meshmod.HexMesh(nodes=g["nodes"], cells=g["cells"], bfaces=g["bfaces"], btags=g["btags"])
    sp = Spatial(Part(m), degree=degree, riemann=riemann, device=0)
    cases.apply_bcs_spatial(sp, cases.MIXED_BCS)
    sp.SetLimiter(*cases.WENO)
    sp.SetSolutionColumn(g["u"])                 # WithLimiter::SetSolutionColumn: set, then limit
    assert relerr(sp.GetSolutionColumn(), g["limited"]) < TOL
    RungeKutta(3).Update(sp, 0.0, cases.GOLDEN_DT)
    assert relerr(sp.GetSolutionColumn(), g["u_rk3"]) < TOL


@pytest.mark.parametrize("degree", [2, 3])
def test_weno_lazy_vs_oracle_other_parameters(degree):
    m = meshmod.scramble(meshmod.box(4, 3, 3, warp=0.2, seed=4), seed=9)
    u = cases.noisy_state(cell_points(m.nodes, m.cells, degree), seed=3, amp=1e-2)
    o = oc.Oracle(m, degree, 0, 1)
    for tag in range(6):
        o.set_bc(tag, 0)
    sp = Spatial(Part(m), degree=degree, riemann=capi.HLLC, device=0)
    for tag in range(6):
        sp.SetInviscidWall(tag)
    sp.SetSolutionColumn(u)
    sp.ctx.limiter_weno_lazy(0.01, 1e-5)
    limited = o.weno_lazy(u, 0.01, 1e-5)
    assert relerr(sp.GetSolutionColumn(), limited) < TOL
    # generic column algebra (rk.hpp) on the limited system == fused stages + limiter
    sp.SetLimiter(0.01, 1e-5)
    o.set_limiter(0.01, 1e-5)
    for order in (1, 2, 3):
        sp.SetSolutionColumn(u)
        RungeKutta(order).Update(sp, 0.0, 1e-3)
        assert relerr(sp.GetSolutionColumn(), o.rk_update(limited, order, 0.0, 1e-3)) < TOL
    sp.SetSolutionColumn(u)
    RungeKutta(3).UpdateGeneric(sp, 0.0, 1e-3)
    assert relerr(sp.GetSolutionColumn(), o.rk_update(limited, 3, 0.0, 1e-3)) < TOL


def test_weno_preserves_what_lies_in_the_modal_space():
    """A quadratic field on a parallelepiped mesh is its own projection, every borrowed candidate is the same
    polynomial, so the blend returns it; a periodic box exercises the centre offsets through the faces."""
    m = meshmod.box(5, 4, 3, hi=(2.5, 2.0, 1.5), periodic=False)
    def quad(x, t=0.0):
        X, Y, Z = x[..., 0], x[..., 1], x[..., 2]
        rho = 1.0 + 0.1 * X + 0.05 * Y * Z - 0.02 * X * X
        out = np.stack([rho, 0.3 * rho + 0.01 * Y, -0.1 + 0.02 * Z * Z, 0.05 * X * Y, 2.5 + 0.1 * Z], axis=-1)
        return out
    sp = Spatial(Part(m), degree=2, riemann=capi.HLLC, device=0)
    for tag in range(6):
        sp.SetInviscidWall(tag)
    sp.Approximate(quad)
    u = sp.GetSolutionColumn()
    sp.ctx.limiter_weno_lazy(*cases.WENO)
    assert np.abs(sp.GetSolutionColumn() - u).max() < 1e-12
    # constants on a periodic, warped mesh
    mp = meshmod.box(6, 6, 6, hi=(1.0, 1.0, 1.0), periodic=True)
    spp = Spatial(Part(mp), degree=2, riemann=capi.HLLC, device=0)
    state = np.array([1.2, 0.3, -0.2, 0.1, 2.5])
    spp.Approximate(lambda x: fields.constant(state)(x, 0.0))
    spp.ctx.limiter_weno_lazy(*cases.WENO)
    assert np.abs(spp.GetSolutionColumn().reshape(-1, 5) - state).max() < 1e-13


def test_weno_damps_an_oscillating_cell_and_keeps_the_mean():
    m = meshmod.box(6, 5, 4, hi=(6.0, 5.0, 4.0))
    sp = Spatial(Part(m), degree=2, riemann=capi.HLLC, device=0)
    for tag in range(6):
        sp.SetInviscidWall(tag)
    pts = cell_points(m.nodes, m.cells, 2)
    u = fields.smooth(0.3)(pts, 0.0)
    rng = np.random.default_rng(0)
    bad = 37
    u[bad] *= 1 + 0.3 * rng.uniform(-1, 1, size=u[bad].shape)
    sp.SetSolutionColumn(u.reshape(-1))
    length, volume = sp.ctx.cell_geometry()
    w = np.array([1 / 3, 4 / 3, 1 / 3])
    W = (w[:, None, None] * w[None, :, None] * w[None, None, :]).reshape(-1) / 8
    mean_before = (u[bad] * W[:, None]).sum(0)
    sp.ctx.limiter_weno_lazy(*cases.WENO)
    v = sp.GetSolutionColumn().reshape(-1, 27, 5)
    mean_after = (v[bad] * W[:, None]).sum(0)
    assert np.abs(mean_after - mean_before).max() < 1e-12 * np.abs(mean_before).max() + 1e-13
    osc = lambda a: np.abs(a - (a * W[:, None]).sum(0)).max()
    assert osc(v[bad]) < 0.5 * osc(u[bad])


@pytest.mark.skipif(not refapi.available(2), reason="oracle/_ref not shipped")
def test_weno_two_parts_with_halo_buffers_match_the_reference_parts():
    """Two parts on one GPU: phase 0 (inner cells), halo of the unlimited state, phase 1 (inter cells), and RK3
    steps with the limiter after every stage, against the reference run on the same two parts."""
    import torch
    from minicfd_b200.halo import _DeviceArray
    m = meshmod.scramble(meshmod.box(6, 4, 2, warp=0.1, seed=2), seed=3)
    parts = (np.arange(m.n_cells) * 2 // m.n_cells).astype(np.int32)
    u = cases.noisy_state(cell_points(m.nodes, m.cells, 2), seed=9, amp=1e-3)
    r = refapi.Ref(m, degree=2, scheme=refapi.FR_WENO_LAZY, riemann=1, cell_parts=parts, n_parts=2, weno=cases.WENO)
    for tag in range(6):
        r.set_bc(tag, refapi.WALL)
    P = [Part(m, parts, k, 2) for k in range(2)]
    r.set_solution_all([p.gather_from_global(u, 135) for p in P])
    S = [Spatial(p, degree=2, riemann=capi.HLLC, device=0) for p in P]
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
        for sp in S:
            sp.ctx.halo_unpack()

    def limit():
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

    for sp, p in zip(S, P):
        sp.ctx.set_solution(p.gather_from_global(u, 135))
    limit()
    for k, sp in enumerate(S):
        assert relerr(sp.GetSolutionColumn(), r.get_solution(k)) < TOL
    r.rk_update(3, 0.0, 1e-3, 2)
    t = 0.0
    for _ in range(2):
        for sp in S:
            sp.ctx.rk_begin(3, t, 1e-3)
        for stage in range(3):
            exchange()
            for sp in S:
                sp.ctx.rk_stage_inner(stage)
                sp.ctx.rk_stage_inter(stage)
            limit()
        t += 1e-3
    for k, sp in enumerate(S):
        assert relerr(sp.GetSolutionColumn(), r.get_solution(k)) < 1e-12


# ---- weno::Eigen and EigenMatrices ------------------------------------------------------
@pytest.mark.parametrize("name,degree,riemann", cases.WENO_EIGEN_GOLDEN_CASES)
def test_weno_eigen_golden(name, degree, riemann):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    m = meshmod.HexMesh(nodes=g["nodes"], cells=g["cells"], bfaces=g["bfaces"], btags=g["btags"])
    sp = Spatial(Part(m), degree=degree, riemann=riemann, device=0)
    cases.apply_bcs_spatial(sp, cases.MIXED_BCS)
    sp.SetLimiter(*cases.WENO, eigen=True)
    sp.SetSolutionColumn(g["u"])
    got = sp.GetSolutionColumn()
    kf = sp.kfields
    touched = np.abs((g["limited"] - g["u"]).reshape(-1, kf)).max(1) > 0
    assert np.array_equal(np.abs((got - g["u"]).reshape(-1, kf)).max(1) > 0, touched)   # same troubled cells
    assert relerr(got, g["limited"]) < TOL
    RungeKutta(3).Update(sp, 0.0, cases.GOLDEN_DT)
    assert relerr(sp.GetSolutionColumn(), g["u_rk3"]) < 1e-12


@pytest.mark.parametrize("degree", [2, 3])
def test_weno_eigen_vs_oracle(degree):
    m = meshmod.scramble(meshmod.box(4, 3, 3, warp=0.2, seed=4), seed=9)
    u = cases.shocked_state(cell_points(m.nodes, m.cells, degree), seed=3)
    o = oc.Oracle(m, degree, 0, 1)
    sp = Spatial(Part(m), degree=degree, riemann=capi.HLLC, device=0)
    sp.ctx.set_solution(u)
    sp.ctx.limiter_weno_eigen(0.01, 1e-5)
    assert relerr(sp.GetSolutionColumn(), o.weno_eigen(u, 0.01, 1e-5)) < TOL


def test_eigen_matrices_on_device():
    """mini::riemann::euler::EigenMatrices (eigen.hpp:27-56): L R = I, and both against the oracle."""
    import ctypes as C
    rng = np.random.default_rng(5)
    n = 64
    rho = rng.uniform(0.5, 2.0, n); vel = rng.uniform(-1, 1, (n, 3)); p = rng.uniform(0.5, 3.0, n)
    cons = np.column_stack([rho, rho[:, None] * vel, p / 0.4 + 0.5 * rho * (vel ** 2).sum(1)])
    q, _ = np.linalg.qr(rng.normal(size=(n, 3, 3)))
    frames = np.ascontiguousarray(np.transpose(q, (0, 2, 1))).reshape(n, 9)    # rows = nu, mu, pi
    sp = Spatial(Part(cases.small_mesh()), degree=2, riemann=capi.HLLC, device=0)
    L, R = sp.ctx.eigen_matrices(cons, frames)
    assert np.abs(L @ R - np.eye(5)).max() < 1e-12
    lib = oc.lib()
    for i in range(n):
        Lo = np.empty(25); Ro = np.empty(25)
        lib.oracle_eigen_matrices(oc._dp(np.ascontiguousarray(cons[i])), oc._dp(np.ascontiguousarray(frames[i])),
                                  oc._dp(Lo), oc._dp(Ro))
        assert np.abs(L[i].reshape(-1) - Lo).max() <= 1e-13 * np.abs(Lo).max()
        assert np.abs(R[i].reshape(-1) - Ro).max() <= 1e-13 * np.abs(Ro).max()


@pytest.mark.skipif(not refapi.available(2), reason="oracle/_ref not shipped")
def test_weno_eigen_two_parts_match_the_reference_parts():
    import torch
    from minicfd_b200.halo import _DeviceArray
    m = meshmod.scramble(meshmod.box(6, 4, 2, warp=0.1, seed=2), seed=3)
    parts = (np.arange(m.n_cells) * 2 // m.n_cells).astype(np.int32)
    u = cases.shocked_state(cell_points(m.nodes, m.cells, 2))
    r = refapi.Ref(m, degree=2, scheme=refapi.FR_WENO_EIGEN, riemann=1, cell_parts=parts, n_parts=2, weno=cases.WENO)
    P = [Part(m, parts, k, 2) for k in range(2)]
    r.set_solution_all([p.gather_from_global(u, 135) for p in P])
    S = [Spatial(p, degree=2, riemann=capi.HLLC, device=0) for p in P]
    bufs = [sum(sp.ctx.halo_buffers(), ()) for sp in S]
    for sp, p in zip(S, P):
        sp.ctx.set_solution(p.gather_from_global(u, 135))
    for sp in S:
        sp.ctx.halo_pack()
    torch.cuda.synchronize()
    for a, b in ((0, 1), (1, 0)):
        src = torch.as_tensor(_DeviceArray(bufs[a][0], bufs[a][1]), device="cuda")
        dst = torch.as_tensor(_DeviceArray(bufs[b][2], bufs[b][3]), device="cuda")
        dst.copy_(src)
    torch.cuda.synchronize()
    for sp in S:
        sp.ctx.limiter_weno_eigen(*cases.WENO, 0)
        sp.ctx.halo_unpack()
        sp.ctx.limiter_weno_eigen(*cases.WENO, 1)
    for k, sp in enumerate(S):
        assert relerr(sp.GetSolutionColumn(), r.get_solution(k)) < TOL


def test_dg_lobatto_with_weno_lazy_vs_oracle():
    """BASELINE config 3's combination: dg::Lobatto + Extrapolation + weno::Lazy(1e-3, 1e-6)."""
    m = cases.small_mesh(seed=31)
    u = cases.noisy_state(cell_points(m.nodes, m.cells, 2), seed=5, amp=1e-2)
    o = oc.Oracle(m, 2, 1, 1)
    cases.apply_bcs_oracle(o)
    o.set_limiter(*cases.WENO)
    limited = o.weno_lazy(u, *cases.WENO)
    sp = Spatial(Part(m), scheme=capi.DG_LOBATTO, degree=2, riemann=capi.HLLC, device=0)
    cases.apply_bcs_spatial(sp, cases.MIXED_BCS)
    sp.SetLimiter(*cases.WENO)
    sp.SetSolutionColumn(u)
    assert relerr(sp.GetSolutionColumn(), limited) < TOL
    RungeKutta(3).Update(sp, 0.0, 1e-3)
    assert relerr(sp.GetSolutionColumn(), o.rk_update(limited, 3, 0.0, 1e-3)) < TOL
