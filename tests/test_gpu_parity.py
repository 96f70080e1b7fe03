"""GPU parity tests: the CUDA path, called through the C ABI (via minicfd_b200.capi / Spatial),
against the oracle on identical inputs.  Tolerance: 1e-12 relative (max norm), the bound stated in
BASELINE.json's north_star; index tables bit-exact."""
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


def make_spatial(m, degree, scheme, riemann, kinds=None, **kw):
    sp = Spatial(Part(m), scheme=scheme, degree=degree, riemann=riemann, device=0, **kw)
    if kinds is None:
        for tag in range(6):
            sp.SetInviscidWall(tag)
    else:
        cases.apply_bcs_spatial(sp, kinds)
    return sp


@pytest.mark.parametrize("name,degree,scheme,riemann", cases.GOLDEN_CASES)
def test_golden_residual_timestep_rk3(name, degree, scheme, riemann):
    """Committed golden vectors of the reference: residual at t, GetTimeStep, one RK3 step."""
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    m = meshmod.HexMesh(nodes=g["nodes"], cells=g["cells"], bfaces=g["bfaces"], btags=g["btags"])
    sp = make_spatial(m, degree, scheme, riemann, cases.MIXED_BCS)
    sp.SetSolutionColumn(g["u"])
    assert np.array_equal(sp.GetSolutionColumn(), g["u"])           # layout round trip is bit-exact
    sp.SetTime(cases.GOLDEN_T)
    assert relerr(sp.GetResidualColumn(), g["residual"]) < TOL
    assert sp.GetTimeStep(1.0, 3) == pytest.approx(float(g["dt"]), rel=1e-13)
    RungeKutta(3).Update(sp, 0.0, cases.GOLDEN_DT)
    assert relerr(sp.GetSolutionColumn(), g["u_rk3"]) < TOL


@pytest.mark.parametrize("degree", [2, 3])
@pytest.mark.parametrize("scheme", [capi.FR_LOBATTO, capi.DG_LOBATTO])
@pytest.mark.parametrize("riemann", [capi.EXACT, capi.HLLC, capi.AUSM])
def test_residual_and_rk_orders_vs_oracle(degree, scheme, riemann):
    """Every (degree, scheme, Riemann) instantiation on a scrambled warped mesh, RK1/2/3, fused stage
    kernels vs the reference's column algebra (UpdateGeneric) vs the oracle."""
    m = meshmod.scramble(meshmod.box(4, 3, 3, warp=0.2, seed=21), seed=8)
    u = cases.noisy_state(cell_points(m.nodes, m.cells, degree), seed=3)
    o = oc.Oracle(m, degree, scheme, riemann)
    cases.apply_bcs_oracle(o)
    sp = make_spatial(m, degree, scheme, riemann, cases.MIXED_BCS)
    sp.SetSolutionColumn(u)
    sp.SetTime(0.3)
    assert relerr(sp.GetResidualColumn(), o.residual(u, 0.3)) < TOL
    assert sp.GetTimeStep(1.0, 3) == pytest.approx(o.time_step(u, 1.0, 3), rel=1e-13)
    for order in (1, 2, 3):
        expect = o.rk_update(u, order, 0.1, 1.5e-3)
        sp.SetSolutionColumn(u)
        RungeKutta(order).Update(sp, 0.1, 1.5e-3)
        fused = sp.GetSolutionColumn()
        sp.SetSolutionColumn(u)
        RungeKutta(order).UpdateGeneric(sp, 0.1, 1.5e-3)
        generic = sp.GetSolutionColumn()
        assert relerr(fused, expect) < TOL and relerr(generic, expect) < TOL


@pytest.mark.skipif(not refapi.available(2), reason="oracle/_ref not shipped")
@pytest.mark.parametrize("degree,riemann", [(2, capi.EXACT), (2, capi.HLLC), (3, capi.AUSM)])
def test_against_reference_binary_many_steps(degree, riemann):
    """Ten RK3 steps with GetTimeStep each step, against the reference's own code (oracle/_ref)."""
    m = meshmod.box(6, 4, 3, hi=(3.0, 2.0, 1.5), warp=0.1, seed=4)
    r = refapi.Ref(m, degree=degree, scheme=refapi.FR, riemann=riemann)
    cases.apply_bcs_ref(r, {t: cases.WALL for t in range(6)})
    sp = make_spatial(m, degree, capi.FR_LOBATTO, riemann)
    # a gentle state and a small dt_guess: the run stays smooth (no clipped negative-pressure
    # states, whose handling in the reference returns uninitialised fluxes, hllc.hpp:34-43)
    u = cases.noisy_state(cell_points(m.nodes, m.cells, degree), seed=5, amp=1e-4)
    r.set_solution(u)
    sp.SetSolutionColumn(u)
    t = 0.0
    for _ in range(10):
        dt_ref, dt = r.time_step(4e-3, 3), sp.GetTimeStep(4e-3, 3)
        assert dt == pytest.approx(dt_ref, rel=1e-12)
        r.rk_update(3, t, dt_ref, 1)
        RungeKutta(3).Update(sp, t, dt_ref)
        t += dt_ref
    # 30 stages; the reference itself moves by 6e-16 under a one-ulp change of the input over these steps
    # (tests/test_oracle_pinned.py::test_reference_drift_under_one_ulp_perturbation)
    assert relerr(sp.GetSolutionColumn(), r.get_solution()) < 1e-12


def test_connectivity_tables_bit_exact():
    """Face tables (holder, sharer, local face ids) against the reference's numbering rules."""
    m = cases.small_mesh(seed=12)
    sp = make_spatial(m, 2, 0, 1)
    holder, sharer, hf, sf, orient = sp.ctx.faces(0)
    if refapi.available(2):
        r = refapi.Ref(m, degree=2)
        rh, rs, rf = r.faces(0, 0)
        assert np.array_equal(holder, rh) and np.array_equal(sharer, rs) and np.array_equal(hf, rf)
        bh, _, bf = r.faces(0, 2)
        mh, _, mf, _, _ = sp.ctx.faces(2)
        assert np.array_equal(mh, bh) and np.array_equal(mf, bf)
        length, volume = sp.ctx.cell_geometry()
        rl, rv = r.cell_geometry()
        assert np.allclose(length, rl, rtol=1e-14, atol=0) and np.allclose(volume, rv, rtol=1e-14, atol=0)
    assert (holder < sharer).all() and len(holder) == 33      # 3x3x2 box: 12+12+9 interior faces
    assert sorted(set(orient)) != []


def test_periodic_equals_embedded_box_and_conserves():
    """Periodic variant (GPU only): equals the interior of an embedded larger box; and the discrete
    conservation property sum_cells sum_nodes w detJ R = 0 holds to rounding."""
    L = 2 * np.pi
    m = meshmod.box(6, 5, 4, hi=(L, L, L), periodic=True)
    sp = Spatial(Part(m), degree=2, riemann=capi.HLLC, device=0)
    sp.Approximate(lambda x: fields.smooth(1.0)(x, 0.0))
    res = sp.GetResidualColumn().reshape(4, 5, 6, 27, 5)
    hx, hy, hz = L / 6, L / 5, L / 4
    mb = meshmod.box(8, 7, 6, lo=(-hx, -hy, -hz), hi=(L + hx, L + hy, L + hz))
    spb = make_spatial(mb, 2, 0, capi.HLLC)
    spb.Approximate(lambda x: fields.smooth(1.0)(x, 0.0))
    resb = spb.GetResidualColumn().reshape(6, 7, 8, 27, 5)
    assert relerr(res, resb[1:5, 1:6, 1:7]) < TOL
    w1 = np.array([1 / 3, 4 / 3, 1 / 3])
    w = np.einsum("i,j,k->ijk", w1, w1, w1).reshape(27)
    total = np.einsum("zyxnc,n->c", res, w)   # detJ is constant on this mesh
    assert np.abs(total).max() < 1e-10 * np.abs(res).max() * m.n_cells


def test_free_stream_preservation_on_warped_mesh():
    """A uniform state has zero residual on a curved (warped trilinear) mesh only up to the metric
    identity error of the scheme; the reference has the same property, so compare with the oracle."""
    m = meshmod.box(4, 4, 4, warp=0.2, seed=6)
    state = fields.primitive_to_conservative(*(np.array(v) for v in (1.1, 0.4, -0.3, 0.2, 0.9)))
    sp = make_spatial(m, 2, 0, capi.AUSM, {t: cases.SMART for t in range(6)})
    for tag in range(6):
        sp.SetSmartBoundary(tag, fields.constant(state), time_dependent=False)
    sp.Approximate(lambda x: fields.constant(state)(x, 0.0))
    o = oc.Oracle(m, 2, 0, 2)
    for tag in range(6):
        o.set_bc(tag, cases.SMART, oc.F_CONST, state)
    ref = o.residual(sp.GetSolutionColumn(), 0.0)
    got = sp.GetResidualColumn()
    assert np.abs(got - ref).max() < 1e-12 * max(1.0, np.abs(ref).max())


def test_two_parts_on_one_gpu_with_halo_buffers():
    """Two parts of a partitioned mesh held by two contexts on one GPU; the halo goes through the
    library's pack -> (device copy standing in for the NVLink transfer) -> unpack path and the staged
    RK interface.  Must equal the single-part result (and the oracle)."""
    import torch
    m = meshmod.scramble(meshmod.box(6, 4, 2, warp=0.1, seed=2), seed=3)
    parts = (np.arange(m.n_cells) * 2 // m.n_cells).astype(np.int32)
    u = cases.noisy_state(cell_points(m.nodes, m.cells, 2), seed=9)
    o = oc.Oracle(m, 2, 0, 1)
    for tag in range(6):
        o.set_bc(tag, 0)
    expect_res = o.residual(u, 0.0).reshape(-1, 135)
    expect_u = o.rk_update(u, 3, 0.0, 1e-3).reshape(-1, 135)
    P = [Part(m, parts, r, 2) for r in range(2)]
    S = [Spatial(p, degree=2, riemann=capi.HLLC, device=0) for p in P]
    for sp in S:
        for tag in range(6):
            sp.SetInviscidWall(tag)
    bufs = []
    for sp in S:
        (s_ptr, s_n), (r_ptr, r_n) = sp.ctx.halo_buffers()
        assert s_n > 0 and r_n > 0
        bufs.append((s_ptr, s_n, r_ptr, r_n))
    assert bufs[0][1] == bufs[1][3] and bufs[1][1] == bufs[0][3]

    def exchange():
        from minicfd_b200.halo import _DeviceArray
        for a, b in ((0, 1), (1, 0)):
            S[a].ctx.halo_pack()
        torch.cuda.synchronize()
        for a, b in ((0, 1), (1, 0)):
            src = torch.as_tensor(_DeviceArray(bufs[a][0], bufs[a][1]), device="cuda")
            dst = torch.as_tensor(_DeviceArray(bufs[b][2], bufs[b][3]), device="cuda")
            dst.copy_(src)
        torch.cuda.synchronize()
        for sp in S:
            sp.ctx.halo_unpack()

    for sp, p in zip(S, P):
        sp.SetSolutionColumn(p.gather_from_global(u, 135))
    exchange()
    for sp, p in zip(S, P):
        assert relerr(sp.ctx.residual(0.0).reshape(-1, 135), expect_res[p.local_global]) < TOL
    for sp in S:
        sp.ctx.rk_begin(3, 0.0, 1e-3)
    for stage in range(3):
        for sp in S:
            sp.ctx.halo_pack()
        torch.cuda.synchronize()
        from minicfd_b200.halo import _DeviceArray
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
        assert relerr(sp.GetSolutionColumn().reshape(-1, 135), expect_u[p.local_global]) < TOL


@pytest.mark.parametrize("riemann", [capi.EXACT, capi.HLLC, capi.AUSM])
def test_riemann_solvers_on_device(riemann):
    """Device Riemann solvers vs the oracle on random rotated problems, incl. the wall flux."""
    rng = np.random.default_rng(31415926)
    n = 2000
    nrm = rng.normal(size=(n, 3)); nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    rho = rng.uniform(0.1, 2.0, size=(2, n)); p = rng.uniform(0.1, 2.0, size=(2, n))
    v = rng.uniform(-1.5, 1.5, size=(2, n, 3))
    U = np.concatenate([rho[..., None], rho[..., None] * v,
                        (p / 0.4 + 0.5 * rho * (v ** 2).sum(-1))[..., None]], axis=-1)
    ctx = capi.Context(riemann=riemann, device=0)
    got = ctx.riemann_flux(nrm, U[0], U[1])
    wall = ctx.riemann_flux(nrm, U[0], U[1], bc_kind=capi.BC_INVISCID_WALL)
    for i in range(0, n, 7):
        t1 = np.cross(nrm[i], [0.3, -0.5, 0.8]); t1 /= np.linalg.norm(t1)
        frame = np.array([nrm[i], t1, np.cross(nrm[i], t1)])
        ref = oc.riemann_flux(riemann, frame, U[0, i], U[1, i])
        assert np.abs(got[i] - ref).max() <= 1e-12 * max(1.0, np.abs(ref).max())
        refw = oc.riemann_flux(riemann, frame, U[0, i], U[1, i], bc=0)
        assert np.abs(wall[i] - refw).max() <= 1e-13 * max(1.0, np.abs(refw).max())


def test_full_size_properties_roofline_case():
    """BASELINE config 2 at full size (42^3 cells, 10 001 880 DOF, periodic): properties that need no
    oracle run — a uniform state is a fixed point of an RK3 step, mass/momentum/energy are conserved
    by a step of the vortex to rounding, and a copy of the mesh shifted by one period gives the
    same residual bit for bit (translation invariance of the index logic)."""
    m = meshmod.box(42, 42, 42, hi=(10.0, 10.0, 10.0), periodic=True)
    sp = Spatial(Part(m), degree=2, riemann=capi.HLLC, device=0)
    assert sp.part.n_local * 135 == 10_001_880
    state = fields.primitive_to_conservative(*(np.array(v) for v in (1.0, 1.0, 1.0, 0.0, 1.0)))
    sp.Approximate(lambda x: fields.constant(state)(x, 0.0))
    u0 = sp.GetSolutionColumn()
    RungeKutta(3).Update(sp, 0.0, 1e-3)
    assert np.abs(sp.GetSolutionColumn() - u0).max() < 1e-13
    sp.Approximate(lambda x: fields.vortex()(x, 0.0))
    u0 = sp.GetSolutionColumn().reshape(-1, 27, 5)
    dt = sp.GetTimeStep(1.0, 3)
    RungeKutta(3).Update(sp, 0.0, dt)
    u1 = sp.GetSolutionColumn().reshape(-1, 27, 5)
    w1 = np.array([1 / 3, 4 / 3, 1 / 3])
    w = np.einsum("i,j,k->ijk", w1, w1, w1).reshape(27)
    before, after = np.einsum("cnk,n->k", u0, w), np.einsum("cnk,n->k", u1, w)
    assert np.abs(after - before).max() < 1e-9 * np.abs(before).max()
    assert np.isfinite(u1).all() and np.abs(u1 - u0).max() > 1e-6


def test_multi_gpu_nccl_halo_matches_single_part():
    """N ranks, one per GPU, NCCL ghost exchange overlapped with the inner cells (needs >= 2 GPUs)."""
    import subprocess
    import sys
    import torch
    n = min(torch.cuda.device_count(), 4)
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for degree, mode in ((2, "euler"), (2, "peer"), (3, "peer"), (3, "euler"), (2, "viscous"), (2, "weno"), (2, "energy")):
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
               "--master-addr", "127.0.0.1", "--master-port", "29533",
               os.path.join(root, "tests", "mgpu_worker.py"), str(degree), mode]
        out = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
        assert out.returncode == 0 and "MGPU OK" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]


def test_device_resident_time_step_and_update():
    """mcfd_time_step_device + mcfd_rk_step_device (dt never leaves the GPU) equal the host-dt path bit for bit."""
    import torch
    from minicfd_b200.spatial import time_step_device
    m = cases.small_mesh(seed=4)
    u = cases.noisy_state(cell_points(m.nodes, m.cells, 2), seed=2)
    a = make_spatial(m, 2, 0, capi.HLLC)
    b = make_spatial(m, 2, 0, capi.HLLC)
    a.SetSolutionColumn(u)
    b.SetSolutionColumn(u)
    dt_dev = torch.zeros(1, dtype=torch.float64, device="cuda")
    for _ in range(3):
        dt = a.GetTimeStep(2e-3, 3)          # small steps: the noisy state stays admissible
        RungeKutta(3).Update(a, 0.0, dt)
        time_step_device(b, 2e-3, 3, dt_dev)
        RungeKutta(3).UpdateDevice(b, 0.0, dt_dev.data_ptr())
        assert float(dt_dev.item()) == dt
    got = b.GetSolutionColumn()
    assert np.isfinite(got).all() and np.array_equal(a.GetSolutionColumn(), got)


@pytest.mark.parametrize("degree", [2, 3])
@pytest.mark.parametrize("kind", [0, 1])
def test_initial_limiters(degree, kind):
    """average::Reconstruct / majority::Reconstruct on the device vs the oracle."""
    from minicfd_b200.spatial import average_reconstruct, majority_reconstruct
    m = cases.small_mesh(seed=5)
    n = (degree + 1) ** 3
    rng = np.random.default_rng(1)
    u = cases.noisy_state(cell_points(m.nodes, m.cells, degree), seed=3).reshape(m.n_cells, n, 5)
    if kind == 1:
        maj = u[:, :1, :].copy()
        mask = rng.random((m.n_cells, n, 1)) < 0.7
        mask[:, : n // 2 + 2] = True
        u = np.where(mask, maj, u)
    u = u.reshape(-1)
    sp = make_spatial(m, degree, 0, capi.HLLC)
    sp.SetSolutionColumn(u)
    (average_reconstruct if kind == 0 else majority_reconstruct)(sp)
    expect = oc.Oracle(m, degree).limiter(u, kind)
    got = sp.GetSolutionColumn()
    if kind == 1:
        assert np.array_equal(got, expect)          # pure selection: bit-exact
    else:
        assert relerr(got, expect) < TOL


@pytest.mark.parametrize("degree,c_next,riemann", [(2, 1.0, capi.HLLC), (2, 0.3, capi.EXACT), (3, 0.8, capi.AUSM)])
def test_fr_general_vincent(degree, c_next, riemann):
    """fr::General(part, c_next): residual, time step (no doubling) and RK3 vs the oracle."""
    m = cases.small_mesh(seed=8)
    u = cases.noisy_state(cell_points(m.nodes, m.cells, degree), seed=2)
    o = oc.Oracle(m, degree, 2, riemann, c_next=c_next)
    cases.apply_bcs_oracle(o)
    sp = Spatial(Part(m), scheme=capi.FR_GENERAL, degree=degree, riemann=riemann, device=0, fr_c_next=c_next)
    cases.apply_bcs_spatial(sp, cases.MIXED_BCS)
    sp.SetSolutionColumn(u)
    sp.SetTime(0.2)
    assert relerr(sp.GetResidualColumn(), o.residual(u, 0.2)) < TOL
    assert sp.GetTimeStep(1.0, 3) == pytest.approx(o.time_step(u, 1.0, 3), rel=1e-13)
    RungeKutta(3).Update(sp, 0.0, 1e-3)
    assert relerr(sp.GetSolutionColumn(), o.rk_update(u, 3, 0.0, 1e-3)) < TOL


def test_contexts_with_different_tables_coexist():
    """The per-degree tables sit in one __constant__ symbol; contexts of different degree / scheme / launch shape
    alive at the same time must each see their own (they are refreshed stream-ordered before a launch)."""
    m = cases.small_mesh(seed=4)
    S, U, O = [], [], []
    for degree, scheme in ((2, capi.FR_LOBATTO), (3, capi.DG_LOBATTO), (2, capi.FR_GENERAL)):
        kw = {"fr_c_next": 0.4} if scheme == capi.FR_GENERAL else {}
        S.append(make_spatial(m, degree, scheme, capi.HLLC, **kw))
        U.append(cases.noisy_state(cell_points(m.nodes, m.cells, degree), seed=degree))
        o = oc.Oracle(m, degree, scheme, 1, c_next=0.4 if scheme == capi.FR_GENERAL else 0.0)
        for tag in range(6):
            o.set_bc(tag, 0)
        O.append(o)
    for sp, u in zip(S, U):
        sp.SetSolutionColumn(u)
    for _ in range(2):
        for sp, u, o in zip(S, U, O):          # interleaved launches
            assert relerr(sp.GetResidualColumn(), o.residual(u, 0.0)) < TOL


def test_contexts_of_different_shape_on_two_streams_at_once():
    """Three contexts whose kernels read different tables (degree 3 slot kernel, fr::General, the node-line kernel with
    its own parameter copy) advance RK3 steps in turn on two streams, nothing synchronising in between: each must see
    its own tables although the launches of one are still in flight when the next one's are issued."""
    import torch
    m = meshmod.scramble(meshmod.box(6, 5, 4, warp=0.1, seed=7), seed=2)
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    S, U, O = [], [], []
    for k, (degree, scheme) in enumerate(((3, capi.DG_LOBATTO), (2, capi.FR_GENERAL), (2, capi.FR_LOBATTO))):
        kw = {"fr_c_next": 0.4} if scheme == capi.FR_GENERAL else {}
        sp = make_spatial(m, degree, scheme, capi.HLLC, **kw)
        sp.ctx.set_stream(streams[k % 2].cuda_stream)
        u = cases.noisy_state(cell_points(m.nodes, m.cells, degree), seed=degree + k)
        sp.SetSolutionColumn(u)
        o = oc.Oracle(m, degree, scheme, 1, c_next=0.4 if scheme == capi.FR_GENERAL else 0.0)
        for tag in range(6):
            o.set_bc(tag, 0)
        S.append(sp); U.append(u); O.append(o)
    for step in range(3):
        for sp in S:                                   # enqueue only: no getter, no synchronisation
            sp.ctx.rk_step(3, 1e-3 * step, 1e-3)
    torch.cuda.synchronize()
    for sp, u, o in zip(S, U, O):
        ref = u
        for step in range(3):
            ref = o.rk_update(ref, 3, 1e-3 * step, 1e-3)
        assert relerr(sp.GetSolutionColumn(), ref) < TOL


def test_full_size_parity_variant_against_reference():
    """BASELINE config 2 at full size in its PARITY variant: 42^3 cells, 10 001 880 DOF, FR-Lobatto p = 2,
    HLLC, the exact isentropic vortex as SetSmartBoundary on all six sides — residual, GetTimeStep and one RK3 step
    against the unmodified reference (oracle/_ref, the box's host threads)."""
    if not refapi.available(2):
        pytest.skip("oracle/_ref not shipped")
    n = 42
    m = meshmod.box(n, n, n, hi=(10.0, 10.0, 10.0))
    vortex = [5.0, 5.0, 1.0, 1.0, 0.0, 5.0]
    n_parts = max(1, min(8, os.cpu_count() or 1))
    px, py, pz = meshmod.factor3(n_parts)
    parts = meshmod.block_partition(m, px, py, pz)
    r = refapi.Ref(m, degree=2, scheme=refapi.FR, riemann=1, cell_parts=parts, n_parts=n_parts)
    for tag in range(6):
        r.set_bc(tag, refapi.SMART, refapi.F_VORTEX, vortex)
    r.approximate(refapi.F_VORTEX, vortex)
    P = [Part(m, parts, k, n_parts) for k in range(n_parts)]

    def whole(columns):
        out = np.zeros(m.n_cells * 135)
        for p, c in zip(P, columns if n_parts > 1 else [columns]):
            p.scatter_to_global(c, 135, out)
        return out
    sp = Spatial(Part(m), degree=2, riemann=capi.HLLC, device=0)
    assert sp.part.n_local * 135 == 10_001_880
    for tag in range(6):
        sp.SetSmartBoundary(tag, fields.vortex(), time_dependent=True)
    u0 = whole([r.get_solution(k) for k in range(n_parts)])
    sp.SetSolutionColumn(u0)
    sp.SetTime(0.0)
    assert relerr(sp.GetResidualColumn(), whole(r.residual(0.0))) < TOL
    dt_ref = r.time_step(1.0, 3)
    assert sp.GetTimeStep(1.0, 3) == pytest.approx(dt_ref, rel=1e-13)
    r.rk_update(3, 0.0, dt_ref, 1)
    RungeKutta(3).Update(sp, 0.0, dt_ref)
    assert relerr(sp.GetSolutionColumn(), whole([r.get_solution(k) for k in range(n_parts)])) < TOL


def test_device_resident_time_step_refuses_time_dependent_boundaries():
    """A device-resident dt leaves the host without the stage times t + dt, t + dt / 2: with a boundary function of
    time the device-dt entry points must fail loudly instead of sampling it at t three times."""
    import torch
    from minicfd_b200.spatial import time_step_device
    m = cases.small_mesh(seed=4)
    sp = make_spatial(m, 2, capi.FR_LOBATTO, capi.HLLC)
    sp.SetSupersonicInlet(meshmod.XMIN, lambda x, t: fields.smooth(1.0)(x, t), time_dependent=True)
    sp.SetSolutionColumn(cases.noisy_state(cell_points(m.nodes, m.cells, 2), seed=2))
    dt_dev = torch.zeros(1, dtype=torch.float64, device="cuda")
    time_step_device(sp, 2e-3, 3, dt_dev)
    with pytest.raises(capi.McfdError, match="time-dependent"):
        RungeKutta(3).UpdateDevice(sp, 0.0, dt_dev.data_ptr())
    dt = sp.GetTimeStep(2e-3, 3)
    RungeKutta(3).Update(sp, 0.0, dt)              # the host-dt path serves the stage times
    assert np.isfinite(sp.GetSolutionColumn()).all()


def test_malformed_mesh_is_an_error_not_a_crash():
    m = cases.small_mesh(seed=4)
    ctx = capi.Context(riemann=capi.HLLC, device=0)
    cells = m.cells.copy()
    cells[3, 5] = len(m.nodes) + 7                  # a node id beyond the coordinate table
    with pytest.raises(capi.McfdError, match="names node"):
        ctx.set_mesh(m.nodes, cells, len(cells), 0, m.bfaces, m.btags, None)
