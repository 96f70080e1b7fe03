"""The second p = 2 stage kernel (csrc/stage_nodes.cuh: persistent, one warp per node, bulk asynchronous copies one tile
ahead), chosen per context by MCFD_STAGE_KERNEL=nodes when the context is created: the same checks as the default
node-line kernel — golden vectors of the reference, every Riemann solver and boundary kind, all RK orders, a mesh of
several tiles per CTA, and the device-side ghost exchange."""
import os

import numpy as np
import pytest

import cases
import oracle_check as oc
from minicfd_b200 import capi, mesh as meshmod, peer
from minicfd_b200.part import Part, cell_points
from minicfd_b200.spatial import RungeKutta, Spatial

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def relerr(a, b):
    return float(np.abs(np.asarray(a) - np.asarray(b)).max() / np.abs(np.asarray(b)).max())


@pytest.fixture(autouse=True)
def node_kernel(monkeypatch):
    monkeypatch.setenv("MCFD_STAGE_KERNEL", "nodes")


@pytest.mark.parametrize("name,degree,scheme,riemann", [c for c in cases.GOLDEN_CASES if c[1] == 2 and c[2] in (0, 1)])
def test_golden_residual_and_rk3(name, degree, scheme, riemann):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    m = meshmod.HexMesh(nodes=g["nodes"], cells=g["cells"], bfaces=g["bfaces"], btags=g["btags"])
    sp = Spatial(Part(m), scheme=scheme, degree=degree, riemann=riemann, device=0)
    cases.apply_bcs_spatial(sp, cases.MIXED_BCS)
    sp.SetSolutionColumn(g["u"])
    sp.SetTime(cases.GOLDEN_T)
    assert relerr(sp.GetResidualColumn(), g["residual"]) < 1e-12
    RungeKutta(3).Update(sp, 0.0, cases.GOLDEN_DT)
    assert relerr(sp.GetSolutionColumn(), g["u_rk3"]) < 1e-12


@pytest.mark.parametrize("riemann", [capi.EXACT, capi.HLLC, capi.AUSM])
def test_affine_mesh_takes_the_per_tile_metric(riemann):
    """A sheared box (every cell the same parallelepiped, no face aligned with an axis): the affine instantiation, all
    boundary kinds of the golden cases, against the oracle."""
    m = meshmod.box(9, 8, 7, hi=(1.0, 0.8, 0.7))
    shear = np.array([[1.0, 0.3, 0.1], [0.0, 1.0, 0.2], [0.0, 0.0, 1.0]])
    m = meshmod.HexMesh(nodes=m.nodes @ shear.T, cells=m.cells, bfaces=m.bfaces, btags=m.btags)
    u = cases.noisy_state(cell_points(m.nodes, m.cells, 2), seed=3)
    o = oc.Oracle(m, 2, 0, riemann)
    cases.apply_bcs_oracle(o)
    sp = Spatial(Part(m), degree=2, riemann=riemann, device=0)
    cases.apply_bcs_spatial(sp, cases.MIXED_BCS)
    sp.SetSolutionColumn(u)
    sp.SetTime(0.3)
    assert relerr(sp.GetResidualColumn(), o.residual(u, 0.3)) < 1e-12
    sp.SetSolutionColumn(u)
    RungeKutta(3).Update(sp, 0.1, 1e-3)
    assert relerr(sp.GetSolutionColumn(), o.rk_update(u, 3, 0.1, 1e-3)) < 1e-12


@pytest.mark.parametrize("riemann", [capi.EXACT, capi.HLLC, capi.AUSM])
def test_many_tiles_per_cta_all_rk_orders(riemann):
    """6 000 cells = 188 tiles on 148 CTAs: every CTA's pipeline (prefetch of the next tile, both buffers, the last,
    partly filled tile) is exercised; walls all around."""
    m = meshmod.scramble(meshmod.box(20, 20, 15, hi=(2.0, 2.0, 1.5), warp=0.1, seed=5), seed=6)
    u = cases.noisy_state(cell_points(m.nodes, m.cells, 2), seed=3)
    o = oc.Oracle(m, 2, 0, riemann)
    for tag in range(6):
        o.set_bc(tag, 0)
    sp = Spatial(Part(m), degree=2, riemann=riemann, device=0)
    for tag in range(6):
        sp.SetInviscidWall(tag)
    for order in (1, 2, 3):
        sp.SetSolutionColumn(u)
        RungeKutta(order).Update(sp, 0.0, 1e-3)
        assert relerr(sp.GetSolutionColumn(), o.rk_update(u, order, 0.0, 1e-3)) < 1e-12


@pytest.mark.parametrize("world,warp", [(2, 0.1), (3, 0.1), (3, 0.0)])
def test_peer_step_on_several_parts(world, warp):
    """mcfd_peer_step_group with the node kernel: ghost states read from the receive stage, the producer thread polls
    for the neighbours' delivery, send cells packed by the kernel.  warp = 0: a mesh of parallelepipeds, i.e. the
    instantiation with the per-tile metric (what the library picks for such a mesh by itself)."""
    import torch
    m = meshmod.scramble(meshmod.box(12, 8, 6, hi=(1.2, 0.8, 0.6), warp=warp, seed=2), seed=3)
    parts = (np.arange(m.n_cells) * world // m.n_cells).astype(np.int32)
    u = cases.noisy_state(cell_points(m.nodes, m.cells, 2), seed=9)
    o = oc.Oracle(m, 2, 0, capi.HLLC)
    for tag in range(6):
        o.set_bc(tag, 0)
    P = [Part(m, parts, r, world) for r in range(world)]
    S = [Spatial(p, degree=2, riemann=capi.HLLC, device=0) for p in P]
    streams = [torch.cuda.Stream() for _ in S]
    for sp, p, st in zip(S, P, streams):
        for tag in range(6):
            sp.SetInviscidWall(tag)
        sp.ctx.set_stream(st.cuda_stream)
        sp.SetSolutionColumn(p.gather_from_global(u, 135))
    peer.connect_local(S, P)
    ref, t = u.copy(), 0.0
    for step in range(4):
        capi.peer_step_group([sp.ctx for sp in S], 3, t, 2e-3)
        dt = o.time_step(ref, 2e-3, 3)
        ref = o.rk_update(ref, 3, t, dt)
        t += dt
    for sp in S:
        sp.ctx.peer_check()
    got = np.zeros(m.n_cells * 135)
    for sp, p in zip(S, P):
        p.scatter_to_global(sp.GetSolutionColumn(), 135, got)
    assert relerr(got, ref) < 1e-12


def test_library_picks_the_kernel_by_the_mesh(monkeypatch):
    """Without MCFD_STAGE_KERNEL: a box whose vertex coordinates are not exact in binary (i * 10 / 7: the bilinear
    coefficients come out as 1e-16, not 0) is still a mesh of parallelepipeds and takes the node kernel with the
    per-tile metric; a warped mesh takes the node-line kernel; p = 3 and viscous contexts their own."""
    monkeypatch.delenv("MCFD_STAGE_KERNEL")
    box = meshmod.box(7, 5, 3, hi=(10.0, 10.0, 10.0))
    assert Spatial(Part(box), degree=2, riemann=capi.HLLC, device=0).ctx.stage_kernel() == "NodeStageKernel<affine>"
    warped = meshmod.box(7, 5, 3, hi=(10.0, 10.0, 10.0), warp=0.1, seed=1)
    assert Spatial(Part(warped), degree=2, riemann=capi.HLLC, device=0).ctx.stage_kernel() == "LineStageKernel<general>"
    assert Spatial(Part(box), degree=3, riemann=capi.HLLC, device=0).ctx.stage_kernel() == "StageKernel"
    monkeypatch.setenv("MCFD_STAGE_KERNEL", "lines")
    assert Spatial(Part(box), degree=2, riemann=capi.HLLC, device=0).ctx.stage_kernel() == "LineStageKernel<affine>"
