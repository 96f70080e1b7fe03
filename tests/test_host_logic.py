"""CPU-only tests of everything around the kernels: the C-ABI library loads and exports what the
header declares, the Part bookkeeping matches the reference's numbering rules, the kernel *source*
(compiled for the host, tests/host_emulation) reproduces the oracle, and the halo exchange logic
works across two gloo ranks."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest

import cases
import emuapi
import oracle_check as oc
import refapi
from minicfd_b200 import capi, fields, mesh as meshmod
from minicfd_b200.part import Part, cell_points, dual_graph

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def relerr(a, b):
    return np.abs(a - b).max() / np.abs(b).max()


def test_library_exports_every_declared_symbol():
    """The C ABI loads without a GPU and exports each function of include/minicfd_b200.h."""
    header = open(os.path.join(ROOT, "include", "minicfd_b200.h")).read()
    declared = set(re.findall(r"\b(mcfd_[a-z_0-9]+)\s*\(", header)) - {"mcfd_time_callback"}
    assert declared == set(capi.SYMBOLS), declared ^ set(capi.SYMBOLS)
    lib = ctypes.CDLL(capi.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), name


def test_no_cpu_fallback_without_gpu():
    """Without a device mcfd_create must fail loudly (no fallback); with one it must succeed."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(capi.McfdError, match="no CUDA device|CUDA"):
        capi.Context()


def test_product_never_touches_the_checkers():
    """Nothing under minicfd_b200/ or include/ may import or link oracle/ or tests/ code."""
    bad = re.compile(r"(import|from)\s+(oracle|refapi|emuapi|oracle_check|cases)\b|libminicfd_ref|"
                     r"libminicfd_oracle|libmcfd_emu|#include\s+\"[^\"]*(oracle|tests)/|dlopen")
    for base in ("minicfd_b200", "include"):
        for dirpath, _, files in os.walk(os.path.join(ROOT, base)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".hpp", ".h", ".cpp")):
                    text = open(os.path.join(dirpath, f)).read()
                    assert not bad.search(text), (f, bad.search(text).group(0))


@pytest.mark.skipif(not refapi.available(2), reason="oracle/_ref not built")
def test_part_numbering_matches_reference_rules():
    m = meshmod.scramble(meshmod.box(4, 3, 3, warp=0.1), seed=1)
    parts = np.random.default_rng(0).integers(0, 3, size=m.n_cells).astype(np.int32)
    r = refapi.Ref(m, degree=2, cell_parts=parts, n_parts=3)
    pairs = dual_graph(m)
    for rank in range(3):
        p = Part(m, parts, rank, 3, pairs=pairs)
        assert np.array_equal(r.cell_global_ids(rank), p.global_ids)      # inner, inter, then ghosts
        assert r.count_local(rank) == p.n_local and r.count_ghost(rank) == p.n_ghost
        sp, sc, scell = r.halo(rank, 1)
        rp, rc, rcell = r.halo(rank, 0)
        assert np.array_equal(sp, p.neighbours) and np.array_equal(rp, p.neighbours)
        assert np.array_equal(scell, np.concatenate(p.send_cells))
        assert np.array_equal(rcell, np.concatenate(p.recv_cells))


def test_block_partition_and_periodic_graph():
    m = meshmod.box(4, 4, 2)
    parts = meshmod.block_partition(m, 2, 2, 1)
    assert sorted(np.bincount(parts)) == [8, 8, 8, 8]
    assert len(dual_graph(meshmod.box(3, 3, 3, periodic=True))) == 81
    assert len(dual_graph(meshmod.box(3, 3, 3))) == 54


@pytest.mark.parametrize("name,degree,scheme,riemann", cases.GOLDEN_CASES)
def test_kernel_source_on_host_matches_golden(name, degree, scheme, riemann):
    """The device code of the stage kernel, run thread by thread on the host, against the reference's
    golden residual and RK3 state (mixed boundary kinds, scrambled warped mesh)."""
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    m = meshmod.HexMesh(nodes=g["nodes"], cells=g["cells"], bfaces=g["bfaces"], btags=g["btags"])
    holder, hf, tags, pts = emuapi.boundary(m, degree)
    def bvals(t):
        out = np.zeros(pts.shape[:2] + (5,))
        for tag, kind in cases.MIXED_BCS.items():
            sel = tags == tag
            out[sel] = cases.bc_function(kind)(pts[sel], t)
        return out
    u = g["u"]
    res = emuapi.stage(m, degree, scheme, riemann, u, tag_kinds=cases.MIXED_BCS, bvals=bvals(cases.GOLDEN_T))
    assert relerr(res, g["residual"]) < 1e-12
    dt = cases.GOLDEN_DT
    kw = dict(tag_kinds=cases.MIXED_BCS, dt=dt)
    u1 = emuapi.stage(m, degree, scheme, riemann, u, bvals=bvals(0.0), mode=1, **kw)
    u2 = emuapi.stage(m, degree, scheme, riemann, u1, bvals=bvals(dt), column_u0=u, mode=2, **kw)
    u3 = emuapi.stage(m, degree, scheme, riemann, u2, bvals=bvals(dt / 2), column_u0=u, mode=3, **kw)
    assert relerr(u3, g["u_rk3"]) < 1e-12


@pytest.mark.parametrize("name,degree,riemann", cases.VISCOUS_GOLDEN_CASES)
def test_viscous_kernel_source_on_host_matches_golden(name, degree, riemann):
    """Trace + viscous stage device code, run on the host, against the reference's Navier-Stokes vectors."""
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    m = meshmod.HexMesh(nodes=g["nodes"], cells=g["cells"], bfaces=g["bfaces"], btags=g["btags"])
    holder, hf, tags, pts = emuapi.boundary(m, degree)
    def bvals(t):
        out = np.zeros(pts.shape[:2] + (5,))
        for tag, kind in cases.VISCOUS_BCS.items():
            sel = tags == tag
            out[sel] = cases.viscous_bc_function(kind)(pts[sel], t)
        return out
    u = g["u"]
    kw = dict(tag_kinds=cases.VISCOUS_BCS, viscous=cases.VISCOUS)
    res = emuapi.stage(m, degree, 0, riemann, u, bvals=bvals(cases.GOLDEN_T), **kw)
    assert relerr(res, g["residual"]) < 1e-12
    dt = cases.GOLDEN_DT
    u1 = emuapi.stage(m, degree, 0, riemann, u, bvals=bvals(0.0), mode=1, dt=dt, **kw)
    u2 = emuapi.stage(m, degree, 0, riemann, u1, bvals=bvals(dt), column_u0=u, mode=2, dt=dt, **kw)
    u3 = emuapi.stage(m, degree, 0, riemann, u2, bvals=bvals(dt / 2), column_u0=u, mode=3, dt=dt, **kw)
    assert relerr(u3, g["u_rk3"]) < 1e-12


@pytest.mark.parametrize("name,degree,riemann", cases.WENO_GOLDEN_CASES)
def test_weno_kernel_source_on_host_matches_golden(name, degree, riemann):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    m = meshmod.HexMesh(nodes=g["nodes"], cells=g["cells"], bfaces=g["bfaces"], btags=g["btags"])
    assert relerr(emuapi.weno_lazy(m, degree, g["u"], *cases.WENO), g["limited"]) < 1e-12


@pytest.mark.parametrize("name,degree,riemann", cases.ISO_GOLDEN_CASES)
def test_cell_viscosity_kernel_source_on_host(name, degree, riemann):
    """Uniform nu against the reference's vectors; a per-cell, per-component table against the C oracle."""
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    m = meshmod.HexMesh(nodes=g["nodes"], cells=g["cells"], bfaces=g["bfaces"], btags=g["btags"])
    holder, hf, tags, pts = emuapi.boundary(m, degree)
    bv = np.zeros(pts.shape[:2] + (5,))
    for tag, kind in cases.VISCOUS_BCS.items():
        sel = tags == tag
        bv[sel] = cases.viscous_bc_function(kind)(pts[sel], cases.GOLDEN_T)
    kw = dict(tag_kinds=cases.VISCOUS_BCS, bvals=bv, viscous=cases.ISO_VISCOUS)
    uniform = np.full((m.n_cells, 5), cases.ISO_VISCOUS[2])
    assert relerr(emuapi.stage(m, degree, 0, riemann, g["u"], cell_nu=uniform, **kw), g["residual"]) < 1e-12
    nu = cases.cell_viscosities(m.n_cells)
    o = oc.Oracle(m, degree, 6, riemann)
    cases.apply_viscous_oracle(o, vis=cases.ISO_VISCOUS)
    o.set_cell_viscosity(nu)
    assert relerr(emuapi.stage(m, degree, 0, riemann, g["u"], cell_nu=nu, **kw), o.residual(g["u"], cases.GOLDEN_T)) < 1e-12


@pytest.mark.parametrize("name,degree,riemann", cases.ENERGY_GOLDEN_CASES)
def test_energy_viscosity_kernel_source_on_host_matches_golden(name, degree, riemann):
    """Probe launches of the trace + stage kernels -> damping matrices; ViscosityCell -> viscosities and min_dt_."""
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    m = meshmod.HexMesh(nodes=g["nodes"], cells=g["cells"], bfaces=g["bfaces"], btags=g["btags"])
    damping, nu, min_dt = emuapi.energy_viscosity(m, degree, cases.ENERGY_BETAS, g["u"])
    assert relerr(damping[:4], g["damping"]) < 1e-12
    assert np.array_equal(nu > 0, g["nu"] > 0)
    assert relerr(nu, g["nu"]) < 1e-12
    cfl = {2: 0.209, 3: 0.130}[degree] * 2
    assert min(1.0, cfl * min_dt) == pytest.approx(float(g["dt"]), rel=1e-13)


@pytest.mark.parametrize("name,degree,riemann,viscous", cases.LOCAL_GOLDEN_CASES)
def test_klocal_kernel_source_on_host_matches_golden(name, degree, riemann, viscous):
    """The device keeps u = U / det; only the traces interpolate U.  Residual columns are det du/dt."""
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    m = meshmod.HexMesh(nodes=g["nodes"], cells=g["cells"], bfaces=g["bfaces"], btags=g["btags"])
    det = g["det"]
    u = (g["u"].reshape(det.shape + (5,)) / det[:, :, None]).reshape(-1)
    holder, hf, tags, pts = emuapi.boundary(m, degree)
    kinds = cases.VISCOUS_BCS if viscous else cases.MIXED_BCS
    bv = np.zeros(pts.shape[:2] + (5,))
    for tag, kind in kinds.items():
        sel = tags == tag
        f = cases.viscous_bc_function(kind) if viscous else cases.bc_function(kind)
        bv[sel] = f(pts[sel], cases.GOLDEN_T)
    kw = dict(tag_kinds=kinds, bvals=bv)
    if viscous:
        kw["viscous"] = list(cases.VISCOUS) + [1.0]
    res = emuapi.stage(m, degree, 0, riemann, u, **kw)
    assert relerr((res.reshape(det.shape + (5,)) * det[:, :, None]).reshape(-1), g["residual"]) < 1e-12


@pytest.mark.parametrize("name,degree,riemann", cases.DGM_GOLDEN_CASES)
def test_dg_general_kernel_source_on_host_matches_golden(name, degree, riemann):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    m = meshmod.HexMesh(nodes=g["nodes"], cells=g["cells"], bfaces=g["bfaces"], btags=g["btags"])
    gp, qp = emuapi.dgm(m, degree, want_points=True)
    assert relerr(emuapi.dgm(m, degree, riemann, values=fields.smooth(1.0)(qp, 0.0)), g["projected"]) < 1e-12
    holder, hf, tags, _ = emuapi.boundary(m, degree)
    def bvals(t):
        out = np.zeros(gp.shape[:2] + (5,))
        for tag, kind in cases.MIXED_BCS.items():
            sel = tags == tag
            out[sel] = cases.bc_function(kind)(gp[sel], t)
        return out
    kw = dict(tag_kinds=cases.MIXED_BCS)
    assert relerr(emuapi.dgm(m, degree, riemann, g["u"], bvals=bvals(cases.GOLDEN_T), **kw), g["residual"]) < 1e-12
    dt = cases.GOLDEN_DT
    u1 = emuapi.dgm(m, degree, riemann, g["u"], bvals=bvals(0.0), mode=1, dt=dt, **kw)
    u2 = emuapi.dgm(m, degree, riemann, u1, bvals=bvals(dt), column_u0=g["u"], mode=2, dt=dt, **kw)
    u3 = emuapi.dgm(m, degree, riemann, u2, bvals=bvals(dt / 2), column_u0=g["u"], mode=3, dt=dt, **kw)
    assert relerr(u3, g["u_rk3"]) < 1e-12


@pytest.mark.parametrize("name,degree,riemann", cases.WENO_EIGEN_GOLDEN_CASES)
def test_weno_eigen_kernel_source_on_host_matches_golden(name, degree, riemann):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    m = meshmod.HexMesh(nodes=g["nodes"], cells=g["cells"], bfaces=g["bfaces"], btags=g["btags"])
    assert relerr(emuapi.weno_lazy(m, degree, g["u"], *cases.WENO, eigen=True), g["limited"]) < 1e-12


@pytest.mark.skipif(not refapi.available(2), reason="oracle/_ref not built (needs /root/reference)")
def test_weno_eigen_two_parts_need_the_global_ids_for_the_face_frames():
    """The tangent of a face frame follows the holder's vertex order, and the reference picks the holder of a
    local face by global cell id (shuffler.hpp:417-420): with mcfd_set_cell_global_ids the parts match."""
    m = meshmod.scramble(meshmod.box(6, 4, 2, warp=0.1, seed=2), seed=3)
    parts = (np.arange(m.n_cells) * 2 // m.n_cells).astype(np.int32)
    u = cases.shocked_state(cell_points(m.nodes, m.cells, 2))
    r = refapi.Ref(m, degree=2, scheme=refapi.FR_WENO_EIGEN, riemann=1, cell_parts=parts, n_parts=2)
    P = [Part(m, parts, k, 2) for k in range(2)]
    r.set_solution_all([p.gather_from_global(u, 135) for p in P])
    for k, p in enumerate(P):
        col = u.reshape(-1, 135)[p.global_ids].reshape(-1)
        class Sub:
            nodes, periodic = m.nodes, None
        got = emuapi.weno_lazy(Sub, 2, col, n_local=p.n_local, n_ghost=p.n_ghost, cells=p.cell_nodes,
                               bfaces=p.bfaces, btags=p.btags, eigen=True, global_ids=p.global_ids)
        assert relerr(got, r.get_solution(k)) < 1e-12


@pytest.mark.skipif(not refapi.available(2), reason="oracle/_ref not built (needs /root/reference)")
def test_weno_two_batches_match_the_reference_parts():
    """Inner cells first, then inter cells borrowing the already limited inner cells and the UNLIMITED ghost
    values (limiter/reconstruct.hpp:46-51): the result depends on the partition, and must match part by part."""
    m = meshmod.scramble(meshmod.box(6, 4, 2, warp=0.1, seed=2), seed=3)
    parts = (np.arange(m.n_cells) * 2 // m.n_cells).astype(np.int32)
    u = cases.noisy_state(cell_points(m.nodes, m.cells, 2), amp=1e-2)
    r = refapi.Ref(m, degree=2, scheme=refapi.FR_WENO_LAZY, riemann=1, cell_parts=parts, n_parts=2)
    P = [Part(m, parts, k, 2) for k in range(2)]
    r.set_solution_all([p.gather_from_global(u, 135) for p in P])
    for k, p in enumerate(P):
        col = u.reshape(-1, 135)[p.global_ids].reshape(-1)
        class Sub:
            nodes, periodic = m.nodes, None
        got = emuapi.weno_lazy(Sub, 2, col, n_local=p.n_local, n_ghost=p.n_ghost, cells=p.cell_nodes,
                               bfaces=p.bfaces, btags=p.btags)
        assert relerr(got, r.get_solution(k)) < 1e-12


def test_kernel_source_periodic_equals_embedded_box():
    """A periodic box must give the same residuals as the interior of a larger box fed with the same
    2*pi-periodic field (GPU-only periodic variant of BASELINE config 2)."""
    L = 2 * np.pi
    m = meshmod.box(4, 4, 4, hi=(L, L, L), periodic=True)
    u = fields.smooth(1.0)(cell_points(m.nodes, m.cells, 2), 0.0).reshape(-1)
    res = emuapi.stage(m, 2, 0, 1, u)
    h = L / 4
    mb = meshmod.box(6, 6, 6, lo=(-h,) * 3, hi=(L + h,) * 3)
    ub = fields.smooth(1.0)(cell_points(mb.nodes, mb.cells, 2), 0.0).reshape(-1)
    resb = emuapi.stage(mb, 2, 0, 1, ub).reshape(6, 6, 6, 135)
    assert relerr(res, resb[1:5, 1:5, 1:5].reshape(-1)) < 1e-12


def test_kernel_source_with_ghost_cells_matches_single_part():
    """A part with ghost cells (fed with the neighbours' coefficients) reproduces the single-part
    residual on its local cells: the ghost-face path of fem.hpp:374-380."""
    m = cases.small_mesh(seed=9)
    parts = (np.arange(m.n_cells) % 2).astype(np.int32)
    u = cases.noisy_state(cell_points(m.nodes, m.cells, 2))
    o = oc.Oracle(m, 2, 0, 1)
    for tag in range(6):
        o.set_bc(tag, 0)
    full = o.residual(u, 0.0).reshape(-1, 135)
    for rank in range(2):
        p = Part(m, parts, rank, 2)
        col = u.reshape(-1, 135)[p.global_ids].reshape(-1)
        class Sub:  # the part's own view of the mesh
            nodes, periodic = m.nodes, None
        res = emuapi.stage(Sub, 2, 0, 1, col, n_local=p.n_local, n_ghost=p.n_ghost, cells=p.cell_nodes,
                           bfaces=p.bfaces, btags=p.btags)
        assert relerr(res.reshape(-1, 135), full[p.local_global]) < 1e-12


_GLOO_WORKER = r'''
import os, sys
sys.path.insert(0, sys.argv[1])
import numpy as np, torch, torch.distributed as dist
from minicfd_b200 import mesh as meshmod, halo
from minicfd_b200.part import Part
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo", rank=rank, world_size=world)
m = meshmod.scramble(meshmod.box(4, 3, 2, warp=0.1), seed=2)
parts = (np.arange(m.n_cells) * world // m.n_cells).astype(np.int32)
p = Part(m, parts, rank, world)
K = 7
field = np.arange(m.n_cells * K, dtype=np.float64).reshape(m.n_cells, K) * 1.5   # global "coefficients"
ex = halo.for_part(p, K)
send = torch.from_numpy(field[p.global_ids[np.concatenate(p.send_cells)]].reshape(-1).copy())
recv = torch.zeros(int(sum(len(c) for c in p.recv_cells)) * K, dtype=torch.float64)
ex.bind(send, recv)
ex.exchange()
expect = field[p.global_ids[np.concatenate(p.recv_cells)]].reshape(-1)
assert np.array_equal(recv.numpy(), expect), "ghost data mismatch"
dt = ex.allreduce_min(1.0 + rank)
assert dt == 1.0
dist.destroy_process_group()
print("OK", rank)
'''


def test_halo_exchange_two_gloo_ranks(tmp_path):
    """world_size-2 gloo run of the exchange used for N > 1 GPUs (pack order = send_cell_ptrs_, unpack
    order = recv_cell_ptrs_; part.hpp:1242-1314) and of the dt all-reduce."""
    script = tmp_path / "worker.py"
    script.write_text(_GLOO_WORKER)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", "29541", str(script), ROOT]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout + out.stderr
    assert out.stdout.count("OK") == 2


@pytest.mark.parametrize("degree,c_next", [(2, 1.0), (3, 0.3)])
def test_kernel_source_fr_general(degree, c_next):
    """General Vincent correction in the kernel source vs the oracle, mixed boundaries."""
    m = cases.small_mesh(seed=8)
    u = cases.noisy_state(cell_points(m.nodes, m.cells, degree), seed=2)
    o = oc.Oracle(m, degree, 2, 1, c_next=c_next)
    cases.apply_bcs_oracle(o)
    holder, hf, tags, pts = emuapi.boundary(m, degree)
    bvals = np.zeros(pts.shape[:2] + (5,))
    for tag, kind in cases.MIXED_BCS.items():
        bvals[tags == tag] = cases.bc_function(kind)(pts[tags == tag], 0.2)
    got = emuapi.stage(m, degree, 2, 1, u, tag_kinds=cases.MIXED_BCS, bvals=bvals, c_next=c_next)
    assert relerr(got, o.residual(u, 0.2)) < 1e-12


@pytest.mark.parametrize("world", [1, 4])
def test_nompi_allreduce_between_processes(tmp_path, world):
    """include/mini/b200/nompi/mpi.h + comm.hpp: MPI_Allreduce(MIN | MAX | SUM), MPI_Barrier, rank and size between
    real processes over POSIX shared memory — what the C++ drivers use where the reference calls MPI
    (demo/euler/sourceless.cpp:60-66, 194)."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = tmp_path / "comm_check"
    subprocess.check_call(["g++", "-std=c++20", "-O1", "-I", os.path.join(root, "include"),
                           "-I", os.path.join(root, "include", "mini", "b200", "nompi"),
                           os.path.join(root, "tests", "host_emulation", "comm_check.cpp"), "-o", str(exe), "-lrt"])
    procs = [subprocess.Popen([str(exe)], env=dict(os.environ, MCFD_SESSION=f"c{os.getpid()}_{world}", MCFD_WORLD_SIZE=str(world),
                                                   MCFD_RANK=str(r)), stdout=subprocess.PIPE, text=True) for r in range(world)]
    for r, p in enumerate(procs):
        out = p.communicate(timeout=120)[0]
        assert p.returncode == 0 and f"rank {r} of {world} ok" in out, out


@pytest.mark.parametrize("name,degree,scheme,riemann", [c for c in cases.GOLDEN_CASES if c[1] == 2 and c[2] in (0, 1)])
def test_node_kernel_source_on_host_matches_golden(name, degree, scheme, riemann):
    """stage_nodes.cuh (one warp per node, persistent; p = 2 lumped Lobatto schemes), its phase functions run thread by
    thread on the host, against the reference's golden residual and RK3 state."""
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    m = meshmod.HexMesh(nodes=g["nodes"], cells=g["cells"], bfaces=g["bfaces"], btags=g["btags"])
    holder, hf, tags, pts = emuapi.boundary(m, degree)
    def bvals(t):
        out = np.zeros(pts.shape[:2] + (5,))
        for tag, kind in cases.MIXED_BCS.items():
            sel = tags == tag
            out[sel] = cases.bc_function(kind)(pts[sel], t)
        return out
    u = g["u"]
    kw = dict(tag_kinds=cases.MIXED_BCS, line_cells=1)
    res = emuapi.stage(m, degree, scheme, riemann, u, bvals=bvals(cases.GOLDEN_T), **kw)
    assert relerr(res, g["residual"]) < 1e-12
    dt = cases.GOLDEN_DT
    u1 = emuapi.stage(m, degree, scheme, riemann, u, bvals=bvals(0.0), mode=1, dt=dt, **kw)
    u2 = emuapi.stage(m, degree, scheme, riemann, u1, bvals=bvals(dt), column_u0=u, mode=2, dt=dt, **kw)
    u3 = emuapi.stage(m, degree, scheme, riemann, u2, bvals=bvals(dt / 2), column_u0=u, mode=3, dt=dt, **kw)
    assert relerr(u3, g["u_rk3"]) < 1e-12


@pytest.mark.parametrize("name,degree,riemann", cases.ENERGY_NS_GOLDEN_CASES)
def test_energy_viscosity_over_navier_stokes_kernel_source_on_host(name, degree, riemann):
    """EnergyBasedViscosity over a DIFFUSIVE base (demo/laminar/shocked.hpp:100-111; viscosity.hpp:76-139): the stage
    kernel's kDiffNsIsotropic instantiation with the reference's viscosities against the reference's residual (viscous
    boundary set-up with no-slip walls), and the rule's set-up, which the base does not enter: the damping matrices
    take the mass row only, where Navier-Stokes adds nothing."""
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    m = meshmod.HexMesh(nodes=g["nodes"], cells=g["cells"], bfaces=g["bfaces"], btags=g["btags"])
    holder, hf, tags, pts = emuapi.boundary(m, degree)
    bv = np.zeros(pts.shape[:2] + (5,))
    for tag, kind in cases.VISCOUS_BCS.items():
        sel = tags == tag
        bv[sel] = cases.viscous_bc_function(kind)(pts[sel], cases.GOLDEN_T)
    res = emuapi.stage(m, degree, 0, riemann, g["u"], tag_kinds=cases.VISCOUS_BCS, bvals=bv,
                       viscous=list(cases.VISCOUS) + [2.0], cell_nu=g["nu"])    # 2: the viscosities on top of Navier-Stokes
    assert relerr(res, g["residual"]) < 1e-12
    damping, nu, min_dt = emuapi.energy_viscosity(m, degree, cases.VISCOUS[:2], g["u"])
    assert relerr(damping[:4], g["damping"]) < 1e-12
    assert np.array_equal(nu > 0, g["nu"] > 0) and relerr(nu, g["nu"]) < 1e-12


def test_shocked_type_composition_compiles_against_the_shim():
    """demo/laminar/shocked.hpp:43-111 spelled against include/mini: EnergyBasedViscosity over ConvectionDiffusion<Euler,
    Direct<NavierStokes>> compiles and selects the library's Navier-Stokes + per-cell-viscosity path."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    subprocess.check_call(["g++", "-std=c++20", "-fsyntax-only", "-I", os.path.join(root, "include"),
                           "-I", os.path.join(root, "include", "mini", "b200", "nompi"),
                           os.path.join(root, "tests", "drivers", "shocked_types.cpp")])
