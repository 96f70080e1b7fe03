#!/usr/bin/env python
"""bench.py — BASELINE.json's metric on BASELINE.json's config.

Metric: DOF-updates/s per RK stage (p=2 Euler, hexa).  One *step* = one `RungeKutta<3>::Update`
(3 stages, each one residual + stage update) preceded by `GetTimeStep`, as in the reference's time
loop (demo/euler/sourceless.cpp:187-215).  value = 3 * n_dof_total * steps / time.

Workload (config 2 of BASELINE.json): synthetic structured box of 42^3 Hexahedron8 cells per GPU
(10 001 880 DOF at p=2, K=5), FR-Lobatto p=2, HLLC, isentropic-vortex initial state, periodic
wrap (the GPU-only performance variant), RK3.  N > 1: weak scaling, one 42^3
block per GPU in a px*py*pz block partition; the ghost rows travel by stores into the neighbours' memory over
NVLink (mcfd_peer_*; `--halo nccl`: the pack -> NCCL send/recv -> unpack path).

Arms:
  (default)          our CUDA path.  `value`: state resident in HBM.  `e2e`: the same step through the
                     System interface with HOST columns: SetSolutionColumn(pinned host) -> GetTimeStep
                     -> Update -> GetSolutionColumn(pinned host) every step.
  --impl reference   the reference's own CPU code (oracle/_ref: unmodified headers behind
                     oracle/ref_harness.cpp; falls back to the C restatement when it was not shipped),
                     all host threads, on a bounded sample of the workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

B_ALG_P2 = 24.0 + 192.0 / 135.0   # algorithmic bytes per DOF-update
SCHEME_NAME = {"fr": "FR-Lobatto", "dg": "DG-Lobatto", "dgm": "dg::General (modal Projection, Gauss-Legendre)"}
RIEMANN = {"exact": 0, "hllc": 1, "ausm": 2}


def stage_kernel_name(args):
    """The kernel behind a stage of this configuration (libmcfd.cu: mcfd_create's use_lines / LaunchStage)."""
    q = args.degree + 1
    if args.scheme == "dgm":
        return f"mcfd::DgmStageKernel (degree {args.degree})"
    if args.viscous is None and not args.energy_viscosity and q == 3:
        # the library's choice (mcfd_set_mesh): the node kernel on meshes of parallelepipeds, unless told otherwise
        which = getattr(args, "stage_kernel", None) or os.environ.get("MCFD_STAGE_KERNEL") or ("nodes" if args.warp == 0.0 else "lines")
        if which == "nodes":
            return f"mcfd::NodeStageKernel<{args.riemann}, {'affine' if args.warp == 0.0 else 'general'}> (persistent, one warp per node)"
        return f"mcfd::LineStageKernel<{q}, {args.riemann}, 16 cells per CTA, {'affine' if args.warp == 0.0 else 'general'}>"
    return f"mcfd::StageKernel<{q}, {args.riemann}>"


def traffic_shape(args, world):
    """Key of profiles/stage_kernel_traffic.json: the configuration whose stage kernel the ncu capture profiled."""
    plain = (args.viscous is None and not args.limiter and not args.energy_viscosity and args.warp == 0.0
             and stage_kernel_name(args).startswith("mcfd::NodeStageKernel"))
    return f"{args.scheme} p{args.degree} {args.riemann} {args.cells}^3 cells per GPU {'inviscid' if plain else 'other'} n{world}"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cells", type=int, default=42, help="cells per direction per GPU")
    ap.add_argument("--degree", type=int, default=2)
    ap.add_argument("--riemann", default="hllc", choices=list(RIEMANN))
    ap.add_argument("--scheme", default="fr", choices=["fr", "dg", "dgm"],
                    help="fr: FR-Lobatto (headline), dg: DGSEM-Lobatto, dgm: dg::General on the modal Projection polynomial")
    ap.add_argument("--cpu-cells", type=int, default=0, help="cells per direction of the CPU sample (0 = auto)")
    ap.add_argument("--viscous", type=float, default=None, metavar="NU",
                    help="run the Navier-Stokes + direct-DG path with this kinematic viscosity (not the headline workload)")
    ap.add_argument("--limiter", action="store_true",
                    help="WithLimiter<.., weno::Lazy(0.001, 1e-6)>: the WENO limiter after every stage (not the headline workload)")
    ap.add_argument("--energy-viscosity", action="store_true",
                    help="WithViscosity<.., EnergyBasedViscosity>: artificial viscosity updated after every stage "
                         "(not the headline workload; single GPU)")
    ap.add_argument("--frame-output", action="store_true",
                    help="also time the frame output's sampling step (mcfd_sample_vtk: values at the 27 points of every "
                         "cell's VTK hexahedron, copied to the host), next to the reference's vtk::Writer::Prepare on the CPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-features", action="store_true",
                    help="skip the short runs of the other configurations (viscous, limiter, ..., BASELINE config 5)")
    ap.add_argument("--halo", default="peer", choices=["peer", "nccl"],
                    help="N > 1: ghost exchange by device-to-device copies into the neighbours' memory (mcfd_peer_*, the default for the "
                         "inviscid nodal schemes) or pack -> NCCL send/recv -> unpack issued from Python")
    ap.add_argument("--warp", type=float, default=0.0,
                    help="displace the interior mesh nodes by this fraction of h (non-affine cells: the general metric path)")
    ap.add_argument("--stage-kernel", default=None, choices=["lines", "nodes"],
                    help="p = 2 inviscid nodal schemes: the node-line kernel or the persistent one-warp-per-node kernel "
                         "(csrc/stage_nodes.cuh); sets MCFD_STAGE_KERNEL for the contexts of this run.  Default: the "
                         "library's choice (nodes on meshes of parallelepipeds, lines elsewhere)")
    ap.add_argument("--no-graph", action="store_true", help="launch every kernel from the host instead of replaying a CUDA graph")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------------
# clocks
# ---------------------------------------------------------------------------------------------------
class ClockSampler:
    """`nvidia-smi -lms` running beside the timed region (B200_PROFILING.md's clocks line); samples
    taken between mark_begin() and mark_end() are the ones "under load"."""
    FIELDS = ("timestamp,clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0, period_ms=20):
        self.index, self.period_ms, self.proc, self.t0, self.t1 = index, period_ms, None, None, None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", str(self.period_ms)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            time.sleep(0.25)   # let it produce its first samples
        except Exception:
            self.proc = None
        return self

    def mark_begin(self):
        self.t0 = time.time()

    def mark_end(self):
        self.t1 = time.time()

    def __exit__(self, *a):
        if self.proc is not None:
            time.sleep(0.05)
            self.proc.terminate()
            try:
                self.out = self.proc.communicate(timeout=5)[0]
            except Exception:
                self.proc.kill()
                self.out = ""

    def summary(self):
        import datetime
        rows = []
        for line in getattr(self, "out", "").splitlines():
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                ts = datetime.datetime.strptime(parts[0], "%Y/%m/%d %H:%M:%S.%f").timestamp()
                rows.append((ts, float(parts[1]), float(parts[2]), parts[3:7]))
            except ValueError:
                continue
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"], "samples": 0}
        load = [r for r in rows if self.t0 is not None and self.t0 - 0.02 <= r[0] <= self.t1 + 0.02] or rows
        sm = sorted(r[1] for r in load)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3][i].lower().startswith("active") for r in load)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": rows[0][2], "reasons": reasons,
                "samples_under_load": len(load), "samples": len(rows)}


# ---------------------------------------------------------------------------------------------------
# reference arm / cpu baseline
# ---------------------------------------------------------------------------------------------------
def cpu_run(args, steps, warmup, cells=None):
    """Times the reference's CPU implementation (oracle/_ref) — or the C restatement when the
    reference binary is absent — on a bounded sample: an n^3 box of the same cells, same scheme, same
    Riemann solver, vortex state, inviscid walls instead of the periodic wrap (the reference has no
    periodic boundary), block-partitioned over all host cores (one thread per part, in-process halo)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_check as oc
    import refapi
    from minicfd_b200 import mesh as meshmod
    from minicfd_b200.part import cell_points
    cores = os.cpu_count() or 1
    scheme = {"fr": 0, "dg": 1, "dgm": 9}[args.scheme]
    if args.viscous is not None:
        scheme = 3
    if args.limiter:
        scheme = 4
    if args.energy_viscosity:
        scheme = 7
    riemann = RIEMANN[args.riemann]
    kf = 5 * (args.degree + 1) ** 3
    if args.scheme == "dgm":
        kf = 5 * (10 if args.degree == 2 else 20)
    if refapi.available(args.degree):
        kind = "reference"
        n_parts = max(1, min(cores, 64))
        if args.energy_viscosity:
            n_parts = 1        # EnergyBasedViscosity keeps its state in class statics: one spatial object per process
        # ~8e6 DOF-updates/s/core; aim at ~1.5 s per step
        n = cells or args.cpu_cells or int(round((n_parts * 8e6 * 1.5 / (3 * kf)) ** (1 / 3)))
        n = max(6, min(n, 48))
        if args.energy_viscosity:
            n = cells or args.cpu_cells or 12      # BuildDampingMatrices costs ~5 ms per cell
        m = meshmod.box(n, n, n, hi=(10.0, 10.0, 10.0))
        px, py, pz = meshmod.factor3(n_parts)
        parts = meshmod.block_partition(m, px, py, pz)
        r = refapi.Ref(m, degree=args.degree, scheme=scheme, riemann=riemann, cell_parts=parts, n_parts=n_parts,
                       viscous=(2.0, 1.0 / 12, 0.0, 0.708) if args.energy_viscosity else None)
        if args.viscous is not None:
            r.set_viscous(2.0, 1.0 / 12, args.viscous, 0.708)
        for tag in range(6):
            r.set_bc(tag, refapi.WALL)
        r.approximate(refapi.F_VORTEX, [5.0, 5.0, 1.0, 1.0, 0.0, 5.0])
        dt = r.time_step(1.0, 3)
        step = lambda: r.rk_update(3, 0.0, r.time_step(1.0, 3), 1)  # noqa: E731
    else:
        kind = "port"
        n_parts = 1
        n = cells or args.cpu_cells or 12
        m = meshmod.box(n, n, n, hi=(10.0, 10.0, 10.0))
        o = oc.Oracle(m, args.degree, 0 if args.limiter else scheme, riemann)
        if args.limiter:
            o.set_limiter(0.001, 1e-6)
        if args.viscous is not None:
            o.set_viscous(2.0, 1.0 / 12, args.viscous, 0.708)
        for tag in range(6):
            o.set_bc(tag, 0)
        state = {"u": o.approximate(oc.F_VORTEX, [5.0, 5.0, 1.0, 1.0, 0.0, 5.0])}

        def step():
            dt = o.time_step(state["u"], 1.0, 3)
            state["u"] = o.rk_update(state["u"], 3, 0.0, dt)
    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt_wall = time.perf_counter() - t0
    n_dof = m.n_cells * kf
    value = 3.0 * n_dof * steps / dt_wall
    sample = (f"{n}^3 Hexahedron8 cells ({n_dof} DOF), {SCHEME_NAME[args.scheme]} p={args.degree}, "
              f"{args.riemann}, RK3 incl. GetTimeStep, inviscid walls, {n_parts} part(s)/thread(s), "
              f"{steps} step(s) after {warmup} warm-up")
    return {"value": value, "unit": "DOF-updates/s", "cores": n_parts, "kind": kind, "sample": sample,
            "host_cores": cores, "ms_per_step": 1e3 * dt_wall / steps}


def metric_name(args):
    what = "Euler" if args.viscous is None else "Navier-Stokes"
    return f"DOF-updates/s per RK stage (p={args.degree} {what}, hexa)"


def l2_policy(args):
    dof = algorithmic_bytes(args)[2]
    mb = 3 * args.cells ** 3 * dof * 8 / 1e6
    if mb > 126.0:
        return f"state (3 x {mb / 3:.0f} MB per GPU) exceeds the 126 MB L2; no flush needed"
    return (f"state (3 x {mb / 3:.0f} MB per GPU) FITS in the 126 MB L2 and nothing is flushed between steps: "
            "an L2-assisted number")


def config_dict(args, n_gpus, dims, part):
    label = ("BASELINE config 2" if (args.degree, args.cells) == (2, 42) else
             "BASELINE config 5" if (args.degree, args.cells) == (3, 32) else "other size")
    return {"workload": f"{label}: synthetic structured box, {dims[0]}x{dims[1]}x{dims[2]} Hexahedron8 "
                        f"cells total ({args.cells}^3 per GPU), "
                        + ("Euler isentropic vortex, " if args.viscous is None else
                           f"Navier-Stokes (nu={args.viscous}, direct-DG face flux) on the isentropic vortex, ") +
                        ("weno::Lazy limiter after every stage, " if args.limiter else "") +
                        ("EnergyBasedViscosity updated after every stage, " if args.energy_viscosity else "") +
                        
                        f"{SCHEME_NAME[args.scheme]} p={args.degree}, {args.riemann.upper()}, RK3, periodic" +
                        (f", interior nodes displaced by {args.warp} h" if getattr(args, "warp", 0.0) else ""),
            "cells_per_gpu": args.cells ** 3, "dof_per_gpu": args.cells ** 3 * 5 * (args.degree + 1) ** 3,
            "step": "GetTimeStep + RungeKutta<3>::Update (3 stages)", "parallelism": f"domain decomposition {part}",
            "stage_kernel": stage_kernel_name(args),
            "l2_policy": l2_policy(args)}


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    # bound the whole run to a few minutes: the sample is sized for ~1.5 s per step
    steps_run, warm_run = min(steps, 10), min(warmup, 2)
    res = cpu_run(args, steps_run, warm_run)
    px, py, pz = __import__("minicfd_b200.mesh", fromlist=["factor3"]).factor3(args.gpus)
    dims = (args.cells * px, args.cells * py, args.cells * pz)
    ours_cfg = config_dict(args, args.gpus, dims, f"{px}x{py}x{pz}")
    ran = {"workload": "reference CPU code on a bounded sample of the workload: " + res["sample"],
           "sample_of": ours_cfg["workload"], "step": ours_cfg["step"],
           "parallelism": f"{res['cores']} host thread(s), one part each, in-process halo"}
    line = {"impl": "reference", "metric": metric_name(args),
            "value": res["value"], "unit": "DOF-updates/s", "n_gpus": args.gpus, "steps": steps_run,
            "warmup": warm_run, "ms_per_step": res["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": ran,
            "cpu_baseline": {k: res[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": res["value"], "unit": "DOF-updates/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


# ---------------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------------
def algorithmic_bytes(args):
    """Algorithmic HBM bytes per DOF-update of one RK stage (+ what follows every stage) of a configuration, and
    how the figure is made up.  DOFs per cell: 5 N for the nodal schemes, 5 M for dg::General."""
    n = (args.degree + 1) ** 3
    dof = 5 * n if args.scheme != "dgm" else 5 * (10 if args.degree == 2 else 20)
    m = 10 if args.degree == 2 else 20
    state = 3 * 8 * dof                       # read the stage's state, read u0, write the next state
    geom = 192                                # 8 vertices x 3 doubles
    bas = (3 + m * (m + 1) // 2 + 3) * 8      # centre, lower-triangular orthonormalisation, lengths
    parts = {"state (read, read u0, write)": state, "vertices": geom}
    if args.scheme == "dgm":
        parts["Taylor form written and read once"] = 2 * 8 * dof
        parts["per-cell basis"] = bas
    if args.viscous is not None or args.energy_viscosity:
        # TraceKernel: gradients at every node (15 N) and Hessian.n at every face point (15 x 6 Q^2), written once and
        # read once.  The neighbours' reads at the face points are second reads of the same bytes (L2) and not counted;
        # the ncu captures of round 1 show 22 KB of DRAM traffic per cell and stage for 22.7 KB counted here.
        q2 = (args.degree + 1) ** 2
        parts["traces (gradients + Hessian.n), written and read once"] = 2 * 8 * (15 * n + 15 * 6 * q2)
    if args.energy_viscosity:
        parts["damping matrix + state read by ViscosityKernel"] = 8 * n * n + 8 * dof
    if args.limiter:
        # ModalKernel: state -> orthonormal + Taylor coefficients; WenoLazyKernel: own + 6 neighbours' Taylor
        # coefficients (each cell's read once algorithmically), basis, limited state written
        parts["limiter: state read, limited state written"] = 2 * 8 * dof
        parts["limiter: modal + Taylor coefficients written and read once"] = 4 * 8 * 5 * m
        parts["limiter: per-cell basis, twice"] = 2 * bas
    total = float(sum(parts.values()))
    return total / dof, {k: v / dof for k, v in parts.items()}, dof


FEATURES = [   # name, overrides of the headline arguments
    ("line_kernel", {"stage_kernel": "lines"}),
    ("viscous", {"viscous": 0.01}),
    ("limiter", {"limiter": True}),
    ("energy_viscosity", {"energy_viscosity": True}),
    ("dgm", {"scheme": "dgm"}),
    ("config5_p3", {"degree": 3, "cells": 32}),
]


def ours(args):
    import copy
    import gc
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the CUDA path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        # NCCL's kernels must not queue behind the thousands of pending CTAs of the inner-cell kernel
        os.environ.setdefault("TORCH_NCCL_HIGH_PRIORITY", "1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    line = measure(args, world, rank, local_rank)
    plain = (args.viscous is None and not args.limiter and not args.energy_viscosity and args.scheme == "fr"
             and args.degree == 2 and args.warp == 0.0 and args.stage_kernel is None)
    if plain and world == 1 and not args.no_features:
        # the other configurations of the path, briefly, each with its own algorithmic-bytes roofline.  One GPU only:
        # under torchrun every one of them is its own run (e.g. `--gpus 8 --degree 3 --cells 32` for BASELINE
        # config 5), so that the headline line never waits on a second rendezvous
        feats = {}
        for name, over in FEATURES:
            fa = copy.copy(args)
            for k, v in over.items():
                setattr(fa, k, v)
            fa.steps, fa.warmup, fa.no_e2e, fa.no_cpu_baseline, fa.frame_output = 12, 3, True, True, False
            gc.collect()
            torch.cuda.empty_cache()
            try:
                res = measure(fa, world, rank, local_rank, nested=True)
            except Exception as e:      # a feature must not take the headline down
                res = {"error": f"{type(e).__name__}: {e}"}
            if rank == 0:
                feats[name] = res if "error" in res else {
                    "metric": res["metric"], "workload": res["config"]["workload"], "value": res["value"],
                    "ms_per_step": res["ms_per_step"], "steps": res["steps"], "gpu_launches": res["gpu_launches"],
                    "cuda_graph": res["config"]["cuda_graph"], "roofline": res["roofline"], "check": res["check"]}
        if rank == 0:
            line["features"] = feats
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        torch.cuda.synchronize()
        dist.barrier()
        dist.destroy_process_group()


def measure(args, world, rank, local_rank, nested=False):
    """One configuration: warm up, time args.steps steps, return the JSON line's dictionary on rank 0."""
    import torch
    import torch.distributed as dist
    from minicfd_b200 import capi, fields, halo, mesh as meshmod
    from minicfd_b200.part import Part
    from minicfd_b200.spatial import RungeKutta, Spatial

    env_before = os.environ.get("MCFD_STAGE_KERNEL")
    if args.stage_kernel is not None:          # read by mcfd_create, i.e. when the Spatial below is built
        os.environ["MCFD_STAGE_KERNEL"] = args.stage_kernel
    px, py, pz = meshmod.factor3(world)
    n = args.cells
    dims = (n * px, n * py, n * pz)
    m = meshmod.box(*dims, hi=(10.0 * px, 10.0 * py, 10.0 * pz), periodic=True, warp=args.warp, seed=7)
    parts = meshmod.block_partition(m, px, py, pz) if world > 1 else None
    part = Part(m, parts, rank, world)
    scheme = {"fr": capi.FR_LOBATTO, "dg": capi.DG_LOBATTO, "dgm": capi.DG_GENERAL}[args.scheme]
    kf_halo = 5 * (args.degree + 1) ** 3 if args.scheme != "dgm" else 5 * (10 if args.degree == 2 else 20)
    comm = halo.for_part(part, kf_halo) if world > 1 else None
    sp = Spatial(part, scheme=scheme, degree=args.degree, riemann=RIEMANN[args.riemann], device=local_rank,
                 comm=comm, diffusion=capi.CELL_VISCOSITY_DDG if args.energy_viscosity else
                 (capi.INVISCID if args.viscous is None else capi.NAVIER_STOKES_DDG))
    if args.stage_kernel is not None:
        if env_before is None:
            del os.environ["MCFD_STAGE_KERNEL"]
        else:
            os.environ["MCFD_STAGE_KERNEL"] = env_before
    if args.energy_viscosity:
        sp.SetDiffusion(2.0, 1.0 / 12, 0.0, 0.708)
        sp.EnableEnergyBasedViscosity()
    if args.viscous is not None:
        sp.SetDiffusion(2.0, 1.0 / 12, args.viscous, 0.708)
    if args.limiter:
        sp.SetLimiter(0.001, 1e-6)
    sp.ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    sp.Approximate(lambda x: fields.vortex()(x, 0.0))
    use_peer = (world > 1 and args.halo == "peer" and args.scheme != "dgm" and args.viscous is None
                and not args.limiter and not args.energy_viscosity)
    if comm is not None:
        comm.exchange()
    if use_peer:
        from minicfd_b200 import peer
        peer.connect(sp, part)
    sums0 = conserved_sums(sp, part, args, world)
    rk = RungeKutta(3)
    n_dof_local = part.n_local * sp.kfields
    n_dof_total = n_dof_local * world

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    from minicfd_b200.spatial import time_step_device
    dt_dev = torch.zeros(1, dtype=torch.float64, device="cuda")

    def step_device(t):
        # GetTimeStep (+ MIN over ranks) and Update without a host round trip: dt stays on the device
        if use_peer:     # the whole step, halo and MIN included, enqueued by the library
            sp.ctx.peer_step(3, t, 1.0)
            return 0.0
        time_step_device(sp, 1.0, 3, dt_dev)
        rk.UpdateDevice(sp, t, dt_dev.data_ptr())
        return 0.0

    t = 0.0
    for _ in range(max(args.warmup, 3)):
        t += step_device(t)
    # THREE steps (each GetTimeStep [+ MIN over ranks] + 3 stages [+ ghost exchange]) captured once as a CUDA graph and
    # replayed: the three state buffers rotate with period three, so the replays march on in time exactly like host
    # launches do.  At N > 1 the peer path is captured the same way (its push counts live on the device); the NCCL
    # path is launched from the host (replay with NCCL point-to-point inside was slower and did not shut down cleanly).
    kGraphSteps = 3
    graph = None
    launches_per_graph = None
    if not args.no_graph and (world == 1 or use_peer):
        try:
            barrier()
            l0 = sp.ctx.launch_count()
            graph = torch.cuda.CUDAGraph()
            cap_stream = torch.cuda.Stream()
            cap_stream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(cap_stream):
                sp.ctx.set_stream(cap_stream.cuda_stream)
                with torch.cuda.graph(graph, stream=cap_stream, capture_error_mode="relaxed"):
                    for _ in range(kGraphSteps):
                        step_device(t)
                    if use_peer:
                        sp.ctx.peer_join()      # the library's side stream rejoins the captured one
                sp.ctx.set_stream(torch.cuda.current_stream().cuda_stream)
            launches_per_graph = sp.ctx.launch_count() - l0
            torch.cuda.current_stream().wait_stream(cap_stream)
            sp.ctx.set_stream(torch.cuda.current_stream().cuda_stream)
            for _ in range(2):
                graph.replay()
            barrier()
        except Exception as e:  # capture not possible: fall back to host launches, and say so
            if rank == 0:
                print(f"bench.py: CUDA graph capture failed ({type(e).__name__}: {e}); launching from the host",
                      file=sys.stderr)
            graph = None
            sp.ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    if world > 1:    # all ranks replay, or none does
        ok = torch.tensor([1.0 if (graph is not None or not use_peer or args.no_graph) else 0.0], device="cuda")
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if ok.item() == 0.0:
            raise SystemExit("bench.py: graph capture succeeded on some ranks only")

    def run_steps(tt, n_steps):
        done = 0
        if graph is not None:
            for _ in range(n_steps // kGraphSteps):
                graph.replay()
            done = n_steps // kGraphSteps * kGraphSteps
        for _ in range(n_steps - done):
            tt += step_device(tt)
        return tt

    launches0 = sp.ctx.launch_count()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank) as clocks:
        barrier()
        clocks.mark_begin()
        e0.record()
        host_t0 = time.perf_counter()
        t = run_steps(t, args.steps)
        host_issue_ms = 1e3 * (time.perf_counter() - host_t0)   # time the host needed to issue the steps
        e1.record()
        barrier()
        clocks.mark_end()
    ms = e0.elapsed_time(e1)
    launches = sp.ctx.launch_count() - launches0
    if graph is not None:
        launches += launches_per_graph * (args.steps // kGraphSteps)
    # duration of the stage kernel: CUDA events around every stage launch, on the launching stream, over
    # the same steps launched from the host right after the timed region (events cannot sit in a graph)
    k_steps = min(args.steps, 20)
    sp.ctx.enable_kernel_timing(True)
    for _ in range(k_steps):
        t += step_device(t)
    barrier()
    kernel_ms, kernel_launches = sp.ctx.kernel_time(reset=True)
    sp.ctx.enable_kernel_timing(False)
    if use_peer:
        sp.ctx.peer_check()
    sums1 = conserved_sums(sp, part, args, world)
    tmax = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms = float(tmax.item())
    value = 3.0 * n_dof_total * args.steps / (ms * 1e-3)

    # ---- e2e: host columns in and out every step ---------------------------------------------------
    e2e = None
    if not args.no_e2e:
        bufs = [torch.empty(n_dof_local, dtype=torch.float64, pin_memory=True) for _ in range(2)]
        bufs[0].numpy()[:] = sp.GetSolutionColumn()
        e_steps = max(3, min(args.steps, 20))

        def step_e2e(tt):
            # the user's loop on host columns: SetSolutionColumn(host) -> GetTimeStep -> Update ->
            # GetSolutionColumn(host); the output buffer of one step is the input of the next
            sp.ctx.set_solution(bufs[0].numpy())          # H2D from pinned memory (ghosts: refreshed in Update)
            if use_peer:
                sp.ctx.peer_push()                        # the new solution's ghost images
                sp.ctx.peer_step(3, tt, 1.0)
                sp.ctx.get_solution(bufs[1].numpy())
                bufs.reverse()
                return 0.0
            dt = sp.GetTimeStep(1.0, 3)                   # includes the D2H read of dt
            rk.Update(sp, tt, dt)
            sp.ctx.get_solution(bufs[1].numpy())          # D2H of the step's result
            bufs.reverse()
            return dt
        for _ in range(2):
            t += step_e2e(t)
        barrier()
        t0 = time.perf_counter()
        e0.record()
        for _ in range(e_steps):
            t += step_e2e(t)
        e1.record()
        barrier()
        ms_e = max(e0.elapsed_time(e1), 1e3 * (time.perf_counter() - t0))
        tm = torch.tensor([ms_e], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        e2e = {"value": 3.0 * n_dof_total * e_steps / (float(tm.item()) * 1e-3), "unit": "DOF-updates/s",
               "h2d_bytes_per_step": n_dof_local * 8, "d2h_bytes_per_step": n_dof_local * 8 + 8,
               "steps": e_steps}

    if rank == 0:
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        else:
            peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
        b_alg, b_parts, _ = algorithmic_bytes(args)
        per_launch_ms = kernel_ms / max(kernel_launches, 1)
        # one launch per stage when there is no ghost layer; with ghosts a stage is two launches
        launches_per_stage = kernel_launches / (3.0 * k_steps)
        achieved = b_alg * n_dof_local / (per_launch_ms * launches_per_stage * 1e-3) / 1e9
        over = "the stage kernel's launches"
        if args.viscous is not None or args.energy_viscosity or args.limiter or args.scheme == "dgm":
            # several kernels per stage (traces, limiter, viscosity, Taylor form): the bytes are those of the whole
            # stage, so the time is the whole step's
            achieved = b_alg * 3.0 * n_dof_local / ((ms / args.steps) * 1e-3) / 1e9
            over = "the whole step (every kernel of the three stages and GetTimeStep)"
        roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                    "traffic": None, "peak_source": peak_src, "kernel": stage_kernel_name(args),
                    "kernel_reported_by_library": sp.ctx.stage_kernel(), "over": over,
                    "algorithmic_bytes_per_dof_update": b_alg, "algorithmic_bytes_made_of": b_parts,
                    "kernel_ms_per_stage": per_launch_ms * launches_per_stage,
                    "kernel_share_of_step": (kernel_ms / k_steps) / (ms / args.steps),
                    "kernel_timing": f"CUDA events around each stage launch over {k_steps} host-launched steps"}
        # DRAM bytes per launch from the committed ncu capture: only for the shape that capture profiled
        traffic_path = os.path.join(ROOT, "profiles", "stage_kernel_traffic.json")
        if os.path.exists(traffic_path):
            measured = json.load(open(traffic_path))
            if measured.get("shape") == traffic_shape(args, world):
                roofline["traffic"] = measured.get("dram_bytes_per_launch")
                roofline["traffic_source"] = measured.get("source")
        line = {"metric": metric_name(args), "value": value, "unit": "DOF-updates/s",
                "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": dict(config_dict(args, world, dims, f"{px}x{py}x{pz}"),
                               cuda_graph=(f"{kGraphSteps} steps per replay" if graph is not None else False)),
                "clocks": clocks.summary(),
                "roofline": roofline, "gpu_launches": int(launches),
                "halo": ("none (one part)" if world == 1 else
                         "stage kernel packs, copy engines deliver into the neighbours' receive stages, epoch words; step "
                         "enqueued by the library (mcfd_peer_step)" if use_peer else
                         "pack -> NCCL send/recv -> unpack, issued per stage from Python"),
                "check": conservation_check(sums0, sums1),
                "host_issue_ms_per_step": host_issue_ms / args.steps}
        if e2e is not None:
            line["e2e"] = e2e
        if args.frame_output:
            line["frame_output"] = frame_output(sp, args)
        if world == 1 and args.warp == 0.0 and not args.no_e2e and args.scheme != "dgm" and args.viscous is None:
            # the box's cells are parallelepipeds and take the kernel instantiation with one metric per cell; the same
            # workload on a warped copy of the mesh (general trilinear metric at every node and face point) beside it
            line["warped_mesh"] = warped_mesh_run(args)
        if world == 1 and not args.no_cpu_baseline:
            res = cpu_run(args, steps=3, warmup=1)
            line["cpu_baseline"] = {k: res[k] for k in ("value", "unit", "cores", "kind", "sample")}
    if world > 1:
        comm.drain()
        torch.cuda.synchronize()
        dist.barrier()
    return line if rank == 0 else None


def warped_mesh_run(args, steps=30, warp=0.15):
    import torch
    from minicfd_b200 import capi, fields, mesh as meshmod
    from minicfd_b200.part import Part
    from minicfd_b200.spatial import RungeKutta, Spatial, time_step_device
    n = args.cells
    m = meshmod.box(n, n, n, hi=(10.0, 10.0, 10.0), periodic=True, warp=warp, seed=7)
    scheme = {"fr": capi.FR_LOBATTO, "dg": capi.DG_LOBATTO}[args.scheme]
    sp = Spatial(Part(m), scheme=scheme, degree=args.degree, riemann=RIEMANN[args.riemann], device=torch.cuda.current_device())
    sp.ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    sp.Approximate(lambda x: fields.vortex()(x, 0.0))
    rk = RungeKutta(3)
    dt_dev = torch.zeros(1, dtype=torch.float64, device="cuda")

    def step():
        time_step_device(sp, 1.0, 3, dt_dev)
        rk.UpdateDevice(sp, 0.0, dt_dev.data_ptr())
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    sp.ctx.enable_kernel_timing(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    kernel_ms, launches = sp.ctx.kernel_time(reset=True)
    ms = e0.elapsed_time(e1) / steps
    n_dof = sp.part.n_local * sp.kfields
    return {"what": f"the same workload on a mesh whose interior nodes are displaced by {warp} h (no cell is a parallelepiped), "
                    f"{steps} host-launched steps", "ms_per_step": ms, "value": 3.0 * n_dof / (ms * 1e-3),
            "kernel_ms_per_stage": kernel_ms / max(launches, 1), "finite": bool(np.isfinite(sp.GetSolutionColumn()).all())}


def conserved_sums(sp, part, args, world):
    """Integrals of the five conserved variables over the whole (periodic) domain: sum over ranks of
    sum_cells sum_nodes w_i w_j w_k det(J) u — a correctness signal that costs one column read.  With a periodic
    wrap the scheme conserves them to rounding, and only if every face sees the same two states on both sides, i.e.
    only if the ghost rows are the neighbours' current values."""
    import torch
    import torch.distributed as dist
    if args.scheme == "dgm":
        return None
    q = args.degree + 1
    w1 = {3: np.array([1.0, 4.0, 1.0]) / 3, 4: np.array([1.0, 5.0, 5.0, 1.0]) / 6}[q]
    w = (w1[:, None, None] * w1[None, :, None] * w1[None, None, :]).reshape(-1)
    det = sp.ctx.jacobian_determinants()                      # [n_local][N]
    u = sp.GetSolutionColumn().reshape(part.n_local, q ** 3, 5)
    sums = torch.from_numpy(np.einsum("cn,n,cnk->k", det, w, u)).cuda()
    if world > 1:
        dist.all_reduce(sums)
    return sums.cpu().numpy()


def conservation_check(s0, s1):
    if s0 is None or s1 is None:
        return None
    scale = np.maximum(np.abs(s0), np.abs(s0).max() * 1e-3)
    drift = float(np.max(np.abs(s1 - s0) / scale))
    return {"what": "relative drift of the global integrals of (rho, rho u, rho v, rho w, E) over the timed run "
                    "(periodic domain: conserved to rounding only if every ghost row is current)",
            "conservation_rel_drift": drift, "finite": bool(np.isfinite(s1).all()), "ok": bool(drift < 1e-10)}


def frame_output(sp, args):
    """The step after the path: one frame written through vtk.Writer -- values sampled and turned into
    text on the device, bytes written by the host -- next to the reference's vtk::Writer<Part>::WriteSolutions."""
    import shutil
    import tempfile
    from minicfd_b200 import mesh as meshmod, vtk
    d = tempfile.mkdtemp()
    w = vtk.Writer(sp)
    w.WriteSolutions(d, "Frame0")               # first frame: encodes the static points and cell arrays
    t0 = time.perf_counter()
    n_frames = 3
    for i in range(n_frames):
        w.WriteSolutions(d, "Frame%d" % (i + 1))
    ms = 1e3 * (time.perf_counter() - t0) / n_frames
    t0 = time.perf_counter()
    sp.ctx.sample_vtk_on_device(False)
    texts = [sp.ctx.vtk_component_text(k) for k in range(5)]
    ms_text = 1e3 * (time.perf_counter() - t0)
    n_cells = sp.ctx.n_local
    size = os.path.getsize(os.path.join(d, "Frame1", "0.vtu"))
    shutil.rmtree(d)
    out = {"what": "vtk.Writer.WriteSolutions: sampling + number formatting on the GPU, file written by the host",
           "ms_per_frame": ms, "cells_per_s": n_cells / (ms * 1e-3), "file_bytes": size,
           "ms_sample_and_format_5_fields": ms_text, "text_bytes_5_fields": sum(len(t) for t in texts)}
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import refapi
    if refapi.available(args.degree):
        n = 16
        m = meshmod.box(n, n, n, hi=(10.0, 10.0, 10.0))
        r = refapi.Ref(m, degree=args.degree, scheme=refapi.FR, riemann=1)
        r.approximate(refapi.F_SMOOTH, [1.0])
        d = tempfile.mkdtemp()
        t0 = time.perf_counter()
        r.write_vtu(d, "Frame0")
        dt = time.perf_counter() - t0
        shutil.rmtree(d)
        out["cpu_reference"] = {"cells_per_s": n ** 3 / dt, "cores": 1, "kind": "reference",
                                "sample": f"vtk::Writer<Part>::WriteSolutions of {n}^3 cells"}
    return out


def main():
    args = parse()
    if args.impl == "reference":
        reference_arm(args)
    else:
        ours(args)


if __name__ == "__main__":
    main()
