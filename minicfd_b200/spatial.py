"""Host-side mirror of the reference's operator interface for the hot path, over the C ABI.

`Spatial` plays the role of `mini::spatial::fr::Lobatto<Part, Riemann>` / `dg::Lobatto<Part, Riemann>`
as a `mini::temporal::System` (include/mini/spatial/fem.hpp:47-436, include/mini/temporal/ode.hpp:17-48):
same method names, argument meaning and call order, so that the parity tests read like the reference's
own tests (test/spatial/fr.cpp).  All arithmetic happens in libminicfd_b200.so on the GPU.
"""
from __future__ import annotations

import numpy as np

from . import capi
from .part import Part, cell_points


class Spatial:
    def __init__(self, part: Part, *, scheme=capi.FR_LOBATTO, degree=2, riemann=capi.HLLC, device=-1,
                 comm=None, fr_c_next=0.0, diffusion=capi.INVISCID, k_local=False):
        self.part = part
        self.degree = degree
        self.scheme = scheme
        self.k_local = bool(k_local)
        self.ctx = capi.Context(scheme, degree, riemann, device, fr_c_next, diffusion, k_local)
        self.kfields = self.ctx.kfields
        m = part.mesh
        self.ctx.set_mesh(m.nodes, part.cell_nodes, part.n_local, part.n_ghost, part.bfaces, part.btags,
                          m.periodic, global_ids=part.global_ids)
        if part.n_ghost:
            self.ctx.set_halo([len(c) for c in part.send_cells],
                              np.concatenate(part.send_cells) if part.send_cells else [],
                              [len(c) for c in part.recv_cells],
                              np.concatenate(part.recv_cells) if part.recv_cells else [])
        self.comm = comm            # halo exchanger (minicfd_b200.halo) or None for a single part
        if comm is not None:
            comm.attach(self)
        self._limiter = None        # (w0, eps) of the weno::Lazy limiter, or None
        self._bc_funcs = {}         # tag -> (kind, func or None, time_dependent)
        self._bc_points = {}
        self._t = 0.0
        self.ctx.set_time_callback(self._on_time)

    # ---- mini::spatial::FiniteElement (fem.hpp:128-154, 196-222) ---------------------------------
    def name(self):
        return "FR::Lobatto" if self.ctx is not None else ""

    def part_ptr(self):
        return self.part

    def Approximate(self, func):
        """fem.hpp:147-154: nodal sampling of func(xyz) on every local cell, then a halo exchange."""
        if self.scheme == capi.DG_GENERAL:   # Projection::Approximate: L2 projection from the Gauss points
            if self.comm is not None:
                self.comm.drain()
            self.ctx.project(np.asarray(func(self.ctx.quadrature_points())))
            return
        pts = cell_points(self.part.mesh.nodes, self.part.cell_nodes[:self.part.n_local], self.degree)
        vals = np.asarray(func(pts))
        if self.k_local:   # Hexahedron::SetValue with kLocal == true: coeff = value * det J (hexahedron.hpp:344-349)
            vals = vals * self.ctx.jacobian_determinants()[:, :, None]
        self.SetSolutionColumn(vals.reshape(-1))

    def _set_bc(self, tag, kind, func=None, time_dependent=True):
        self.ctx.set_boundary(tag, kind)
        self._bc_funcs[tag] = (kind, func, time_dependent)
        # a device-resident time step (UpdateDevice, peer_step) cannot serve stage times to such a function
        self.ctx.set_time_dependent(any(f is not None and td for _, f, td in self._bc_funcs.values()))
        if func is not None:
            self._bc_points[tag] = self.ctx.boundary_points(tag)
            self.ctx.set_boundary_values(tag, func(self._bc_points[tag], self._t))

    def SetSmartBoundary(self, tag, func, time_dependent=True):
        self._set_bc(tag, capi.BC_SMART, func, time_dependent)

    def SetSupersonicInlet(self, tag, func, time_dependent=True):
        self._set_bc(tag, capi.BC_SUPERSONIC_INLET, func, time_dependent)

    def SetSubsonicInlet(self, tag, func, time_dependent=True):
        self._set_bc(tag, capi.BC_SUBSONIC_INLET, func, time_dependent)

    def SetSubsonicOutlet(self, tag, func, time_dependent=True):
        self._set_bc(tag, capi.BC_SUBSONIC_OUTLET, func, time_dependent)

    def SetNoSlipWall(self, tag, func=None, time_dependent=False):
        """fem.hpp:219-222.  func(xyz, t) -> (unused, wall velocity u v w, normal temperature gradient),
        demo/laminar/cavity.cpp:30-38; ignored by inviscid systems (fr/general.hpp:455-459)."""
        self._set_bc(tag, capi.BC_NO_SLIP_WALL, func, time_dependent)

    def EnableEnergyBasedViscosity(self, time_scale=1.0):
        """Makes this object behave like `WithViscosity<General>` over `EnergyBasedViscosity<Part, Riemann>`
        (spatial/with_viscosity.hpp:20-73, demo/euler/sourceless.hpp:82-103): damping matrices now
        (Viscosity::Initialize), UpdateProperties after every SetSolutionColumn and Runge-Kutta stage (with whatever
        the ghost cells hold then, like the reference), the ghost viscosities exchanged before every residual, and the
        time step from its min_dt_."""
        self.ctx.build_damping_matrices()
        self._energy_time_scale = float(time_scale)
        self._nu_comm = None
        if self.comm is not None:          # ShareGhostCellProperties / UpdateGhostCellProperties: 5 doubles per cell
            from . import halo
            self._nu_comm = halo.for_part(self.part, 5, self.comm.group)
            self._nu_comm.attach_viscosity(self)

    def ShareViscosity(self):
        """with_viscosity.hpp:45-50: the ghost cells' viscosities before a residual."""
        if getattr(self, "_nu_comm", None) is not None:
            self._nu_comm.exchange()

    def UpdateProperties(self):
        """EnergyBasedViscosity::UpdateProperties (viscosity.hpp:556-572).  The reference does not exchange anything
        here: the jumps on ghost faces use the ghost cells' coefficients as last received.  The stages write into
        alternating state buffers, so those values are unpacked (again) into the buffer that is current now."""
        if getattr(self, "_energy_time_scale", None) is not None:
            if self.comm is not None and self.comm.received:
                self.comm.drain()
                self.ctx.halo_unpack()
            self.ctx.update_cell_viscosity(self._energy_time_scale)

    def SetCellViscosity(self, nu):
        """EnergyBasedViscosity::properties_ (viscosity.hpp:57, 564-566): nu[n_local + n_ghost][5]."""
        self.ctx.set_cell_viscosity(nu)

    def SetLimiter(self, w0=0.001, eps=1e-6, eigen=False):
        """Makes this object behave like `WithLimiter<General, limiter::weno::Lazy<Cell>>(&limiter, &part)` with
        `limiter(w0, eps)` (spatial/with_limiter.hpp:21-68, demo/euler/sourceless.cpp:114-120): the limiter runs
        after every SetSolutionColumn / Approximate and after every Runge-Kutta stage.  None switches it off;
        eigen=True selects `weno::Eigen<Cell, Riemann>` (the characteristic variant with its jump detector)."""
        self._limiter = None if w0 is None else (float(w0), float(eps))
        self._limit_call = self.ctx.limiter_weno_eigen if eigen else self.ctx.limiter_weno_lazy

    def Reconstruct(self):
        """mini::limiter::Reconstruct(part, limiter) (limiter/reconstruct.hpp:21-54): ShareGhostCellCoeffs, limit
        the inner cells, UpdateGhostCellCoeffs, limit the inter cells."""
        if self._limiter is None:
            return
        w0, eps = self._limiter
        if self.comm is None:
            self._limit_call(w0, eps, -1)
            return
        self.comm.drain()
        self.comm.start()                      # the unlimited values travel
        self._limit_call(w0, eps, 0)
        self.comm.finish()
        self._limit_call(w0, eps, 1)

    def SetDiffusion(self, beta_0, beta_1, nu, prandtl):
        """Riemann::Diffusion::SetBetaValues + SetProperty (direct.hpp:80-83, navier_stokes.hpp:57-60)."""
        self.ctx.set_diffusion(beta_0, beta_1, nu, prandtl)

    def SetInviscidWall(self, tag):
        self._set_bc(tag, capi.BC_INVISCID_WALL)

    def SetSupersonicOutlet(self, tag):
        self._set_bc(tag, capi.BC_SUPERSONIC_OUTLET)

    def _on_time(self, t):
        """Re-sample time-dependent boundary functions (the reference evaluates its std::function at
        (xyz, t_curr_) inside the face loops, fr/lobatto.hpp:350)."""
        self._t = t
        for tag, (kind, func, time_dependent) in self._bc_funcs.items():
            if func is not None and time_dependent:
                self.ctx.set_boundary_values(tag, func(self._bc_points[tag], t))

    # ---- mini::temporal::System (ode.hpp:17-48; fem.hpp:225-258) -----------------------------------
    def SetTime(self, t):
        self._t = t

    def SetSolutionColumn(self, column):
        if self.comm is not None:
            self.comm.drain()              # an exchange of the old state may still be in flight
        self.ctx.set_solution(column)
        self.Reconstruct()                 # with_limiter.hpp:57-61
        self.UpdateProperties()            # with_viscosity.hpp:52-55

    def GetSolutionColumn(self):
        return self.ctx.get_solution()

    def GetResidualColumn(self):
        """fem.hpp:248-258: halo exchange, then volume + face + boundary terms (one fused kernel)."""
        if self.comm is not None:
            self.comm.exchange()
        self.ShareViscosity()
        return self.ctx.residual(self._t)

    def GetTimeStep(self, dt_guess, rk_order):
        """fem.hpp:417-435 / fr/lobatto.hpp:220-222, then MPI_Allreduce(MIN) as in sourceless.cpp:194."""
        dt = self.ctx.time_step(dt_guess, rk_order)
        if self.comm is not None:
            dt = self.comm.allreduce_min(dt)
        return dt


def time_step_device(system: Spatial, dt_guess, rk_order, dt_tensor):
    """GetTimeStep into a 1-element CUDA tensor, min-reduced over the ranks on the device."""
    system.ctx.time_step_device(dt_guess, rk_order, dt_tensor.data_ptr())
    if system.comm is not None:
        system.comm.allreduce_min_tensor(dt_tensor)


def average_reconstruct(system: Spatial):
    """mini::limiter::average::Reconstruct(part) (limiter/average.hpp:11-28)."""
    system.ctx.limiter_reconstruct(0)


def majority_reconstruct(system: Spatial):
    """mini::limiter::majority::Reconstruct(part) (limiter/majority.hpp:14-42)."""
    system.ctx.limiter_reconstruct(1)


class RungeKutta:
    """mini::temporal::RungeKutta<kOrders> (include/mini/temporal/rk.hpp:10-57).  `Update` uses the
    fused stage kernels (mcfd_rk_*); `UpdateGeneric` spells the reference's column algebra out through
    the System interface and is used by tests to check the fused path against it."""

    def __init__(self, order):
        assert order in (1, 2, 3)
        self.order = order

    def Update(self, system: Spatial, t, dt):
        ctx = system.ctx
        if system._limiter is not None or getattr(system, "_energy_time_scale", None) is not None:
            ctx.rk_begin(self.order, t, dt)
            self._stages_with_limiter(system)
            return
        if system.comm is None:
            ctx.rk_step(self.order, t, dt)
            return
        ctx.rk_begin(self.order, t, dt)
        self._stages_with_halo(system)

    def _stages_with_limiter(self, system: Spatial):
        """rk.hpp:16-49 on a WithLimiter system: every SetSolutionColumn of a stage result is followed by the
        limiter (with_limiter.hpp:57-61), whose own ghost exchange carries the unlimited values; the next
        residual exchanges the limited ones (fem.hpp:249-256)."""
        ctx, comm = system.ctx, system.comm
        for stage in range(self.order):
            if comm is not None:
                comm.exchange()
            system.ShareViscosity()
            ctx.rk_stage_inner(stage)
            ctx.rk_stage_inter(stage)
            system.Reconstruct()
            system.UpdateProperties()

    def _stages_with_halo(self, system: Spatial):
        """Every stage: the inner cells advance on the main stream (fem.hpp:249-254) while a high-priority
        side stream waits for the ghost data, unpacks it, advances the inter cells (fem.hpp:255-256) and at
        once packs and posts the exchange of the *next* stage input — the send cells are inter cells, so
        their new values exist as soon as the small inter kernel is done, long before the inner kernel is."""
        import torch
        ctx, comm = system.ctx, system.comm
        main = torch.cuda.current_stream()
        if getattr(system, "_side_stream", None) is None:
            system._side_stream = torch.cuda.Stream(priority=-1)
        side = system._side_stream
        if not comm.inflight:                  # first stage after a SetSolutionColumn: post the exchange now
            comm.start()
        for stage in range(self.order):
            inner_input_ready = torch.cuda.Event()
            inner_input_ready.record(main)     # everything queued on main so far (incl. the previous inner kernel)
            ctx.rk_stage_inner(stage)          # main
            with torch.cuda.stream(side):
                ctx.set_stream(side.cuda_stream)
                side.wait_event(inner_input_ready)
                comm.finish()                  # wait + unpack (UpdateGhostCellCoeffs)
                ctx.rk_stage_inter(stage)
                inter_done = torch.cuda.Event()
                inter_done.record(side)
                comm.start()                   # pack the new inter-cell values, post the next exchange
                ctx.set_stream(main.cuda_stream)
            main.wait_event(inter_done)

    def UpdateDevice(self, system: Spatial, t, dt_ptr: int):
        """Update with the time step resident in device memory (no host round trip)."""
        ctx = system.ctx
        if system._limiter is not None or getattr(system, "_energy_time_scale", None) is not None:
            ctx.rk_begin_device(self.order, t, dt_ptr)
            self._stages_with_limiter(system)
            return
        if system.comm is None:
            ctx.rk_step_device(self.order, t, dt_ptr)
            return
        ctx.rk_begin_device(self.order, t, dt_ptr)
        self._stages_with_halo(system)

    def UpdateGeneric(self, system: Spatial, t, dt):
        def euler_next(tt):
            system.SetTime(tt)
            u = system.GetSolutionColumn()
            r = system.GetResidualColumn()
            r *= dt
            u += r
            return u
        if self.order == 1:
            system.SetSolutionColumn(euler_next(t))
        elif self.order == 2:
            u0 = system.GetSolutionColumn()
            system.SetSolutionColumn(euler_next(t))
            u = euler_next(t + dt)
            u += u0
            u *= 0.5
            system.SetSolutionColumn(u)
        else:
            u0 = system.GetSolutionColumn()
            system.SetSolutionColumn(euler_next(t))
            u = u0 * 3
            u += euler_next(t + dt)
            u /= 4
            system.SetSolutionColumn(u)
            u = euler_next(t + dt / 2)
            u *= 2
            u += u0
            u /= 3
            system.SetSolutionColumn(u)
