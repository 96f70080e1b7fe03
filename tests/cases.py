"""Shared parity cases: meshes, boundary set-ups and fields used by the golden generator, the oracle
pinning tests, the host-logic tests and the GPU parity tests."""
from __future__ import annotations

import numpy as np

from minicfd_b200 import fields, mesh

# bc kinds (same numbering in refapi, oracle_check and capi)
WALL, SUP_OUT, SUP_IN, SUB_IN, SUB_OUT, SMART, NO_SLIP = range(7)

MIXED_BCS = {0: WALL, 1: SUP_OUT, 2: SUP_IN, 3: SMART, 4: SUB_OUT, 5: SUB_IN}
SUBSONIC_INLET_STATE = [2.0, 0.0, 0.0, -1.0, 0.01]   # p_total, direction cosines (into zmax), T_total
SMOOTH_A = 1.3


def small_mesh(seed=3):
    """3x3x2 warped box with every cell randomly rotated and the cell order permuted."""
    return mesh.scramble(mesh.box(3, 3, 2, warp=0.15, seed=seed), seed=seed)


def noisy_state(pts, seed=0, amp=1e-3):
    """Smooth admissible field + multiplicative noise, so that neighbouring cells never hold identical
    face states (the Exact solver's Newton start is a sign test on f(p_left), which is rounding noise
    for identical states: exact.hpp:140-152)."""
    u = fields.smooth(1.0)(pts, 0.0)
    rng = np.random.default_rng(seed)
    return (u * (1 + amp * rng.uniform(-1, 1, size=u.shape))).reshape(-1)


def apply_bcs_ref(r, kinds=MIXED_BCS):
    import refapi
    for tag, k in kinds.items():
        if k == SUB_IN:
            r.set_bc(tag, k, refapi.F_CONST, SUBSONIC_INLET_STATE)
        else:
            r.set_bc(tag, k, refapi.F_SMOOTH, [SMOOTH_A])


def apply_bcs_oracle(o, kinds=MIXED_BCS):
    import oracle_check as oc
    for tag, k in kinds.items():
        if k == SUB_IN:
            o.set_bc(tag, k, oc.F_CONST, SUBSONIC_INLET_STATE)
        else:
            o.set_bc(tag, k, oc.F_SMOOTH, [SMOOTH_A])


def bc_function(kind):
    if kind == SUB_IN:
        return fields.constant(SUBSONIC_INLET_STATE)
    return fields.smooth(SMOOTH_A)


def apply_bcs_spatial(sp, kinds=MIXED_BCS):
    for tag, k in kinds.items():
        f = bc_function(k)
        if k == WALL:
            sp.SetInviscidWall(tag)
        elif k == SUP_OUT:
            sp.SetSupersonicOutlet(tag)
        elif k == SUP_IN:
            sp.SetSupersonicInlet(tag, f)
        elif k == SUB_IN:
            sp.SetSubsonicInlet(tag, f)
        elif k == SUB_OUT:
            sp.SetSubsonicOutlet(tag, f)
        else:
            sp.SetSmartBoundary(tag, f)


GOLDEN_CASES = [  # (name, degree, scheme, riemann)
    ("fr_p2_exact", 2, 0, 0), ("fr_p2_hllc", 2, 0, 1), ("fr_p2_ausm", 2, 0, 2),
    ("dg_p2_hllc", 2, 1, 1), ("fr_p3_hllc", 3, 0, 1), ("dg_p3_ausm", 3, 1, 2),
]
GOLDEN_T = 0.25
GOLDEN_DT = 1e-3


# ---- Navier-Stokes with the direct-DG face flux ----------------------------------
VISCOUS = (2.0, 1.0 / 12, 0.02, 0.708)            # beta_0, beta_1 (demo/laminar/cavity.cpp:50), nu, Prandtl
VISCOUS_BCS = {0: NO_SLIP, 1: WALL, 2: NO_SLIP, 3: SMART, 4: SUP_OUT, 5: SUP_IN}
WALL_VALUE = [0.0, 0.1, -0.2, 0.05, 0.3]          # (unused, wall velocity, normal temperature gradient)
VISCOUS_GOLDEN_CASES = [("ns_p2_hllc", 2, 1), ("ns_p2_exact", 2, 0), ("ns_p3_ausm", 3, 2)]   # (name, degree, riemann)


def apply_viscous_ref(r, kinds=VISCOUS_BCS, vis=VISCOUS):
    import refapi
    r.set_viscous(*vis)
    for tag, k in kinds.items():
        if k == NO_SLIP:
            r.set_bc(tag, k, refapi.F_CONST, WALL_VALUE)
        else:
            r.set_bc(tag, k, refapi.F_SMOOTH, [SMOOTH_A])


def apply_viscous_oracle(o, kinds=VISCOUS_BCS, vis=VISCOUS):
    import oracle_check as oc
    o.set_viscous(*vis)
    for tag, k in kinds.items():
        if k == NO_SLIP:
            o.set_bc(tag, k, oc.F_CONST, WALL_VALUE)
        else:
            o.set_bc(tag, k, oc.F_SMOOTH, [SMOOTH_A])


def viscous_bc_function(kind):
    return fields.constant(WALL_VALUE) if kind == NO_SLIP else fields.smooth(SMOOTH_A)


def apply_viscous_spatial(sp, kinds=VISCOUS_BCS, vis=VISCOUS):
    sp.SetDiffusion(*vis)
    for tag, k in kinds.items():
        if k == NO_SLIP:
            sp.SetNoSlipWall(tag, viscous_bc_function(k))
        elif k == WALL:
            sp.SetInviscidWall(tag)
        elif k == SUP_OUT:
            sp.SetSupersonicOutlet(tag)
        elif k == SUP_IN:
            sp.SetSupersonicInlet(tag, viscous_bc_function(k))
        else:
            sp.SetSmartBoundary(tag, viscous_bc_function(k))


# ---- weno::Lazy limiter -------------------------------------------------------------
WENO = (0.001, 1e-6)                                   # demo/euler/sourceless.cpp:114
WENO_GOLDEN_CASES = [("weno_p2_hllc", 2, 1), ("weno_p3_hllc", 3, 1)]   # (name, degree, riemann)
WENO_EIGEN_GOLDEN_CASES = [("weno_eigen_p2_hllc", 2, 1), ("weno_eigen_p3_hllc", 3, 1)]


def shocked_state(pts, seed=0, amp=1e-2, jump=3.0, x0=0.5):
    """noisy_state with the whole state multiplied by `jump` beyond x = x0: some cells trip the detector of
    weno::Eigen (limiter/weno.hpp:230-258), the others stay untouched."""
    u = noisy_state(pts, seed=seed, amp=amp).reshape(pts.shape[:-1] + (5,)).copy()
    u[pts[..., 0] > x0] *= jump
    return u.reshape(-1)


# ---- component-wise isotropic diffusion / per-cell viscosity (the flux path of SURVEY row a24) ------------------
ISO_GOLDEN_CASES = [("iso_p2_hllc", 2, 1), ("iso_p3_exact", 3, 0)]   # uniform nu: Direct<Isotropic<Scalar, 5>>
ISO_VISCOUS = (2.0, 1.0 / 12, 0.03, 0.708)


def cell_viscosities(n_cells, seed=1):
    """A per-cell, per-component viscosity table with zeros in it, like EnergyBasedViscosity produces."""
    rng = np.random.default_rng(seed)
    nu = rng.uniform(0.0, 0.05, (n_cells, 5))
    nu[rng.uniform(size=nu.shape) < 0.3] = 0.0
    return nu


# ---- EnergyBasedViscosity + WithViscosity ----------------------------------------------------------
ENERGY_GOLDEN_CASES = [("energy_p2_hllc", 2, 1), ("energy_p3_hllc", 3, 1)]
ENERGY_NS_GOLDEN_CASES = [("energy_ns_p2_hllc", 2, 1)]   # EnergyBasedViscosity over the Navier-Stokes base
ENERGY_BETAS = (2.0, 1.0 / 12)
ENERGY_AMP = 5e-2          # noise level at which some, not all, (cell, component) pairs get a viscosity


# ---- dg::General on the modal Projection polynomial with Gauss-Legendre rules ------------------
DGM_GOLDEN_CASES = [("dgm_p2_hllc", 2, 1), ("dgm_p2_exact", 2, 0), ("dgm_p3_ausm", 3, 2)]
DGM_WENO_GOLDEN_CASES = [("dgm_weno_p2_hllc", 2, 1), ("dgm_weno_p3_hllc", 3, 1)]   # DGFEM + LIMITER of the reference


# ---- kLocal == true expansions: columns hold det(J) u (demo/laminar/shockless.hpp:66-69; BASELINE config 4) -------
LOCAL_GOLDEN_CASES = [("local_euler_p2_hllc", 2, 1, False), ("local_ns_p2_hllc", 2, 1, True),
                      ("local_ns_p3_exact", 3, 0, True)]    # (name, degree, riemann, viscous)

# frame output (mini::mesh::vtk::Writer): (name, degree, reference scheme id, what the polynomial is)
VTK_GOLDEN_CASES = [("vtk_nodal_p2", 2, 0, "nodal"), ("vtk_nodal_p3", 3, 0, "nodal"),
                    ("vtk_local_p2", 2, 11, "local"), ("vtk_local_p3", 3, 11, "local"),
                    ("vtk_extrapolation_p2", 2, 4, "extrapolation"), ("vtk_extrapolation_p3", 3, 4, "extrapolation"),
                    ("vtk_projection_p2", 2, 9, "projection"), ("vtk_projection_p3", 3, 9, "projection")]
