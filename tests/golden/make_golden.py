"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref, built from /root/reference
by `make -C oracle ref`).  Run in the build container only: `python tests/golden/make_golden.py`.

Each file holds, for one (degree, scheme, Riemann solver): the input column, the reference's residual
at t = GOLDEN_T with the mixed boundary set-up of tests/cases.py, its time step, and the state after
one RungeKutta<3>::Update from t = 0 — all on tests/cases.small_mesh().
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import cases  # noqa: E402
import refapi  # noqa: E402
from minicfd_b200.part import cell_points  # noqa: E402


def main(only=None):
    m = cases.small_mesh()
    if only == "energy_ns":      # `python tests/golden/make_golden.py energy_ns`: just the files added last
        for k in ("GOLDEN_CASES", "VISCOUS_GOLDEN_CASES", "ISO_GOLDEN_CASES", "ENERGY_GOLDEN_CASES", "DGM_GOLDEN_CASES",
                  "DGM_WENO_GOLDEN_CASES", "LOCAL_GOLDEN_CASES", "WENO_GOLDEN_CASES", "WENO_EIGEN_GOLDEN_CASES"):
            setattr(cases, k, [])
    for name, degree, scheme, riemann in cases.GOLDEN_CASES:
        r = refapi.Ref(m, degree=degree, scheme=scheme, riemann=riemann)
        cases.apply_bcs_ref(r)
        u = cases.noisy_state(cell_points(m.nodes, m.cells, degree))
        r.set_solution(u)
        res = r.residual(cases.GOLDEN_T)
        dt = r.time_step(1.0, 3)
        r.rk_update(3, 0.0, cases.GOLDEN_DT, 1)
        u_rk3 = r.get_solution()
        np.savez_compressed(os.path.join(HERE, name + ".npz"), u=u, residual=res, dt=dt, u_rk3=u_rk3,
                            nodes=m.nodes, cells=m.cells, bfaces=m.bfaces, btags=m.btags)
        print(name, np.linalg.norm(res), dt)
    # fr::Lobatto on ConvectionDiffusion<rotated::Euler<..>, Direct<NavierStokes>> with no-slip walls
    for name, degree, riemann in cases.VISCOUS_GOLDEN_CASES:
        r = refapi.Ref(m, degree=degree, scheme=refapi.FR_VISCOUS, riemann=riemann)
        cases.apply_viscous_ref(r)
        u = cases.noisy_state(cell_points(m.nodes, m.cells, degree), amp=1e-4)
        r.set_solution(u)
        res = r.residual(cases.GOLDEN_T)
        dt = r.time_step(1.0, 3)
        r.rk_update(3, 0.0, cases.GOLDEN_DT, 1)
        u_rk3 = r.get_solution()
        np.savez_compressed(os.path.join(HERE, name + ".npz"), u=u, residual=res, dt=dt, u_rk3=u_rk3,
                            nodes=m.nodes, cells=m.cells, bfaces=m.bfaces, btags=m.btags)
        print(name, np.linalg.norm(res), dt)
    # fr::Lobatto on ConvectionDiffusion<rotated::Euler<..>, Direct<Isotropic<Scalar, 5>>>, uniform nu
    for name, degree, riemann in cases.ISO_GOLDEN_CASES:
        r = refapi.Ref(m, degree=degree, scheme=refapi.FR_ISOTROPIC, riemann=riemann)
        cases.apply_viscous_ref(r, vis=cases.ISO_VISCOUS)
        u = cases.noisy_state(cell_points(m.nodes, m.cells, degree), amp=1e-4)
        r.set_solution(u)
        res = r.residual(cases.GOLDEN_T)
        r.rk_update(3, 0.0, cases.GOLDEN_DT, 1)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), u=u, residual=res, u_rk3=r.get_solution(),
                            nodes=m.nodes, cells=m.cells, bfaces=m.bfaces, btags=m.btags)
        print(name, np.linalg.norm(res))
    # WithViscosity<fr::Lobatto<Part, EnergyBasedViscosity<Part, ConvectionDiffusion<Euler, Direct<Isotropic>>>>>
    for name, degree, riemann in cases.ENERGY_GOLDEN_CASES:
        r = refapi.Ref(m, degree=degree, scheme=refapi.FR_ENERGY_VISCOSITY, riemann=riemann,
                       viscous=cases.ENERGY_BETAS + (0.0, 0.708))
        cases.apply_bcs_ref(r)
        u = cases.noisy_state(cell_points(m.nodes, m.cells, degree), amp=cases.ENERGY_AMP)
        damping = r.damping_matrices()          # zeroes the solution
        r.set_solution(u)                       # SetSolutionColumn -> UpdateProperties
        nu = r.cell_viscosity()
        dt = r.time_step(1.0, 3)
        res = r.residual(cases.GOLDEN_T)
        r.rk_update(3, 0.0, cases.GOLDEN_DT, 1)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), u=u, damping=damping[:4], nu=nu, dt=dt, residual=res,
                            u_rk3=r.get_solution(), nu_rk3=r.cell_viscosity(),
                            nodes=m.nodes, cells=m.cells, bfaces=m.bfaces, btags=m.btags)
        print(name, int((nu > 0).sum()), "of", nu.size, "viscosities non-zero; dt", dt)
    # the same rule over the Navier-Stokes base (demo/laminar/shocked.hpp:100-111), viscous boundary set-up, p = 2
    for name, degree, riemann in cases.ENERGY_NS_GOLDEN_CASES:
        r = refapi.Ref(m, degree=degree, scheme=refapi.FR_ENERGY_VISCOSITY_NS, riemann=riemann, viscous=cases.VISCOUS)
        cases.apply_viscous_ref(r)
        u = cases.noisy_state(cell_points(m.nodes, m.cells, degree), amp=cases.ENERGY_AMP)
        damping = r.damping_matrices()          # zeroes the solution
        r.set_solution(u)                       # SetSolutionColumn -> UpdateProperties
        nu = r.cell_viscosity()
        dt = r.time_step(1.0, 3)
        res = r.residual(cases.GOLDEN_T)
        r.rk_update(3, 0.0, cases.GOLDEN_DT, 1)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), u=u, damping=damping[:4], nu=nu, dt=dt, residual=res,
                            u_rk3=r.get_solution(), nu_rk3=r.cell_viscosity(),
                            nodes=m.nodes, cells=m.cells, bfaces=m.bfaces, btags=m.btags)
        print(name, int((nu > 0).sum()), "of", nu.size, "viscosities non-zero; dt", dt)
    # dg::General<Part<Projection<Scalar, 3, P, 5>>, Riemann> with Legendre integrators
    for name, degree, riemann in cases.DGM_GOLDEN_CASES:
        r = refapi.Ref(m, degree=degree, scheme=refapi.DG_GENERAL, riemann=riemann)
        cases.apply_bcs_ref(r)
        r.approximate(refapi.F_SMOOTH, [1.0])
        projected = r.get_solution()
        u = projected * (1 + 1e-3 * np.random.default_rng(degree).uniform(-1, 1, projected.shape))
        r.set_solution(u)
        res = r.residual(cases.GOLDEN_T)
        dt = r.time_step(1.0, 3)
        r.rk_update(3, 0.0, cases.GOLDEN_DT, 1)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), projected=projected, u=u, residual=res, dt=dt,
                            u_rk3=r.get_solution(), nodes=m.nodes, cells=m.cells, bfaces=m.bfaces, btags=m.btags)
        print(name, np.linalg.norm(res), dt)
    # WithLimiter<dg::General<Part<Projection>>, weno::Lazy>: demo/euler/sourceless.hpp under DGFEM + LIMITER
    for name, degree, riemann in cases.DGM_WENO_GOLDEN_CASES:
        r = refapi.Ref(m, degree=degree, scheme=refapi.DG_GENERAL_WENO_LAZY, riemann=riemann, weno=cases.WENO)
        cases.apply_bcs_ref(r)
        r.approximate(refapi.F_SMOOTH, [1.0])
        base = r.get_solution()
        u = base * (1 + 5e-2 * np.random.default_rng(degree).uniform(-1, 1, base.shape))
        r.set_solution_all(u)
        limited = r.get_solution()
        r.rk_update(3, 0.0, cases.GOLDEN_DT, 1)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), u=u, limited=limited, u_rk3=r.get_solution(),
                            nodes=m.nodes, cells=m.cells, bfaces=m.bfaces, btags=m.btags)
        print(name, np.abs(limited - u).max())
    # fr::Lobatto on Hexahedron<..., true>: columns are det(J) u; Euler and Navier-Stokes + direct DG
    for name, degree, riemann, viscous in cases.LOCAL_GOLDEN_CASES:
        r = refapi.Ref(m, degree=degree, scheme=refapi.FR_LOCAL_VISCOUS if viscous else refapi.FR_LOCAL, riemann=riemann)
        if viscous:
            cases.apply_viscous_ref(r)
        else:
            cases.apply_bcs_ref(r)
        r.approximate(refapi.F_CONST, [1, 1, 1, 1, 1])          # SetValue(1): the column is det J
        det = r.get_solution().reshape(m.n_cells, -1, 5)[:, :, 0]
        u = cases.noisy_state(cell_points(m.nodes, m.cells, degree), amp=1e-4).reshape(m.n_cells, -1, 5)
        column = (u * det[:, :, None]).reshape(-1)
        r.set_solution(column)
        res = r.residual(cases.GOLDEN_T)
        dt = r.time_step(1.0, 3)
        r.rk_update(3, 0.0, cases.GOLDEN_DT, 1)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), det=det, u=column, residual=res, dt=dt,
                            u_rk3=r.get_solution(), nodes=m.nodes, cells=m.cells, bfaces=m.bfaces, btags=m.btags)
        print(name, np.linalg.norm(res), dt)
    # WithLimiter<fr::Lobatto, weno::Lazy>: the limiter's action on a column, and an RK3 step with it
    for name, degree, riemann in cases.WENO_GOLDEN_CASES:
        r = refapi.Ref(m, degree=degree, scheme=refapi.FR_WENO_LAZY, riemann=riemann, weno=cases.WENO)
        cases.apply_bcs_ref(r)
        u = cases.noisy_state(cell_points(m.nodes, m.cells, degree), amp=1e-2)
        r.set_solution_all(u)
        limited = r.get_solution()
        r.rk_update(3, 0.0, cases.GOLDEN_DT, 1)
        u_rk3 = r.get_solution()
        np.savez_compressed(os.path.join(HERE, name + ".npz"), u=u, limited=limited, u_rk3=u_rk3,
                            nodes=m.nodes, cells=m.cells, bfaces=m.bfaces, btags=m.btags)
        print(name, np.abs(limited - u).max())
    # WithLimiter<fr::Lobatto, weno::Eigen>: detector + characteristic reconstruction, alone and inside RK3
    for name, degree, riemann in cases.WENO_EIGEN_GOLDEN_CASES:
        r = refapi.Ref(m, degree=degree, scheme=refapi.FR_WENO_EIGEN, riemann=riemann, weno=cases.WENO)
        cases.apply_bcs_ref(r)
        u = cases.shocked_state(cell_points(m.nodes, m.cells, degree))
        r.set_solution_all(u)
        limited = r.get_solution()
        r.rk_update(3, 0.0, cases.GOLDEN_DT, 1)
        u_rk3 = r.get_solution()
        np.savez_compressed(os.path.join(HERE, name + ".npz"), u=u, limited=limited, u_rk3=u_rk3,
                            nodes=m.nodes, cells=m.cells, bfaces=m.bfaces, btags=m.btags)
        kf = 5 * (degree + 1) ** 3
        print(name, int((np.abs((limited - u).reshape(-1, kf)).max(1) > 0).sum()), "troubled cells of", m.n_cells)


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else None)
