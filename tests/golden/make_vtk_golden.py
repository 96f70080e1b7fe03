"""Generates tests/golden/vtk_*.npz from the UNMODIFIED reference (oracle/_ref): what
`mini::mesh::vtk::Writer<Part>::Prepare` collects for every cell (points and values of the 27 nodes of its VTK
tri-quadratic hexahedron, full precision) and the `.vtu` / `.pvtu` text `WriteSolutions` produces from them, for the
four kinds of polynomial the drivers use.  Run in the build container only: `python tests/golden/make_vtk_golden.py`."""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import cases  # noqa: E402
import oracle_check as oc  # noqa: E402
import refapi  # noqa: E402
from minicfd_b200.part import cell_points  # noqa: E402


def reference_state(m, degree, scheme, kind):
    """A reference solver holding a rough admissible state; returns (solver, the column it was given)."""
    r = refapi.Ref(m, degree=degree, scheme=scheme, riemann=1, weno=cases.WENO)
    cases.apply_bcs_ref(r)
    n = (degree + 1) ** 3
    if kind == "projection":
        r.approximate(refapi.F_SMOOTH, [1.0])
        base = r.get_solution()
        u = base * (1 + 5e-2 * np.random.default_rng(degree).uniform(-1, 1, base.shape))
    else:
        u = cases.noisy_state(cell_points(m.nodes, m.cells, degree), seed=7, amp=1e-2)
        if kind == "local":       # columns of Hexahedron<..., true> hold det(J) u
            det = oc.Oracle(m, degree, 0, 1).jacobian_determinants()
            u = (u.reshape(-1, n, 5) * det[:, :, None]).reshape(-1)
    r.set_solution_all(u)        # under a limiter: followed by Reconstruct, like the driver's SetSolutionColumn
    return r, u


def main():
    m = cases.small_mesh(seed=11)
    for name, degree, scheme, kind in cases.VTK_GOLDEN_CASES:
        r, u = reference_state(m, degree, scheme, kind)
        points, values = r.sample_vtk()
        with tempfile.TemporaryDirectory() as d:
            r.write_vtu(d, "Frame0")
            vtu = open(os.path.join(d, "Frame0", "0.vtu")).read()
            pvtu = open(os.path.join(d, "Frame0.pvtu")).read()
        np.savez_compressed(os.path.join(HERE, name + ".npz"), u=u, state=r.get_solution(), points=points, values=values,
                            vtu=np.array(vtu), pvtu=np.array(pvtu),
                            nodes=m.nodes, cells=m.cells, bfaces=m.bfaces, btags=m.btags)
        print(name, points.shape, np.abs(values).max(), len(vtu))


if __name__ == "__main__":
    main()
