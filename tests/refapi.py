"""ctypes access to oracle/_ref/libminicfd_ref_p{2,3}.so (TEST INFRASTRUCTURE: the unmodified reference
headers behind oracle/ref_harness.cpp).  Never imported by the product package."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_libs: dict[int, C.CDLL] = {}

CB = C.CFUNCTYPE(None, C.c_int, C.POINTER(C.c_double), C.c_double, C.POINTER(C.c_double), C.c_void_p)

FR, DG, FR_GENERAL, FR_VISCOUS, FR_WENO_LAZY, FR_WENO_EIGEN, FR_ISOTROPIC, FR_ENERGY_VISCOSITY, FR_GENERAL_VISCOUS, DG_GENERAL, DG_GENERAL_WENO_LAZY, FR_LOCAL, FR_LOCAL_VISCOUS, DG_WENO_LAZY = 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13
FR_ENERGY_VISCOSITY_NS = 20   # EnergyBasedViscosity over the Navier-Stokes base (demo/laminar/shocked.hpp:100-111); HLLC, one part
EXACT, HLLC, AUSM = 0, 1, 2
# boundary kinds, order of fem.hpp:196-222 as used by ref_harness.cpp
WALL, SUPERSONIC_OUTLET, SUPERSONIC_INLET, SUBSONIC_INLET, SUBSONIC_OUTLET, SMART, NO_SLIP = range(7)
# field kinds of ref_harness.cpp::EvalField
F_CONST, F_VORTEX, F_CALLBACK, F_SMOOTH = range(4)


def lib_path(degree: int) -> str:
    return os.path.join(ROOT, "oracle", "_ref", f"libminicfd_ref_p{degree}.so")


def available(degree: int = 2) -> bool:
    return os.path.exists(lib_path(degree))


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int64))


def _i32(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


class Ref:
    def __init__(self, mesh, degree=2, scheme=FR, riemann=HLLC, cell_parts=None, n_parts=1, c_next=0.0,
                 weno=(0.001, 1e-6), viscous=None):
        if degree not in _libs:
            _libs[degree] = C.CDLL(lib_path(degree))
        self.lib = _libs[degree]
        self.degree = degree
        self.pre = f"ref_p{degree}_"
        self.n_nodes_per_cell = (degree + 1) ** 3
        self.kfields = 5 * self.n_nodes_per_cell
        if scheme in (9, 10):   # dg::General on Projection: K x M modal coefficients per cell
            self.kfields = 5 * {2: 10, 3: 20}[degree]
        self._f("set_vincent")(C.c_double(c_next))
        self._f("set_weno")(C.c_double(weno[0]), C.c_double(weno[1]))
        if viscous is not None:   # class statics: EnergyBasedViscosity builds its damping matrices at construction
            self.set_viscous(*viscous)
        f = self._f("create")
        f.restype = C.c_void_p
        nodes = np.ascontiguousarray(mesh.nodes, dtype=np.float64)
        cells = np.ascontiguousarray(mesh.cells, dtype=np.int64)
        bf = np.ascontiguousarray(mesh.bfaces, dtype=np.int64)
        bt = np.ascontiguousarray(mesh.btags, dtype=np.int32)
        parts = None if cell_parts is None else np.ascontiguousarray(cell_parts, dtype=np.int32)
        self.h = C.c_void_p(f(C.c_int(scheme), C.c_int(riemann), C.c_int64(len(nodes)), _dp(nodes),
                              C.c_int64(len(cells)), _ip(cells), None if parts is None else _i32(parts),
                              C.c_int(n_parts), C.c_int64(len(bf)), _ip(bf), _i32(bt)))
        assert self.h.value
        self.n_parts = n_parts
        self._keep = []

    def _f(self, name):
        return getattr(self.lib, self.pre + name)

    def close(self):
        if self.h:
            self._f("destroy")(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def count_local(self, r=0):
        f = self._f("count_local"); f.restype = C.c_int64
        return int(f(self.h, C.c_int(r)))

    def count_ghost(self, r=0):
        f = self._f("count_ghost"); f.restype = C.c_int64
        return int(f(self.h, C.c_int(r)))

    def cell_global_ids(self, r=0):
        out = np.empty(self.count_local(r) + self.count_ghost(r), dtype=np.int64)
        self._f("cell_global_ids")(self.h, C.c_int(r), _ip(out))
        return out

    def faces(self, r, kind):
        f = self._f("count_faces"); f.restype = C.c_int64
        n = int(f(self.h, C.c_int(r), C.c_int(kind)))
        holder = np.empty(n, dtype=np.int64); sharer = np.empty(n, dtype=np.int64)
        hf = np.empty(n, dtype=np.int32)
        self._f("face_table")(self.h, C.c_int(r), C.c_int(kind), _ip(holder), _ip(sharer), _i32(hf))
        return holder, sharer, hf

    def halo(self, r, send):
        n = C.c_int64(0)
        self._f("halo_table")(self.h, C.c_int(r), C.c_int(send), C.byref(n), None, None, None)
        parts = np.empty(n.value, dtype=np.int64); counts = np.empty(n.value, dtype=np.int64)
        self._f("halo_table")(self.h, C.c_int(r), C.c_int(send), C.byref(n), _ip(parts), _ip(counts), None)
        cells = np.empty(int(counts.sum()), dtype=np.int64)
        self._f("halo_table")(self.h, C.c_int(r), C.c_int(send), C.byref(n), _ip(parts), _ip(counts), _ip(cells))
        return parts, counts, cells

    def _field(self, kind, params, cb):
        p = np.zeros(16, dtype=np.float64)
        if params is not None:
            p[:len(params)] = params
        c = CB(cb) if cb is not None else C.cast(None, CB)
        self._keep.append(c)
        return C.c_int(kind), _dp(p), C.c_int(16), c, None

    def set_viscous(self, beta0, beta1, nu, prandtl):
        """class statics of the reference (direct.hpp:80-83, navier_stokes.hpp:57-60): process-wide"""
        self._f("set_viscous")(C.c_double(beta0), C.c_double(beta1), C.c_double(nu), C.c_double(prandtl))

    def set_time_scale(self, time_scale):
        """EnergyBasedViscosity::SetTimeScale (scheme FR_ENERGY_VISCOSITY; after construction)"""
        self._f("set_time_scale")(C.c_double(time_scale))

    def set_bc(self, tag, bc_kind, field_kind=F_CONST, params=None, cb=None):
        self._f("set_bc")(self.h, C.c_int(tag), C.c_int(bc_kind), *self._field(field_kind, params, cb))

    def approximate(self, field_kind, params=None, cb=None):
        self._f("approximate")(self.h, *self._field(field_kind, params, cb))

    def set_solution(self, col, r=0):
        col = np.ascontiguousarray(col, dtype=np.float64)
        assert col.size == self.count_local(r) * self.kfields
        self._f("set_solution")(self.h, C.c_int(r), _dp(col))

    def set_solution_all(self, cols):
        """Every part's column at once (list, or one concatenated array): under a limiter the reference's
        SetSolutionColumn exchanges ghost cells, so all parts must take part."""
        col = np.ascontiguousarray(np.concatenate([np.asarray(c, dtype=np.float64).reshape(-1) for c in cols])
                                   if isinstance(cols, (list, tuple)) else cols, dtype=np.float64)
        assert col.size == sum(self.count_local(r) for r in range(self.n_parts)) * self.kfields
        self._f("set_solution_all")(self.h, _dp(col))

    def cell_viscosity(self):
        """EnergyBasedViscosity::properties_ of the local cells (scheme FR_ENERGY_VISCOSITY): [n_local][5]"""
        out = np.empty((self.count_local(0), 5))
        assert self._f("cell_viscosity")(self.h, _dp(out)) == 0
        return out

    def damping_matrices(self):
        """EnergyBasedViscosity::BuildDampingMatrices(): [n_local][N][N]; leaves the solution zeroed"""
        n = self.n_nodes_per_cell
        out = np.empty((self.count_local(0), n, n))
        assert self._f("damping_matrices")(self.h, _dp(out)) == 0
        return out

    def get_solution(self, r=0):
        out = np.empty(self.count_local(r) * self.kfields)
        self._f("get_solution")(self.h, C.c_int(r), _dp(out))
        return out

    def residual(self, t=0.0):
        self._f("residual_all")(self.h, C.c_double(t))
        outs = []
        for r in range(self.n_parts):
            out = np.empty(self.count_local(r) * self.kfields)
            self._f("get_residual")(self.h, C.c_int(r), _dp(out))
            outs.append(out)
        return outs if self.n_parts > 1 else outs[0]

    def rk_update(self, order, t, dt, n_steps=1):
        self._f("rk_update")(self.h, C.c_int(order), C.c_double(t), C.c_double(dt), C.c_int(n_steps))

    def time_step(self, dt_guess, rk_order):
        f = self._f("time_step"); f.restype = C.c_double
        return float(f(self.h, C.c_double(dt_guess), C.c_int(rk_order)))

    def sample_vtk(self, r=0):
        """What vtk::Writer<Part>::Prepare collects: (points [n_local][27][3], values [n_local][27][5])"""
        n = self.count_local(r)
        pts, val = np.empty((n, 27, 3)), np.empty((n, 27, 5))
        self._f("sample_vtk")(self.h, C.c_int(r), _dp(pts), _dp(val))
        return pts, val

    def write_vtu(self, directory, soln_name):
        """mini::mesh::vtk::Writer<Part>::WriteSolutions(part, soln_name) for every part, files under `directory`"""
        self._f("write_vtu")(self.h, C.c_char_p(str(directory).encode()), C.c_char_p(soln_name.encode()))

    def limiter(self, kind):
        self._f("limiter")(self.h, C.c_int(kind))

    def cell_geometry(self, r=0):
        n = self.count_local(r)
        length = np.empty(n); volume = np.empty(n)
        self._f("cell_geometry")(self.h, C.c_int(r), _dp(length), _dp(volume))
        return length, volume

    def riemann_flux(self, kind, frame, ul, ur, bc=-1):
        out = np.empty(5)
        fr = np.ascontiguousarray(frame, dtype=np.float64)
        ul = np.ascontiguousarray(ul, dtype=np.float64); ur = np.ascontiguousarray(ur, dtype=np.float64)
        self._f("riemann_flux")(C.c_int(kind), C.c_int(bc), _dp(fr), _dp(ul), _dp(ur), _dp(out))
        return out
