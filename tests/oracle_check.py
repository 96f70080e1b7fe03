"""ctypes access to oracle/libminicfd_oracle.so (the plain-C restatement; TEST INFRASTRUCTURE) plus
helpers that pick the strongest checker available: the reference itself (oracle/_ref) when it was
built, else the C restatement."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "oracle", "libminicfd_oracle.so")
_lib = None

F_CONST, F_VORTEX, F_SMOOTH = 0, 1, 3


def available():
    return os.path.exists(LIB)


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(LIB)
        _lib.oracle_create.restype = C.c_void_p
        _lib.oracle_time_step.restype = C.c_double
    return _lib


def _dp(a):
    return None if a is None else a.ctypes.data_as(C.POINTER(C.c_double))


class Oracle:
    def __init__(self, mesh, degree=2, scheme=0, riemann=1, c_next=0.0):
        nodes = np.ascontiguousarray(mesh.nodes, dtype=np.float64)
        cells = np.ascontiguousarray(mesh.cells, dtype=np.int64)
        bf = np.ascontiguousarray(mesh.bfaces, dtype=np.int64)
        bt = np.ascontiguousarray(mesh.btags, dtype=np.int32)
        self.n_cells = len(cells)
        self.kfields = 5 * (degree + 1) ** 3
        if scheme == 9:   # dg::General on the modal Projection polynomial: K x M coefficients per cell
            self.kfields = 5 * {2: 10, 3: 20}[degree]
        self.h = C.c_void_p(lib().oracle_create(
            C.c_int(degree), C.c_int(scheme), C.c_int(riemann), C.c_int64(len(nodes)), _dp(nodes),
            C.c_int64(len(cells)), cells.ctypes.data_as(C.POINTER(C.c_int64)), C.c_int64(len(bf)),
            bf.ctypes.data_as(C.POINTER(C.c_int64)), bt.ctypes.data_as(C.POINTER(C.c_int32))))
        lib().oracle_set_vincent(self.h, C.c_double(c_next))

    def __del__(self):
        try:
            if self.h:
                lib().oracle_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def set_local(self, klocal=True):
        """Hexahedron<..., true>: columns hold det(J) u"""
        lib().oracle_set_local(self.h, C.c_int(1 if klocal else 0))

    def jacobian_determinants(self):
        """det J at the nodes, [n_cells][N]: U = det u"""
        self.set_local(False)
        ones = np.ones(self.n_cells * self.kfields)
        n = self.kfields // 5
        out = np.empty((self.n_cells, n))
        lib().oracle_node_determinants(self.h, _dp(out))
        return out

    def set_viscous(self, beta0, beta1, nu, prandtl):
        lib().oracle_set_viscous(self.h, C.c_double(beta0), C.c_double(beta1), C.c_double(nu), C.c_double(prandtl))

    def set_cell_viscosity(self, nu):
        """scheme 6: one viscosity per cell and component ([n_cells][5]); None = the uniform set_viscous value"""
        a = None if nu is None else np.ascontiguousarray(nu, dtype=np.float64).reshape(self.n_cells, 5)
        self._nu_keep = a
        lib().oracle_set_cell_viscosity(self.h, _dp(a))

    def damping_matrices(self):
        """EnergyBasedViscosity::BuildDampingMatrices: [n_cells][N][N] (uses the beta values of set_viscous)"""
        n = self.kfields // 5
        out = np.empty((self.n_cells, n, n))
        lib().oracle_damping_matrices(self.h, _dp(out))
        return out

    def update_viscosity(self, u, damping, time_scale=1.0):
        """EnergyBasedViscosity::UpdateProperties: (nu[n_cells][5], min_dt)"""
        u = np.ascontiguousarray(u, dtype=np.float64)
        d = np.ascontiguousarray(damping, dtype=np.float64)
        nu = np.empty((self.n_cells, 5))
        f = lib().oracle_update_viscosity
        f.restype = C.c_double
        min_dt = f(self.h, _dp(u), _dp(d), C.c_double(time_scale), _dp(nu))
        return nu, float(min_dt)

    def set_bc(self, tag, bc_kind, field_kind=F_CONST, params=None):
        p = np.zeros(8)
        if params is not None:
            p[:len(params)] = params
        lib().oracle_set_bc(self.h, C.c_int(tag), C.c_int(bc_kind), C.c_int(field_kind), _dp(p))

    def approximate(self, field_kind, params=None):
        p = np.zeros(8)
        if params is not None:
            p[:len(params)] = params
        out = np.empty(self.n_cells * self.kfields)
        lib().oracle_approximate(self.h, C.c_int(field_kind), _dp(p), _dp(out))
        return out

    def residual(self, u, t=0.0):
        u = np.ascontiguousarray(u, dtype=np.float64)
        out = np.empty_like(u)
        lib().oracle_residual(self.h, C.c_double(t), _dp(u), _dp(out))
        return out

    def rk_update(self, u, order, t, dt):
        u = np.array(u, dtype=np.float64, copy=True)
        lib().oracle_rk_update(self.h, C.c_int(order), C.c_double(t), C.c_double(dt), _dp(u))
        return u

    def limiter(self, u, kind):
        u = np.array(u, dtype=np.float64, copy=True)
        lib().oracle_limiter(self.h, C.c_int(kind), _dp(u))
        return u

    def set_limiter(self, w0=0.001, eps=1e-6, on=True, eigen=False):
        """rk_update then applies weno::Lazy (or weno::Eigen) after every stage, like WithLimiter
        (with_limiter.hpp:57-61)."""
        lib().oracle_set_limiter(self.h, C.c_int((2 if eigen else 1) if on else 0), C.c_double(w0), C.c_double(eps))

    def weno_eigen(self, u, w0=0.001, eps=1e-6):
        u = np.array(u, dtype=np.float64, copy=True)
        lib().oracle_weno_eigen(self.h, C.c_double(w0), C.c_double(eps), _dp(u))
        return u

    def weno_lazy(self, u, w0=0.001, eps=1e-6):
        u = np.array(u, dtype=np.float64, copy=True)
        f = lib().oracle_dgm_weno_lazy if self.kfields in (50, 100) else lib().oracle_weno_lazy
        f(self.h, C.c_double(w0), C.c_double(eps), _dp(u))
        return u

    def time_step(self, u, dt_guess, rk_order):
        u = np.ascontiguousarray(u, dtype=np.float64)
        return float(lib().oracle_time_step(self.h, _dp(u), C.c_double(dt_guess), C.c_int(rk_order)))


def riemann_flux(kind, frame, ul, ur, bc=-1):
    out = np.empty(5)
    fr = np.ascontiguousarray(frame, dtype=np.float64)
    ul = np.ascontiguousarray(ul, dtype=np.float64); ur = np.ascontiguousarray(ur, dtype=np.float64)
    lib().oracle_riemann_flux(C.c_int(kind), C.c_int(bc), _dp(fr), _dp(ul), _dp(ur), _dp(out))
    return out


def riemann_1d(kind, pl, pr):
    out = np.empty(5)
    pl = np.ascontiguousarray(pl, dtype=np.float64); pr = np.ascontiguousarray(pr, dtype=np.float64)
    lib().oracle_riemann_1d(C.c_int(kind), _dp(pl), _dp(pr), _dp(out))
    return out


def which():
    import refapi
    return "reference (oracle/_ref)" if refapi.available(2) else "C restatement (oracle/minicfd_oracle.c)"


def residual_and_rk3(mesh, u0, *, degree, scheme, riemann, dt, walls=True):
    """Residual at t=0 and the state after one RK3 step, all boundaries inviscid walls."""
    import refapi
    if refapi.available(degree):
        r = refapi.Ref(mesh, degree=degree, scheme=scheme, riemann=riemann)
        for tag in range(6):
            r.set_bc(tag, refapi.WALL)
        r.set_solution(u0)
        res = r.residual(0.0)
        r.rk_update(3, 0.0, dt, 1)
        return res, r.get_solution()
    o = Oracle(mesh, degree, scheme, riemann)
    for tag in range(6):
        o.set_bc(tag, 0)
    return o.residual(u0, 0.0), o.rk_update(u0, 3, 0.0, dt)
