"""ctypes access to tests/host_emulation/libmcfd_emu.so: the kernel *source* compiled for the host
(TEST INFRASTRUCTURE; lets `-m "not gpu"` tests check the kernel's index logic against the
reference).  Never imported by the product package."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "host_emulation", "emu.cpp")
LIB = os.path.join(HERE, "host_emulation", "libmcfd_emu.so")
_lib = None


def build(force=False):
    deps = [SRC] + [os.path.join(HERE, "..", "minicfd_b200", "csrc", f)
                    for f in ("kernels.cuh", "stage_lines.cuh", "physics.cuh", "tables.hpp", "format.cuh")]
    if force or not os.path.exists(LIB) or any(os.path.getmtime(d) > os.path.getmtime(LIB) for d in deps):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-ffp-contract=off",
                               "-o", LIB, SRC])


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(LIB)
    return _lib


def _dp(a):
    return None if a is None else a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a):
    return None if a is None else a.ctypes.data_as(C.POINTER(C.c_int64))


def _i32(a):
    return None if a is None else a.ctypes.data_as(C.POINTER(C.c_int32))


def stage(mesh, degree, scheme, riemann, column_in, *, n_local=None, n_ghost=0, tag_kinds=None,
          bvals=None, column_u0=None, mode=0, dt=0.0, cells=None, bfaces=None, btags=None, c_next=0.0,
          viscous=None, cell_nu=None, line_cells=-1):
    cells = np.ascontiguousarray(mesh.cells if cells is None else cells, dtype=np.int64)
    bfaces = np.ascontiguousarray(mesh.bfaces if bfaces is None else bfaces, dtype=np.int64)
    btags = np.ascontiguousarray(mesh.btags if btags is None else btags, dtype=np.int32)
    nodes = np.ascontiguousarray(mesh.nodes, dtype=np.float64)
    if n_local is None:
        n_local = len(cells)
    F = 5 * (degree + 1) ** 3
    column_in = np.ascontiguousarray(column_in, dtype=np.float64)
    assert column_in.size == (n_local + n_ghost) * F
    u0 = column_in[:n_local * F].copy() if column_u0 is None else np.ascontiguousarray(column_u0)
    kinds = np.zeros(16, dtype=np.int32)
    if tag_kinds is not None:
        for k, v in tag_kinds.items():
            kinds[k] = v
    pairs = None if mesh.periodic is None else np.ascontiguousarray(mesh.periodic, dtype=np.int64)
    out = np.empty(n_local * F)
    err = C.create_string_buffer(256)
    bv = None if bvals is None else np.ascontiguousarray(bvals, dtype=np.float64)
    rc = lib().emu_stage(C.c_int(degree), C.c_int(scheme), C.c_int(riemann), C.c_int64(len(nodes)),
                         _dp(nodes), C.c_int64(n_local), C.c_int64(n_ghost), _ip(cells),
                         C.c_int64(len(bfaces)), _ip(bfaces), _i32(btags),
                         C.c_int64(0 if pairs is None else len(pairs)), _ip(pairs), _i32(kinds), _dp(bv),
                         _dp(column_in), _dp(u0), C.c_int(mode), C.c_double(dt), _dp(out), err, C.c_int(256),
                         C.c_double(c_next),
                         None if viscous is None else _dp(np.ascontiguousarray(
                             (list(viscous) + [0.0])[:5], dtype=np.float64)),
                         None if cell_nu is None else _dp(np.ascontiguousarray(cell_nu, dtype=np.float64)),
                         C.c_int(line_cells))
    if rc != 0:
        raise RuntimeError(err.value.decode())
    return out


def weno_lazy(mesh, degree, column_in, w0=0.001, eps=1e-6, *, n_local=None, n_ghost=0, cells=None, bfaces=None,
              btags=None, eigen=False, global_ids=None):
    cells = np.ascontiguousarray(mesh.cells if cells is None else cells, dtype=np.int64)
    bfaces = np.ascontiguousarray(mesh.bfaces if bfaces is None else bfaces, dtype=np.int64)
    btags = np.ascontiguousarray(mesh.btags if btags is None else btags, dtype=np.int32)
    nodes = np.ascontiguousarray(mesh.nodes, dtype=np.float64)
    if n_local is None:
        n_local = len(cells)
    F = 5 * (degree + 1) ** 3
    column_in = np.ascontiguousarray(column_in, dtype=np.float64)
    assert column_in.size == (n_local + n_ghost) * F
    pairs = None if mesh.periodic is None else np.ascontiguousarray(mesh.periodic, dtype=np.int64)
    out = np.empty(n_local * F)
    rc = lib().emu_weno_lazy(C.c_int(degree), C.c_int64(len(nodes)), _dp(nodes), C.c_int64(n_local), C.c_int64(n_ghost),
                             _ip(cells), C.c_int64(len(bfaces)), _ip(bfaces), _i32(btags),
                             C.c_int64(0 if pairs is None else len(pairs)), _ip(pairs), C.c_double(w0), C.c_double(eps),
                             _dp(column_in), _dp(out), C.c_int(1 if eigen else 0),
                             None if global_ids is None else _ip(np.ascontiguousarray(global_ids, dtype=np.int64)))
    if rc != 0:
        raise RuntimeError("emu_weno_lazy failed")
    return out


def energy_viscosity(mesh, degree, betas, column=None, time_scale=1.0):
    """(damping[n_cells][N][N], nu[n_cells][5] or None, min_dt or None) from the device code run on the host."""
    cells = np.ascontiguousarray(mesh.cells, dtype=np.int64)
    bfaces = np.ascontiguousarray(mesh.bfaces, dtype=np.int64)
    btags = np.ascontiguousarray(mesh.btags, dtype=np.int32)
    nodes = np.ascontiguousarray(mesh.nodes, dtype=np.float64)
    n = (degree + 1) ** 3
    damping = np.empty((len(cells), n, n)); nu = np.empty((len(cells), 5)); min_dt = C.c_double(0)
    b = np.ascontiguousarray(betas, dtype=np.float64)
    col = None if column is None else np.ascontiguousarray(column, dtype=np.float64)
    rc = lib().emu_energy_viscosity(C.c_int(degree), C.c_int64(len(nodes)), _dp(nodes), C.c_int64(len(cells)), _ip(cells),
                                    C.c_int64(len(bfaces)), _ip(bfaces), _i32(btags), _dp(b), _dp(col),
                                    C.c_double(time_scale), _dp(damping), _dp(nu), C.byref(min_dt))
    if rc != 0:
        raise RuntimeError("emu_energy_viscosity failed")
    return damping, (None if column is None else nu), (None if column is None else min_dt.value)


def dgm(mesh, degree, riemann=1, column_in=None, *, tag_kinds=None, bvals=None, column_u0=None, mode=0, dt=0.0,
        values=None, want_points=False):
    """dg::General on the modal polynomial (single part): one stage / residual, or the L2 projection of `values`
    given at the cells' Gauss points; want_points=True returns (boundary Gauss points, cell Gauss points) only."""
    cells = np.ascontiguousarray(mesh.cells, dtype=np.int64)
    bfaces = np.ascontiguousarray(mesh.bfaces, dtype=np.int64)
    btags = np.ascontiguousarray(mesh.btags, dtype=np.int32)
    nodes = np.ascontiguousarray(mesh.nodes, dtype=np.float64)
    q = degree + 1
    F = 5 * (10 if degree == 2 else 20)
    kinds = np.zeros(16, dtype=np.int32)
    for k, v in (tag_kinds or {}).items():
        kinds[k] = v
    gp = np.zeros((len(bfaces), q * q, 3)); qp = np.zeros((len(cells), q ** 3, 3))
    out = None if want_points else np.empty(len(cells) * F)
    arr = lambda a: None if a is None else np.ascontiguousarray(a, dtype=np.float64)
    ci, cu, bv, vals = arr(column_in), arr(column_u0), arr(bvals), arr(values)
    rc = lib().emu_dgm_stage(C.c_int(degree), C.c_int(riemann), C.c_int64(len(nodes)), _dp(nodes), C.c_int64(len(cells)),
                             C.c_int64(0), _ip(cells), C.c_int64(len(bfaces)), _ip(bfaces), _i32(btags), _i32(kinds),
                             _dp(bv), _dp(ci), _dp(cu), C.c_int(mode), C.c_double(dt), _dp(vals), _dp(out), _dp(gp), _dp(qp))
    if rc != 0:
        raise RuntimeError("emu_dgm_stage failed")
    return (gp, qp) if want_points else out


def riemann(kind, n, ul, ur, bc=-1):
    out = np.empty(5)
    n = np.ascontiguousarray(n, dtype=np.float64)
    ul = np.ascontiguousarray(ul, dtype=np.float64); ur = np.ascontiguousarray(ur, dtype=np.float64)
    lib().emu_riemann(C.c_int(kind), C.c_int(bc), _dp(n), _dp(ul), _dp(ur), _dp(out))
    return out


def boundary(mesh, degree, *, n_local=None, n_ghost=0, cells=None, bfaces=None, btags=None):
    """Boundary-face table in the library's order: (holder, holder_face, tag, points[nb, Q*Q, 3])."""
    cells = np.ascontiguousarray(mesh.cells if cells is None else cells, dtype=np.int64)
    bfaces = np.ascontiguousarray(mesh.bfaces if bfaces is None else bfaces, dtype=np.int64)
    btags = np.ascontiguousarray(mesh.btags if btags is None else btags, dtype=np.int32)
    nodes = np.ascontiguousarray(mesh.nodes, dtype=np.float64)
    if n_local is None:
        n_local = len(cells)
    nb = len(bfaces)
    fq = (degree + 1) ** 2
    holder = np.empty(nb, dtype=np.int64); hf = np.empty(nb, dtype=np.int32); tg = np.empty(nb, dtype=np.int32)
    pts = np.zeros((nb, fq, 3))
    f = lib().emu_boundary
    f.restype = C.c_int64
    n = f(C.c_int(degree), C.c_int64(len(nodes)), _dp(nodes), C.c_int64(n_local), C.c_int64(n_ghost),
          _ip(cells), C.c_int64(nb), _ip(bfaces), _i32(btags), _ip(holder), _i32(hf), _i32(tg), _dp(pts))
    assert n >= 0
    return holder[:n], hf[:n], tg[:n], pts[:n]


def format_g6(values):
    """FormatG6 (format.cuh) on the host: (list of token bytes, unsure flags)."""
    v = np.ascontiguousarray(values, dtype=np.float64).reshape(-1)
    slots = np.zeros((v.size, 16), dtype=np.uint8)
    lens = np.zeros(v.size, dtype=np.int32)
    unsure = np.zeros(v.size, dtype=np.uint8)
    lib().emu_format_g6(C.c_int64(v.size), _dp(v), slots.ctypes.data_as(C.c_char_p), _i32(lens),
                        unsure.ctypes.data_as(C.c_char_p))
    return [bytes(slots[i, :lens[i]]) for i in range(v.size)], unsure.astype(bool)


def format_integer(values):
    v = np.ascontiguousarray(values, dtype=np.int64).reshape(-1)
    slots = np.zeros((v.size, 16), dtype=np.uint8)
    lens = np.zeros(v.size, dtype=np.int32)
    lib().emu_format_integer(C.c_int64(v.size), _ip(v), slots.ctypes.data_as(C.c_char_p), _i32(lens))
    return [bytes(slots[i, :lens[i]]) for i in range(v.size)]
