"""ctypes binding of libminicfd_b200.so (include/minicfd_b200.h).  No CPU fallback: importing works
anywhere, but every compute call needs the built CUDA library and a GPU and raises otherwise."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libminicfd_b200.so")

FR_LOBATTO, DG_LOBATTO, FR_GENERAL, DG_GENERAL = 0, 1, 2, 3
EXACT, HLLC, AUSM = 0, 1, 2
INVISCID, NAVIER_STOKES_DDG, CELL_VISCOSITY_DDG, NAVIER_STOKES_CELL_VISCOSITY_DDG = 0, 1, 2, 3
(BC_INVISCID_WALL, BC_SUPERSONIC_OUTLET, BC_SUPERSONIC_INLET, BC_SUBSONIC_INLET, BC_SUBSONIC_OUTLET,
 BC_SMART, BC_NO_SLIP_WALL) = range(7)

TIME_CALLBACK = C.CFUNCTYPE(None, C.c_void_p, C.c_double)


class Config(C.Structure):
    _fields_ = [("scheme", C.c_int32), ("degree", C.c_int32), ("riemann", C.c_int32), ("device", C.c_int32),
                ("fr_c_next", C.c_double), ("diffusion", C.c_int32), ("k_local", C.c_int32)]


class McfdError(RuntimeError):
    pass


_lib = None

# every symbol include/minicfd_b200.h declares: (name, restype, argtypes)
_P = C.c_void_p
_D = C.POINTER(C.c_double)
_I64 = C.POINTER(C.c_int64)
_I32 = C.POINTER(C.c_int32)
SYMBOLS = {
    "mcfd_create": (C.c_int, [C.POINTER(Config), C.POINTER(_P)]),
    "mcfd_destroy": (C.c_int, [_P]),
    "mcfd_last_error": (C.c_char_p, [_P]),
    "mcfd_set_stream": (C.c_int, [_P, _P]),
    "mcfd_set_mesh": (C.c_int, [_P, C.c_int64, _D, C.c_int64, C.c_int64, _I64, C.c_int64, _I64, _I32,
                                C.c_int64, _I64]),
    "mcfd_set_cell_global_ids": (C.c_int, [_P, C.c_int64, _I64]),
    "mcfd_get_faces": (C.c_int64, [_P, C.c_int, _I64, _I64, _I32, _I32, _I32]),
    "mcfd_get_cell_geometry": (C.c_int, [_P, _D, _D]),
    "mcfd_set_halo": (C.c_int, [_P, C.c_int32, _I64, _I64, C.c_int32, _I64, _I64]),
    "mcfd_halo_buffers": (C.c_int, [_P, C.POINTER(_P), _I64, C.POINTER(_P), _I64]),
    "mcfd_halo_pack": (C.c_int, [_P]),
    "mcfd_halo_unpack": (C.c_int, [_P]),
    "mcfd_set_boundary": (C.c_int, [_P, C.c_int32, C.c_int32]),
    "mcfd_count_boundary_faces": (C.c_int64, [_P, C.c_int32]),
    "mcfd_get_boundary_points": (C.c_int, [_P, C.c_int32, _D]),
    "mcfd_set_boundary_values": (C.c_int, [_P, C.c_int32, _D]),
    "mcfd_set_diffusion": (C.c_int, [_P, C.c_double, C.c_double, C.c_double, C.c_double]),
    "mcfd_set_cell_viscosity": (C.c_int, [_P, C.c_int64, _D]),
    "mcfd_get_cell_viscosity": (C.c_int, [_P, _D]),
    "mcfd_viscosity_halo_buffers": (C.c_int, [_P, C.POINTER(_P), _I64, C.POINTER(_P), _I64]),
    "mcfd_viscosity_halo_pack": (C.c_int, [_P]),
    "mcfd_viscosity_halo_unpack": (C.c_int, [_P]),
    "mcfd_build_damping_matrices": (C.c_int, [_P]),
    "mcfd_get_damping_matrices": (C.c_int, [_P, _D]),
    "mcfd_update_cell_viscosity": (C.c_int, [_P, C.c_double]),
    "mcfd_get_quadrature_points": (C.c_int, [_P, _D]),
    "mcfd_get_jacobian_determinants": (C.c_int, [_P, _D]),
    "mcfd_sample_vtk": (C.c_int, [_P, C.c_int32, _D, _D]),
    "mcfd_get_node_parts": (C.c_int, [C.c_int32, C.c_int64, C.c_int64, _I64, _I64, _I64, C.c_int64, _I64]),
    "mcfd_get_new_order": (C.c_int, [C.c_int32, C.c_int64, _I64, _I64, _I64]),
    "mcfd_shuffle_data": (C.c_int, [C.c_int32, C.c_int64, C.c_int64, _I64, _D]),
    "mcfd_shuffle_connectivity": (C.c_int, [C.c_int32, C.c_int64, C.c_int32, C.c_int64, _I64, _I64, _I64]),
    "mcfd_vtk_component_text": (C.c_int, [_P, C.c_int32, C.c_char_p, C.c_int64, _I64]),
    "mcfd_format_doubles": (C.c_int, [_P, C.c_int64, _D, C.c_char_p, C.c_int64, _I64]),
    "mcfd_format_integers": (C.c_int, [_P, C.c_int64, C.c_int64, C.c_int64, C.c_char_p, C.c_int64, _I64]),
    "mcfd_project": (C.c_int, [_P, _D]),
    "mcfd_set_time_callback": (C.c_int, [_P, TIME_CALLBACK, _P]),
    "mcfd_set_time_dependent": (C.c_int, [_P, C.c_int]),
    "mcfd_count_local_cells": (C.c_int64, [_P]),
    "mcfd_cell_data_size": (C.c_int64, [_P]),
    "mcfd_set_solution": (C.c_int, [_P, _D]),
    "mcfd_get_solution": (C.c_int, [_P, _D]),
    "mcfd_set_solution_device": (C.c_int, [_P, _P]),
    "mcfd_get_solution_device": (C.c_int, [_P, _P]),
    "mcfd_residual": (C.c_int, [_P, C.c_double, _D, _P]),
    "mcfd_rk_step": (C.c_int, [_P, C.c_int, C.c_double, C.c_double]),
    "mcfd_rk_begin": (C.c_int, [_P, C.c_int, C.c_double, C.c_double]),
    "mcfd_rk_stage_inner": (C.c_int, [_P, C.c_int]),
    "mcfd_rk_stage_inter": (C.c_int, [_P, C.c_int]),
    "mcfd_rk_stage_time": (C.c_double, [_P, C.c_int]),
    "mcfd_time_step": (C.c_int, [_P, C.c_double, C.c_int, _D]),
    "mcfd_time_step_device": (C.c_int, [_P, C.c_double, C.c_int, _P]),
    "mcfd_rk_begin_device": (C.c_int, [_P, C.c_int, C.c_double, _P]),
    "mcfd_rk_step_device": (C.c_int, [_P, C.c_int, C.c_double, _P]),
    "mcfd_time_step_static": (C.c_int, [C.c_int32, C.c_int32, C.c_int64, _P, _P, C.c_double, C.c_int32, _P, _P]),
    "mcfd_limiter_reconstruct": (C.c_int, [_P, C.c_int]),
    "mcfd_limiter_weno_lazy": (C.c_int, [_P, C.c_double, C.c_double, C.c_int]),
    "mcfd_limiter_weno_eigen": (C.c_int, [_P, C.c_double, C.c_double, C.c_int]),
    "mcfd_eigen_matrices": (C.c_int, [_P, C.c_int64, _D, _D, _D, _D]),
    "mcfd_launch_count": (C.c_int64, [_P]),
    "mcfd_stage_kernel": (C.c_char_p, [_P]),
    "mcfd_enable_kernel_timing": (C.c_int, [_P, C.c_int]),
    "mcfd_kernel_time": (C.c_int, [_P, _D, _I64, C.c_int]),
    "mcfd_riemann_flux": (C.c_int, [_P, C.c_int32, C.c_int64, _D, _D, _D, _D]),
    "mcfd_peer_export": (C.c_int, [_P, C.c_int32, _P, C.POINTER(_P)]),
    "mcfd_peer_open": (C.c_int, [_P, _P, C.POINTER(_P)]),
    "mcfd_peer_connect": (C.c_int, [_P, C.c_int32, C.c_int32, C.c_int32, C.POINTER(_P), _I32, C.POINTER(_P), _I64, _I64,
                                    C.POINTER(_P), C.POINTER(_P)]),
    "mcfd_peer_push": (C.c_int, [_P]),
    "mcfd_peer_wait": (C.c_int, [_P]),
    "mcfd_peer_min_time_step": (C.c_int, [_P, _P]),
    "mcfd_peer_step": (C.c_int, [_P, C.c_int, C.c_double, C.c_double]),
    "mcfd_peer_step_group": (C.c_int, [C.POINTER(_P), C.c_int32, C.c_int, C.c_double, C.c_double]),
    "mcfd_peer_join": (C.c_int, [_P]),
    "mcfd_peer_check": (C.c_int, [_P]),
}


def load():
    """Load the CUDA library; raises McfdError when it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise McfdError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`"
                            " (there is no CPU fallback)")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def _dp(a):
    return None if a is None else a.ctypes.data_as(_D)


def _ip(a):
    return None if a is None else a.ctypes.data_as(_I64)


def _i32(a):
    return None if a is None else a.ctypes.data_as(_I32)


def time_step_static(scheme, degree, n_cells, column_ptr, length_ptr, dt_guess, rk_order, dt_ptr, stream=0):
    """mcfd_time_step_static: the time step from caller-owned device buffers, without any allocation
    (fixed 1 MiB scratch inside the library, at most 1 048 576 cells).  All pointers are device addresses."""
    lib = load()
    rc = lib.mcfd_time_step_static(scheme, degree, n_cells, _P(column_ptr), _P(length_ptr), dt_guess, rk_order,
                                   _P(dt_ptr), _P(stream))
    if rc != 0:
        raise McfdError(f"libminicfd_b200 error {rc}: {lib.mcfd_last_error(None).decode()}")


def peer_step_group(contexts, order, t, dt_guess):
    """mcfd_peer_step_group: the contexts of this process advance one step in lockstep."""
    lib = load()
    arr = (_P * len(contexts))(*[c.h for c in contexts])
    rc = lib.mcfd_peer_step_group(arr, len(contexts), order, t, dt_guess)
    if rc != 0:
        raise McfdError(f"libminicfd_b200 error {rc}: " + "; ".join(lib.mcfd_last_error(c.h).decode() for c in contexts))


class Context:
    """One mcfd_ctx: a part of the mesh resident on one GPU."""

    def __init__(self, scheme=FR_LOBATTO, degree=2, riemann=HLLC, device=-1, fr_c_next=0.0, diffusion=INVISCID,
                 k_local=False):
        self.lib = load()
        self.h = _P()
        cfg = Config(scheme, degree, riemann, device, fr_c_next, diffusion, 1 if k_local else 0)
        rc = self.lib.mcfd_create(C.byref(cfg), C.byref(self.h))
        if rc != 0:
            raise McfdError(f"mcfd_create failed ({rc}): {self.lib.mcfd_last_error(None).decode()}")
        self.degree = degree
        self.q = degree + 1
        self.kfields = 5 * self.q ** 3
        if scheme == DG_GENERAL:       # K x M modal coefficients per cell
            self.kfields = 5 * (10 if degree == 2 else 20)
        self.n_local = 0
        self._cb = None

    def limiter_weno_lazy(self, w0, eps, phase=-1):
        self._check(self.lib.mcfd_limiter_weno_lazy(self.h, w0, eps, phase))

    def limiter_weno_eigen(self, w0, eps, phase=-1):
        self._check(self.lib.mcfd_limiter_weno_eigen(self.h, w0, eps, phase))

    def eigen_matrices(self, conservative, frames):
        u = np.ascontiguousarray(conservative, dtype=np.float64).reshape(-1, 5)
        f = np.ascontiguousarray(frames, dtype=np.float64).reshape(-1, 9)
        left = np.empty((len(u), 5, 5)); right = np.empty((len(u), 5, 5))
        self._check(self.lib.mcfd_eigen_matrices(self.h, len(u), _dp(u), _dp(f), _dp(left), _dp(right)))
        return left, right

    def set_diffusion(self, beta_0, beta_1, nu, prandtl):
        self._check(self.lib.mcfd_set_diffusion(self.h, beta_0, beta_1, nu, prandtl))

    def viscosity_halo_buffers(self):
        sp, rp = _P(), _P()
        sn, rn = C.c_int64(0), C.c_int64(0)
        self._check(self.lib.mcfd_viscosity_halo_buffers(self.h, C.byref(sp), C.byref(sn), C.byref(rp), C.byref(rn)))
        return (sp.value or 0, sn.value), (rp.value or 0, rn.value)

    def viscosity_halo_pack(self):
        self._check(self.lib.mcfd_viscosity_halo_pack(self.h))

    def viscosity_halo_unpack(self):
        self._check(self.lib.mcfd_viscosity_halo_unpack(self.h))

    def build_damping_matrices(self):
        self._check(self.lib.mcfd_build_damping_matrices(self.h))

    def damping_matrices(self):
        n = self.q ** 3
        out = np.empty((self.n_local, n, n))
        self._check(self.lib.mcfd_get_damping_matrices(self.h, _dp(out)))
        return out

    def update_cell_viscosity(self, time_scale=1.0):
        self._check(self.lib.mcfd_update_cell_viscosity(self.h, time_scale))

    def cell_viscosity(self, n_cells=None):
        """[n_local][5] by default; pass n_local + n_ghost to see the ghost cells' values too."""
        out = np.empty((self.n_cells, 5))        # the library always writes local and ghost cells
        self._check(self.lib.mcfd_get_cell_viscosity(self.h, _dp(out)))
        return out[:self.n_local if n_cells is None else n_cells].copy()

    def jacobian_determinants(self):
        out = np.empty((self.n_local, self.q ** 3))
        self._check(self.lib.mcfd_get_jacobian_determinants(self.h, _dp(out)))
        return out

    def sample_vtk(self, modal=False, points=True):
        """mcfd_sample_vtk: (points [n_local][27][3] or None, values [n_local][27][5])"""
        pts = np.empty((self.n_local, 27, 3)) if points else None
        values = np.empty((self.n_local, 27, 5))
        self._check(self.lib.mcfd_sample_vtk(self.h, 1 if modal else 0, _dp(pts) if points else None, _dp(values)))
        return pts, values

    def sample_vtk_on_device(self, modal=False):
        """mcfd_sample_vtk with the values kept on the device (for vtk_component_text)."""
        self._check(self.lib.mcfd_sample_vtk(self.h, 1 if modal else 0, None, None))

    def _text(self, n, call):
        buf = np.empty(16 * max(int(n), 1), dtype=np.uint8)      # not zero-filled: only `length` bytes get touched
        length = C.c_int64(0)
        self._check(call(buf.ctypes.data_as(C.c_char_p), C.c_int64(buf.size), C.byref(length)))
        return buf[:length.value].data                            # a bytes-like view, no copy

    def vtk_component_text(self, component):
        """bytes: `vtu << v << ' '` over component k of the last samples (mcfd_vtk_component_text)"""
        return self._text(self.n_local * 27, lambda b, c, l: self.lib.mcfd_vtk_component_text(self.h, component, b, c, l))

    def format_doubles(self, values):
        v = np.ascontiguousarray(values, dtype=np.float64).reshape(-1)
        return self._text(v.size, lambda b, c, l: self.lib.mcfd_format_doubles(self.h, v.size, _dp(v), b, c, l))

    def format_integers(self, n, first, step):
        return self._text(n, lambda b, c, l: self.lib.mcfd_format_integers(self.h, n, first, step, b, c, l))

    def quadrature_points(self):
        out = np.empty((self.n_local, self.q ** 3, 3))
        self._check(self.lib.mcfd_get_quadrature_points(self.h, _dp(out)))
        return out

    def project(self, values):
        v = np.ascontiguousarray(values, dtype=np.float64)
        assert v.size == self.n_local * self.q ** 3 * 5
        self._check(self.lib.mcfd_project(self.h, _dp(v)))

    def set_cell_viscosity(self, nu):
        nu = np.ascontiguousarray(nu, dtype=np.float64).reshape(-1, 5)
        self._check(self.lib.mcfd_set_cell_viscosity(self.h, len(nu), _dp(nu)))

    def _check(self, rc):
        if rc != 0:
            raise McfdError(f"libminicfd_b200 error {rc}: {self.lib.mcfd_last_error(self.h).decode()}")

    def close(self):
        if self.h:
            self.lib.mcfd_destroy(self.h)
            self.h = _P()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_stream(self, cuda_stream: int):
        self._check(self.lib.mcfd_set_stream(self.h, _P(cuda_stream)))

    def set_mesh(self, nodes, cell_nodes, n_local, n_ghost, bface_nodes, bface_tags, periodic=None, global_ids=None):
        if global_ids is not None:
            ids = np.ascontiguousarray(global_ids, dtype=np.int64)
            self._check(self.lib.mcfd_set_cell_global_ids(self.h, len(ids), _ip(ids)))
        nodes = np.ascontiguousarray(nodes, dtype=np.float64)
        cell_nodes = np.ascontiguousarray(cell_nodes, dtype=np.int64)
        bface_nodes = np.ascontiguousarray(bface_nodes, dtype=np.int64).reshape(-1, 4)
        bface_tags = np.ascontiguousarray(bface_tags, dtype=np.int32)
        assert cell_nodes.shape == (n_local + n_ghost, 8)
        pairs = None if periodic is None else np.ascontiguousarray(periodic, dtype=np.int64)
        self._check(self.lib.mcfd_set_mesh(self.h, len(nodes), _dp(nodes), n_local, n_ghost, _ip(cell_nodes),
                                           len(bface_nodes), _ip(bface_nodes), _i32(bface_tags),
                                           0 if pairs is None else len(pairs), _ip(pairs)))
        self.n_local = n_local
        self.n_cells = n_local + n_ghost

    def faces(self, kind):
        n = self.lib.mcfd_get_faces(self.h, kind, None, None, None, None, None)
        if n < 0:
            self._check(int(n))
        holder = np.empty(n, dtype=np.int64); sharer = np.empty(n, dtype=np.int64)
        hf = np.empty(n, dtype=np.int32); sf = np.empty(n, dtype=np.int32); o = np.empty(n, dtype=np.int32)
        self.lib.mcfd_get_faces(self.h, kind, _ip(holder), _ip(sharer), _i32(hf), _i32(sf), _i32(o))
        return holder, sharer, hf, sf, o

    def cell_geometry(self):
        length = np.empty(self.n_local); volume = np.empty(self.n_local)
        self._check(self.lib.mcfd_get_cell_geometry(self.h, _dp(length), _dp(volume)))
        return length, volume

    def set_halo(self, send_counts, send_cells, recv_counts, recv_cells):
        sc = np.ascontiguousarray(send_counts, dtype=np.int64); scl = np.ascontiguousarray(send_cells, dtype=np.int64)
        rc = np.ascontiguousarray(recv_counts, dtype=np.int64); rcl = np.ascontiguousarray(recv_cells, dtype=np.int64)
        self._check(self.lib.mcfd_set_halo(self.h, len(sc), _ip(sc), _ip(scl), len(rc), _ip(rc), _ip(rcl)))

    def halo_buffers(self):
        sp, rp = _P(), _P()
        sn, rn = C.c_int64(), C.c_int64()
        self._check(self.lib.mcfd_halo_buffers(self.h, C.byref(sp), C.byref(sn), C.byref(rp), C.byref(rn)))
        return (sp.value or 0, sn.value), (rp.value or 0, rn.value)

    # ---- halo through peer-mapped memory (mcfd_peer_*) ----
    def set_time_dependent(self, flag):
        self._check(self.lib.mcfd_set_time_dependent(self.h, 1 if flag else 0))

    def peer_export(self, world):
        """-> (three 64-byte IPC handles, three device pointers): epoch words, time-step slots, receive stage."""
        handles = (C.c_ubyte * (3 * 64))()
        ptrs = (_P * 3)()
        self._check(self.lib.mcfd_peer_export(self.h, world, C.cast(handles, _P), ptrs))
        raw = bytes(handles)
        return [raw[64 * i:64 * (i + 1)] for i in range(3)], [p or 0 for p in ptrs]

    def peer_open(self, handle: bytes) -> int:
        buf = (C.c_ubyte * 64).from_buffer_copy(handle)
        out = _P()
        self._check(self.lib.mcfd_peer_open(self.h, C.cast(buf, _P), C.byref(out)))
        return out.value

    def peer_connect(self, rank, world, neighbour_flags, slot_at_neighbour, neighbour_recv, neighbour_recv_items,
                     offset_at_neighbour, world_flags, world_dt_slots):
        n = len(neighbour_flags)
        fl = (_P * max(1, n))(*neighbour_flags)
        rv = (_P * max(1, n))(*neighbour_recv)
        slots = np.ascontiguousarray(slot_at_neighbour, dtype=np.int32)
        items = np.ascontiguousarray(neighbour_recv_items, dtype=np.int64)
        offs = np.ascontiguousarray(offset_at_neighbour, dtype=np.int64)
        wf = (_P * world)(*world_flags)
        ws = (_P * world)(*world_dt_slots)
        self._check(self.lib.mcfd_peer_connect(self.h, rank, world, n, fl, slots.ctypes.data_as(_I32), rv, _ip(items),
                                               _ip(offs), wf, ws))

    def peer_join(self):
        self._check(self.lib.mcfd_peer_join(self.h))

    def peer_push(self):
        self._check(self.lib.mcfd_peer_push(self.h))

    def peer_wait(self):
        self._check(self.lib.mcfd_peer_wait(self.h))

    def peer_min_time_step(self, dt_ptr=0):
        self._check(self.lib.mcfd_peer_min_time_step(self.h, _P(dt_ptr)))

    def peer_step(self, order, t, dt_guess):
        self._check(self.lib.mcfd_peer_step(self.h, order, t, dt_guess))

    def peer_check(self):
        self._check(self.lib.mcfd_peer_check(self.h))

    def halo_pack(self):
        self._check(self.lib.mcfd_halo_pack(self.h))

    def halo_unpack(self):
        self._check(self.lib.mcfd_halo_unpack(self.h))

    def set_boundary(self, tag, kind):
        self._check(self.lib.mcfd_set_boundary(self.h, tag, kind))

    def boundary_points(self, tag):
        n = self.lib.mcfd_count_boundary_faces(self.h, tag)
        if n < 0:
            self._check(int(n))
        xyz = np.empty((n, self.q * self.q, 3))
        if n:
            self._check(self.lib.mcfd_get_boundary_points(self.h, tag, _dp(xyz)))
        return xyz

    def set_boundary_values(self, tag, values):
        values = np.ascontiguousarray(values, dtype=np.float64)
        if values.size:
            self._check(self.lib.mcfd_set_boundary_values(self.h, tag, _dp(values)))

    def set_time_callback(self, fn):
        """fn(t) is called before every residual evaluation."""
        if fn is None:
            self._cb = C.cast(None, TIME_CALLBACK)
        else:
            self._cb = TIME_CALLBACK(lambda user, t: fn(t))
        self._check(self.lib.mcfd_set_time_callback(self.h, self._cb, None))

    def set_solution(self, column):
        column = np.ascontiguousarray(column, dtype=np.float64)
        assert column.size == self.n_local * self.kfields
        self._check(self.lib.mcfd_set_solution(self.h, _dp(column)))

    def get_solution(self, out=None):
        out = np.empty(self.n_local * self.kfields) if out is None else out
        self._check(self.lib.mcfd_get_solution(self.h, _dp(out)))
        return out

    def set_solution_device(self, ptr: int):
        self._check(self.lib.mcfd_set_solution_device(self.h, _P(ptr)))

    def get_solution_device(self, ptr: int):
        self._check(self.lib.mcfd_get_solution_device(self.h, _P(ptr)))

    def residual(self, t=0.0, out=None):
        out = np.empty(self.n_local * self.kfields) if out is None else out
        self._check(self.lib.mcfd_residual(self.h, t, _dp(out), None))
        return out

    def rk_step(self, order, t, dt):
        self._check(self.lib.mcfd_rk_step(self.h, order, t, dt))

    def rk_begin(self, order, t, dt):
        self._check(self.lib.mcfd_rk_begin(self.h, order, t, dt))

    def rk_stage_inner(self, stage):
        self._check(self.lib.mcfd_rk_stage_inner(self.h, stage))

    def rk_stage_inter(self, stage):
        self._check(self.lib.mcfd_rk_stage_inter(self.h, stage))

    def time_step(self, dt_guess, rk_order):
        dt = C.c_double()
        self._check(self.lib.mcfd_time_step(self.h, dt_guess, rk_order, C.byref(dt)))
        return dt.value

    def time_step_device(self, dt_guess, rk_order, dt_ptr: int):
        self._check(self.lib.mcfd_time_step_device(self.h, dt_guess, rk_order, _P(dt_ptr)))

    def rk_begin_device(self, order, t, dt_ptr: int):
        self._check(self.lib.mcfd_rk_begin_device(self.h, order, t, _P(dt_ptr)))

    def rk_step_device(self, order, t, dt_ptr: int):
        self._check(self.lib.mcfd_rk_step_device(self.h, order, t, _P(dt_ptr)))

    def limiter_reconstruct(self, kind):
        self._check(self.lib.mcfd_limiter_reconstruct(self.h, kind))

    def launch_count(self):
        return int(self.lib.mcfd_launch_count(self.h))

    def stage_kernel(self):
        """Which kernel a stage of this context launches (mcfd_stage_kernel)."""
        return self.lib.mcfd_stage_kernel(self.h).decode()

    def enable_kernel_timing(self, on=True):
        self._check(self.lib.mcfd_enable_kernel_timing(self.h, int(on)))

    def kernel_time(self, reset=True):
        ms, n = C.c_double(), C.c_int64()
        self._check(self.lib.mcfd_kernel_time(self.h, C.byref(ms), C.byref(n), int(reset)))
        return ms.value, n.value

    def riemann_flux(self, normals, ul, ur, bc_kind=-1):
        normals = np.ascontiguousarray(normals, dtype=np.float64).reshape(-1, 3)
        ul = np.ascontiguousarray(ul, dtype=np.float64).reshape(-1, 5)
        ur = np.ascontiguousarray(ur, dtype=np.float64).reshape(-1, 5)
        out = np.empty_like(ul)
        self._check(self.lib.mcfd_riemann_flux(self.h, bc_kind, len(ul), _dp(normals), _dp(ul), _dp(ur), _dp(out)))
        return out
