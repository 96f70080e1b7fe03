"""Analytic fields used as initial / boundary states (vectorised numpy; host side).

`smooth` and `vortex` evaluate the same formulas as the test harness's built-in fields so that the
GPU path and the reference can be fed identical boundary functions.
"""
from __future__ import annotations

import numpy as np

GAMMA = 1.4


def primitive_to_conservative(rho, u, v, w, p):
    """IdealGas::PrimitiveToConservative (include/mini/riemann/euler/types.hpp:302-311)."""
    e = p / (GAMMA - 1) + 0.5 * rho * (u * u + v * v + w * w)
    return np.stack([rho, rho * u, rho * v, rho * w, e], axis=-1)


def constant(state):
    state = np.asarray(state, dtype=np.float64)

    def f(xyz, t):
        return np.broadcast_to(state, xyz.shape[:-1] + (5,)).copy()
    return f


def smooth(a=1.0):
    """A smooth, strictly admissible test state varying in every direction."""
    def f(xyz, t):
        x, y, z = xyz[..., 0], xyz[..., 1], xyz[..., 2]
        rho = 1.0 + 0.2 * np.sin(a * (x + 2 * y - z) + 0.3 * t)
        u = 0.5 + 0.3 * np.cos(a * (x - y) + 0.1)
        v = -0.2 + 0.25 * np.sin(a * (y + z) + 0.7)
        w = 0.1 + 0.2 * np.cos(a * (z - 2 * x) - 0.4)
        p = 1.0 + 0.3 * np.sin(a * (x + y + z) + 1.1)
        e = p / 0.4 + 0.5 * rho * (u * u + v * v + w * w)
        return np.stack([rho, rho * u, rho * v, rho * w, e], axis=-1)
    return f


def vortex(x0=5.0, y0=5.0, u_inf=1.0, v_inf=1.0, w_inf=0.0, beta=5.0):
    """Isentropic vortex of BASELINE.json config 2, exact in time."""
    def f(xyz, t):
        x, y = xyz[..., 0], xyz[..., 1]
        dx = x - x0 - u_inf * t
        dy = y - y0 - v_inf * t
        r2 = dx * dx + dy * dy
        e = np.exp(0.5 * (1 - r2))
        du = -beta / (2 * np.pi) * e * dy
        dv = beta / (2 * np.pi) * e * dx
        dT = -(GAMMA - 1) * beta * beta / (8 * GAMMA * np.pi * np.pi) * e * e
        T = 1 + dT
        rho = T ** (1 / (GAMMA - 1))
        p = rho ** GAMMA
        u = u_inf + du
        v = v_inf + dv
        w = np.full_like(u, w_inf)
        en = p / (GAMMA - 1) + 0.5 * rho * (u * u + v * v + w * w)
        return np.stack([rho, rho * u, rho * v, rho * w, en], axis=-1)
    return f
