"""Seeded synthetic initial conditions: Plummer sphere in Hénon units.

The reference ships no inputs (its ``input/`` submodule is missing), so every config
in BASELINE.json runs on models generated here.  Aarseth–Hénon–Wielen (1974) sampling: radii from
the inverted cumulative mass profile, speeds by von Neumann rejection on g(q) = q²(1−q²)^{7/2},
scaled to Hénon units (G = M = 1, E = −1/4, length scale 3π/16), moved to the centre-of-mass frame.

Masses are rounded to float32 because the reference reader stores them as float
(reference: src/include/NbodySystem.cpp:174, src/include/common.hpp:216,326): every mass the
integrator ever sees is a float-representable double.
"""
from __future__ import annotations

import numpy as np

R_CUT = 22.8042468  # model units; standard cut used by Aarseth's generator


def salpeter_masses(n: int, rng, m_lo=0.1, m_hi=50.0, alpha=2.35):
    """dN/dm ∝ m^-alpha on [m_lo, m_hi], renormalised to Σm = 1."""
    u = rng.random(n)
    a = 1.0 - alpha
    m = (m_lo**a + u * (m_hi**a - m_lo**a)) ** (1.0 / a)
    return m / m.sum()


def plummer(n: int, seed: int = 42, mass_spectrum: str = "equal"):
    """Return (m[n], r[n,3], v[n,3]) float64 arrays; m is float32-representable."""
    rng = np.random.default_rng(seed)
    if mass_spectrum == "equal":
        m = np.full(n, 1.0 / n)
    elif mass_spectrum == "salpeter":
        m = salpeter_masses(n, rng)
    else:
        raise ValueError("mass_spectrum must be 'equal' or 'salpeter'")
    m = m.astype(np.float32).astype(np.float64)

    # radii
    rad = np.empty(n)
    todo = np.arange(n)
    while todo.size:
        x1 = rng.random(todo.size)
        x1 = np.clip(x1, 1e-12, 1.0)
        rr = 1.0 / np.sqrt(x1 ** (-2.0 / 3.0) - 1.0 + 1e-300)
        ok = rr < R_CUT
        rad[todo[ok]] = rr[ok]
        todo = todo[~ok]

    def iso(mag):
        z = (1.0 - 2.0 * rng.random(n)) * mag
        phi = 2.0 * np.pi * rng.random(n)
        s = np.sqrt(np.maximum(mag * mag - z * z, 0.0))
        return np.stack([s * np.cos(phi), s * np.sin(phi), z], axis=1)

    r = iso(rad)

    # speeds
    q = np.empty(n)
    todo = np.arange(n)
    while todo.size:
        x4 = rng.random(todo.size)
        x5 = rng.random(todo.size)
        ok = 0.1 * x5 < x4 * x4 * (1.0 - x4 * x4) ** 3.5
        q[todo[ok]] = x4[ok]
        todo = todo[~ok]
    vel = q * np.sqrt(2.0) * (1.0 + rad * rad) ** (-0.25)
    v = iso(vel)

    # Hénon units
    sf = 3.0 * np.pi / 16.0
    r *= sf
    v /= np.sqrt(sf)

    # centre-of-mass frame
    mt = m.sum()
    r -= (m[:, None] * r).sum(axis=0) / mt
    v -= (m[:, None] * v).sum(axis=0) / mt
    return m, np.ascontiguousarray(r), np.ascontiguousarray(v)


def default_seed(n: int) -> int:
    """Seed convention of SURVEY.md §8d: 42 + log2 N."""
    return 42 + int(round(np.log2(max(n, 1))))
