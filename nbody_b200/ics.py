"""Synthetic initial conditions (fp32 SoA), seeded and reproducible.

The reference generates ICs on the host with libstdc++ <random> (src/cuda/particle_init.cu:290-357):
UNIFORM box +-10, SPHERICAL ball R=10, DISK R=10; mass == 1, v == 0 (disk: rotation).  The
distribution *shapes* are restated here with numpy's PCG64 stream (the libstdc++ stream is
implementation-defined); they enter the engine exactly like the reference's
own ICs do, through ParticleSystem.set_state / ParticleData upload.

The reference has no Plummer model; BASELINE.json asks for one, so it lives here
(Aarseth, Henon & Wielen 1974 sampling, Henon units: M = 1, G = 1, E = -1/4).
"""
from __future__ import annotations

import numpy as np

FIELDS = ["pos_x", "pos_y", "pos_z", "vel_x", "vel_y", "vel_z", "acc_x", "acc_y", "acc_z",
          "acc_old_x", "acc_old_y", "acc_old_z", "mass"]


def empty_state(n: int) -> dict:
    """13 zeroed fp32 arrays in include/nbody/types.hpp:234-262 order."""
    return {k: np.zeros(n, np.float32) for k in FIELDS}


def _state(pos, vel, mass) -> dict:
    n = mass.size
    s = empty_state(n)
    for a, k in enumerate("xyz"):
        s["pos_" + k] = np.ascontiguousarray(pos[:, a], dtype=np.float32)
        s["vel_" + k] = np.ascontiguousarray(vel[:, a], dtype=np.float32)
    s["mass"] = np.ascontiguousarray(mass, dtype=np.float32)
    return s


def plummer(n: int, seed: int = 42, rmax_over_a: float = 10.0) -> dict:
    """Plummer sphere in Henon units (scale a = 3*pi/16, total mass 1, m = 1/N, virial velocities)."""
    rng = np.random.default_rng(seed)
    a = 3.0 * np.pi / 16.0
    # radii: M(r)/M = u -> r = a / sqrt(u^(-2/3) - 1), clipped at rmax
    r = np.empty(n)
    filled = 0
    while filled < n:
        u = rng.random(n - filled)
        u = np.clip(u, 1e-12, 1.0 - 1e-12)
        rr = a / np.sqrt(u ** (-2.0 / 3.0) - 1.0)
        rr = rr[rr <= rmax_over_a * a]
        r[filled:filled + rr.size] = rr
        filled += rr.size
    pos = _iso(rng, n) * r[:, None]
    # speeds: q = v/v_esc from g(q) = q^2 (1-q^2)^(7/2) by rejection
    q = np.empty(n)
    filled = 0
    while filled < n:
        m = 2 * (n - filled) + 16
        x = rng.random(m)
        y = rng.random(m) * 0.1
        ok = x[y < x * x * (1.0 - x * x) ** 3.5]
        take = min(ok.size, n - filled)
        q[filled:filled + take] = ok[:take]
        filled += take
    vesc = np.sqrt(2.0) * (r * r + a * a) ** -0.25  # G = M = 1
    vel = _iso(rng, n) * (q * vesc)[:, None]
    mass = np.full(n, 1.0 / n)
    pos -= pos.mean(axis=0)
    vel -= vel.mean(axis=0)
    return _state(pos, vel, mass)


def _iso(rng, n):
    z = 1.0 - 2.0 * rng.random(n)
    phi = 2.0 * np.pi * rng.random(n)
    s = np.sqrt(np.maximum(0.0, 1.0 - z * z))
    return np.stack([s * np.cos(phi), s * np.sin(phi), z], axis=1)


def uniform_sphere(n: int, radius: float = 10.0, mass: float = 1.0, seed: int = 42) -> dict:
    """Shape of ParticleInitializer::initSpherical (src/cuda/particle_init.cu:311-331):
    r = cbrt(u) R, theta = 2 pi u, phi = acos(2u - 1); v = 0."""
    rng = np.random.default_rng(seed)
    r = np.cbrt(rng.random(n)) * radius
    theta = rng.random(n) * 2.0 * np.pi
    phi = np.arccos(2.0 * rng.random(n) - 1.0)
    pos = np.stack([r * np.sin(phi) * np.cos(theta), r * np.sin(phi) * np.sin(theta), r * np.cos(phi)], axis=1)
    return _state(pos, np.zeros((n, 3)), np.full(n, mass))


def uniform_box(n: int, lo: float = -10.0, hi: float = 10.0, mass: float = 1.0, seed: int = 42) -> dict:
    """Shape of ParticleInitializer::initUniform (src/cuda/particle_init.cu:290-309)."""
    rng = np.random.default_rng(seed)
    pos = lo + (hi - lo) * rng.random((n, 3))
    return _state(pos, np.zeros((n, 3)), np.full(n, mass))


def disk(n: int, radius: float = 10.0, thickness: float = 1.0, rotation_speed: float = 0.5,
         mass: float = 1.0, seed: int = 42) -> dict:
    """Shape of ParticleInitializer::initDisk (src/cuda/particle_init.cu:333-357)."""
    rng = np.random.default_rng(seed)
    r = np.sqrt(rng.random(n)) * radius
    theta = rng.random(n) * 2.0 * np.pi
    z = (rng.random(n) - 0.5) * thickness
    v = rotation_speed * np.sqrt(r)
    pos = np.stack([r * np.cos(theta), r * np.sin(theta), z], axis=1)
    vel = np.stack([-v * np.sin(theta), v * np.cos(theta), np.zeros(n)], axis=1)
    return _state(pos, vel, np.full(n, mass))
