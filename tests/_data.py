"""Seeded synthetic inputs shared by CPU and GPU tests (host side, numpy)."""
import numpy as np


def zeldovich_host(ngrid, L=100.0, seed=1234, start=0, count=None, modes=16, rms_cells=1.5):
    """Small host twin of the device generator's *shape* (not bit-identical to
    it): Lagrangian grid id = ix + N*iy + N^2*iz, sum-of-sines displacement,
    positions wrapped into [0, L), velocities = 100*psi. Planar SoA float32.
    Returns (pos[3,count], vel[3,count])."""
    total = ngrid ** 3
    count = total - start if count is None else count
    ids = np.arange(start, start + count, dtype=np.int64)
    ix, iy, iz = ids % ngrid, (ids // ngrid) % ngrid, ids // (ngrid * ngrid)
    q = np.stack([(ix + 0.5), (iy + 0.5), (iz + 0.5)]).astype(np.float64) * (L / ngrid)
    rng = np.random.default_rng(seed)
    k = rng.integers(-4, 5, size=(modes, 3))
    k[(k == 0).all(axis=1)] = (1, 0, 0)
    phase = rng.uniform(0, 2 * np.pi, modes)
    k2 = (k * k).sum(axis=1).astype(np.float64)
    amp = 1.0 / k2
    psi = np.zeros((3, count))
    for m in range(modes):
        arg = 2 * np.pi * (k[m][:, None] * q).sum(axis=0) / L + phase[m]
        psi += (amp[m] * k[m] / np.sqrt(k2[m]))[:, None] * np.sin(arg)
    norm = np.sqrt((amp ** 2).sum() / 2.0)
    psi *= rms_cells * (L / ngrid) / norm
    pos = np.mod(q + psi, L).astype(np.float32)
    pos[pos >= np.float32(L)] = 0.0
    vel = (100.0 * psi).astype(np.float32)
    return np.ascontiguousarray(pos), np.ascontiguousarray(vel)


def uniform_component(n, seed=0, L=1.0):
    """BASELINE config 0: uniform floats, shifted by 1.5 and wrapped into [0,L)
    (test/compress_util_test.c:624-627), used as all three components."""
    rng = np.random.default_rng(seed)
    x = rng.random(3 * n, dtype=np.float32) * np.float32(L)
    x = np.mod(x + np.float32(1.5), np.float32(L)).astype(np.float32)
    x[x >= np.float32(L)] = 0.0
    return x
