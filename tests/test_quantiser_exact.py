"""The encoder's division-free quantiser against the
reference's arithmetic, util_UniformBinIndex (/root/reference/src/util.c:184-191):
t = RN_f(a/dx); q = 0 if t < 0, top if t >= 1, else trunc(t * 2^depth).

CPU only (numpy float32/float64 are IEEE, round to nearest even): the identity is
what the CUDA kernel relies on; the GPU parity tests check the kernel itself.
"""
import numpy as np
import pytest


def reference_index(a, dx, depth):
    t = (a / dx).astype(np.float32)                      # IEEE float division
    top = np.uint32((1 << depth) - 1)
    q = np.trunc(t * np.float32(1 << depth)).astype(np.int64)
    q = np.where(t < 0, 0, np.where(t >= 1, top, q))
    return q.astype(np.uint32)


def fast_index(a, dx, depth):
    scale = np.float64(1 << depth) / np.float64(dx)      # per field, once
    s = (a.astype(np.float64) * scale).astype(np.float32)
    top = np.float32((1 << depth) - 1)
    s = np.minimum(s, top)                                # the t >= 1 clamp
    return np.trunc(s).astype(np.uint32)


def edge_values(dx, depth, rng, count):
    """Values within a few ulps of bin edges k*dx/2^depth, plus the ends."""
    k = rng.integers(0, (1 << depth) + 1, size=count)
    a = (k.astype(np.float64) * np.float64(dx) / np.float64(1 << depth)).astype(np.float32)
    steps = rng.integers(-3, 4, size=count)
    for _ in range(3):
        up = steps > 0
        dn = steps < 0
        a = np.where(up, np.nextafter(a, np.float32(np.inf)), a)
        a = np.where(dn, np.nextafter(a, np.float32(-np.inf)), a)
        steps = steps - np.sign(steps)
    a = np.clip(a, np.float32(0), np.float32(dx))
    return np.concatenate([a, np.array([0.0, dx], np.float32)])


@pytest.mark.parametrize("depth", list(range(1, 25)))
def test_division_free_index_matches_reference(depth):
    rng = np.random.default_rng(1000 + depth)
    for trial in range(6):
        dx = np.float32(rng.uniform(1e-3, 1e4)) if trial else np.float32(100.0)
        a = np.concatenate([edge_values(dx, depth, rng, 40000),
                            rng.uniform(0, dx, 20000).astype(np.float32)])
        np.testing.assert_array_equal(fast_index(a, dx, depth), reference_index(a, dx, depth))


@pytest.mark.parametrize("depth", [1, 7, 17, 23])
def test_round_toward_zero_add_truncates(depth):
    """For 0 <= s < 2^23, RZ(s + 2^23) has the bit pattern 0x4B000000 + floor(s)."""
    rng = np.random.default_rng(depth)
    s = rng.uniform(0, (1 << depth) - 1, 100000).astype(np.float32)
    # float64 holds s + 2^23 exactly; rounding toward zero to 24 bits = floor on this range
    exact = s.astype(np.float64) + 8388608.0
    rz = np.floor(exact).astype(np.float32)
    bits = rz.view(np.uint32)
    np.testing.assert_array_equal(bits - np.uint32(0x4B000000), np.floor(s).astype(np.uint32))
