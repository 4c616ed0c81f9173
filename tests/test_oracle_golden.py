"""The oracle's primitives against the reference's own golden tables
(/root/reference/test/compress_util_test.c) and against the reference's
compiled objects (oracle/_ref). CPU only."""
import numpy as np
import pytest

import _oracle

FLAVOURS = _oracle.flavours()
IDS = [o.flavour for o in FLAVOURS]

# test/compress_util_test.c:446-460  (unpacked, packed, width)
PACK_ROWS = [
    ([], [], 32),
    ([0x12345678], [], 0),
    ([0x12345678], [0x8], 4),
    ([0x12345678], [0x78], 8),
    ([0x12345678], [0x678], 12),
    ([0x12345678], [0x12345678], 32),
    ([0xaa, 0xbb], [0xbbaa], 8),
    ([0xaa, 0xbb], [0xbb00aa], 16),
    ([0xaa, 0xbb], [0xbb0aa], 12),
    ([3, 2, 3], [5], 1),
    ([0xaaaaaaaa, 0xbbbbbbbb], [0xaaaaaaaa, 0xbbbbbbbb], 32),
    ([0x87654321, 0x87654321], [0x17654321, 0x765432], 28),
    ([0xaaaaaa, 0xbbbbbb, 0xcccccc, 0xdddddd], [0xbbaaaaaa, 0xccccbbbb, 0xddddddcc], 24),
]

# test/compress_util_test.c:352-362  (x, level, idx, x0, dx)
BIN_ROWS = [
    ([], 0, [], 0, 1),
    ([0], 0, [0], 0, 1),
    ([0.25], 1, [0], 0, 1),
    ([0.75], 1, [1], 0, 1),
    ([0.5, 1.5, 3.5], 2, [0, 1, 3], 0, 4),
    ([3.5, 1.5, 0.5], 2, [3, 1, 0], 0, 4),
    ([3, 5, 7, 9], 2, [0, 1, 2, 3], 2, 8),
    ([3, 5, 7, 9], 3, [1, 3, 5, 7], 2, 8),
    ([-0.5, -1.5, -3.5], 2, [0, 1, 3], 0, -4),
]


@pytest.mark.parametrize("orc", FLAVOURS, ids=IDS)
def test_pack_golden(orc):
    for unpacked, packed, width in PACK_ROWS:
        got = orc.pack(np.array(unpacked, np.uint32), width)
        assert got.tolist() == packed, (unpacked, width)
        mask = 0xFFFFFFFF if width == 32 else (1 << width) - 1
        back = orc.unpack(np.array(packed, np.uint32), width, len(unpacked))
        assert back.tolist() == [v & mask for v in unpacked]


@pytest.mark.parametrize("orc", FLAVOURS, ids=IDS)
def test_pack_roundtrip_every_width(orc):
    # test/compress_util_test.c:517-550, with lengths that actually straddle
    rng = np.random.default_rng(0)
    for width in range(0, 33):
        for n in (0, 1, 2, 31, 32, 33, 1000, 1024):
            x = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
            mask = 0xFFFFFFFF if width == 32 else (1 << width) - 1
            words = orc.pack(x, width)
            assert words.size == (width * n + 31) // 32
            assert np.array_equal(orc.unpack(words, width, n), x & np.uint32(mask))


@pytest.mark.parametrize("orc", FLAVOURS, ids=IDS)
def test_uniform_bin_index_golden(orc):
    for x, level, idx, x0, dx in BIN_ROWS:
        got = orc.uniform_bin_index(np.array(x, np.float32), level, x0, dx)
        assert got.tolist() == idx


@pytest.mark.parametrize("orc", FLAVOURS, ids=IDS)
def test_undo_uniform_bin_index_requantizes(orc):
    # test/compress_util_test.c:398-425: dithered value falls back in its bin
    state = orc.rand_seed(0, 1).reshape(-1).copy()
    for idx, level, x0, dx in [([0], 0, 0, 16), (list(range(8)), 4, -4, 8)]:
        x = orc.undo_uniform_bin_index(np.array(idx, np.uint32), level, x0, dx, state)
        assert orc.uniform_bin_index(x, level, x0, dx).tolist() == idx


@pytest.mark.parametrize("orc", FLAVOURS, ids=IDS)
def test_rand_known_answers(orc):
    # rand.c has no tests of its own (src/rand.c:6-8). These values were
    # printed by the reference's rand.c compiled here (oracle/_ref); the state
    # words are splitmix64's published first two outputs for seed 0.
    st = orc.rand_seed(0, 1).reshape(-1).copy()
    assert st.tolist() == [0xE220A8397B1DCDAF, 0x6E789E6AA1B965F4]
    assert orc.lib.orc_rand_u64(st) == 0x509946A41CD733A3
    assert orc.lib.orc_rand_u64(st) == 0x00885667B1934BFA
    assert orc.lib.orc_rand_float(st) == 0.5618565678596497


@pytest.mark.parametrize("orc", FLAVOURS, ids=IDS)
def test_checksum_known_answer(orc):
    assert orc.checksum(np.array([1, 2, 3], np.uint8)) == 0x60000004
    assert orc.checksum(np.zeros(0, np.uint8)) == 1  # src/util.c:439


@pytest.mark.parametrize("orc", FLAVOURS, ids=IDS)
def test_depth_rows(orc):
    # depth column of test/old_tests/algo_test.c:117-122
    assert orc.delta_to_depth(2.0, 3.0, 3.0) == 0
    assert orc.delta_to_depth(2.0, 3.0, 5.0) == 1
    assert orc.delta_to_depth(0.5, 1.0, 8.0) == 4
    assert orc.delta_to_depth(1e-9, 0.0, 100.0) == -1  # reference Panics


@pytest.mark.parametrize("orc", FLAVOURS, ids=IDS)
def test_fast_uniform_compress_property(orc):
    # test/compress_util_test.c:610-659: whole-path round trip, error <= L/2^level mod L
    n, level, L = 100000, 15, np.float32(10)
    st = orc.rand_seed(0, 1).reshape(-1).copy()
    x = np.array([orc.lib.orc_rand_float(st) for _ in range(2000)], np.float32)
    x = np.resize(x, n) + np.float32(1.5)
    x = orc.periodic(x, L)
    u = orc.undo_periodic(x, L)
    lo, hi = orc.minmax(u)
    q = orc.uniform_bin_index(u, level, lo, hi - lo)
    words = orc.pack(q, level)
    q2 = orc.unpack(words, level, n)
    assert np.array_equal(q, q2)
    y = orc.periodic(orc.undo_uniform_bin_index(q2, level, lo, hi - lo, st), L)
    err = np.abs(x.astype(np.float64) - y)
    err = np.minimum(err, float(L) - err)
    assert err.max() <= float(L) / (1 << level) * 1.001


@pytest.mark.skipif(_oracle.reference() is None, reason="oracle/_ref not built")
def test_port_matches_reference_primitives():
    port, ref = _oracle.port(), _oracle.reference()
    rng = np.random.default_rng(7)
    for trial in range(20):
        n = int(rng.integers(1, 5000))
        L = np.float32(rng.uniform(1, 200))
        x = rng.uniform(0, L, n).astype(np.float32)
        assert np.array_equal(port.undo_periodic(x, L), ref.undo_periodic(x, L))
        y = (x + np.float32(rng.uniform(-L, L))).astype(np.float32)
        assert np.array_equal(port.periodic(y, L), ref.periodic(y, L))
        assert port.minmax(x) == ref.minmax(x)
        level = int(rng.integers(0, 25))
        lo, hi = port.minmax(x)
        qa = port.uniform_bin_index(x, level, lo, hi - lo)
        qb = ref.uniform_bin_index(x, level, lo, hi - lo)
        assert np.array_equal(qa, qb)
        sa = port.rand_seed(trial, 3).copy()
        sb = ref.rand_seed(trial, 3).copy()
        assert np.array_equal(sa, sb)
        fa = port.undo_uniform_bin_index(qa, level, lo, hi - lo, sa[2])
        fb = ref.undo_uniform_bin_index(qa, level, lo, hi - lo, sb[2])
        assert np.array_equal(fa.view(np.uint32), fb.view(np.uint32))
        assert np.array_equal(sa, sb)
        w = int(rng.integers(0, 33))
        v = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
        assert np.array_equal(port.pack(v, w), ref.pack(v, w))
        b = rng.integers(0, 256, n, dtype=np.uint64).astype(np.uint8)
        assert port.checksum(b) == ref.checksum(b)
        d = np.float32(10.0 ** rng.uniform(-6, 1))
        assert port.delta_to_depth(d, lo, hi) == ref.delta_to_depth(d, lo, hi)
