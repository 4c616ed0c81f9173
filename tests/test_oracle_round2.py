"""CPU: the oracle pieces added in round 2 -- quantiser alone, particle-ID and
unsigned-int fields, the segment image with per-block checksums -- on both
flavours (the restated port and, where built, the reference's own objects)."""
import numpy as np
import pytest

import _oracle
from _data import zeldovich_host

FLAVOURS = _oracle.flavours()


@pytest.fixture(params=FLAVOURS, ids=[o.flavour for o in FLAVOURS])
def orc(request):
    return request.param


def grid_ids(ngrid, start, count, rng=None, jitter=0):
    ids = np.arange(start, start + count, dtype=np.uint64)
    if rng is not None and jitter:
        ids = ids[rng.permutation(count)[:count]]
    return ids


def test_dequantize_is_the_tail_of_decode(orc):
    pos, _ = zeldovich_host(16)
    payload, q = orc.encode_field(_oracle.FIELD_POSN, pos.reshape(-1), 1e-3, 100.0, seed=5, want_q=True)
    hd = _oracle.parse_header(payload)
    a = orc.decode_field(payload)
    b = orc.dequantize_field(hd, q.astype(np.uint64))
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))


def test_id_quantizer_matches_between_flavours():
    """a slab of a 64^3 grid that wraps in x and crosses the box edge in z"""
    w = 64
    ix, iy, iz = np.meshgrid(np.arange(50, 50 + 30) % w, np.arange(10, 20), np.arange(60, 60 + 8) % w, indexing="ij")
    ids = (ix + w * iy + w * w * iz).astype(np.uint64).T.reshape(-1)
    res = [o.id_quantize(ids, w) for o in FLAVOURS]
    for x0, x1, q in res[1:]:
        assert np.array_equal(x0, res[0][0]) and np.array_equal(x1, res[0][1]) and np.array_equal(q, res[0][2])
    x0, x1, q = res[0]
    assert q.max() < 30 and (x1 - x0).tolist() == [29, 9, 7]
    for o in FLAVOURS:
        assert np.array_equal(o.id_dequantize(q, w, x0), ids)


def test_id_field_roundtrip(orc):
    rng = np.random.default_rng(3)
    w = 128
    ids = np.arange(5 * w * w + 17, 5 * w * w + 17 + 40000, dtype=np.uint64)
    payload = orc.encode_id_field(ids, w)
    assert np.array_equal(orc.decode_id_field(payload), ids)
    assert payload.size < 0.2 * ids.nbytes
    ids2 = ids[rng.permutation(ids.size)]
    assert np.array_equal(orc.decode_id_field(orc.encode_id_field(ids2, w)), ids2)
    assert np.array_equal(_oracle.port().encode_id_field(ids2, w), orc.encode_id_field(ids2, w))


def test_uint_field_roundtrip(orc):
    rng = np.random.default_rng(4)
    for n in (1, 2, 3, 1000, 3073, 10000):
        data = (np.uint64(1) << np.uint64(40)) + rng.integers(0, 1 << 20, n).astype(np.uint64)
        payload = orc.encode_uint_field(data)
        assert np.array_equal(orc.decode_uint_field(payload, n), data)
        assert np.array_equal(_oracle.port().encode_uint_field(data), payload)
    with pytest.raises(ValueError):
        orc.encode_uint_field(np.array([0, 1 << 30], np.uint64))  # range beyond 24 bits


def test_segment_image_roundtrip_and_block_damage(orc):
    pos, vel = zeldovich_host(16)
    n = pos.shape[1]
    ids = np.arange(n, dtype=np.uint64)
    fields = [(_oracle.FIELD_POSN, n, orc.encode_field(_oracle.FIELD_POSN, pos.reshape(-1), 1e-3, 100.0, seed=1)),
              (_oracle.FIELD_VELC, n, orc.encode_field(_oracle.FIELD_VELC, vel.reshape(-1), 0.5, 0.0, seed=2)),
              (_oracle.FIELD_PTID, n, orc.encode_id_field(ids, 16))]
    image = orc.seg_to_bytes(fields, n)
    nb = (n + 1023) // 1024
    assert image.size == 16 + 16 * 3 + 8 * 3 * (1 + 3 * nb) + sum(f[2].size for f in fields)
    assert image.size % 8 == 0
    hdr = image[:16].view(np.int32)
    assert hdr[1] == 3 * (1 + 3 * nb) and hdr[2] == 3 and hdr[3] == n
    assert int(image[:4].view(np.uint32)[0]) == orc.checksum(image[4:16 + 48 + 8 * hdr[1]])
    back, pn = orc.seg_from_bytes(image)
    assert pn == n and len(back) == 3
    for (code, m, payload, valid), want in zip(back, fields):
        assert valid and code == want[0] and m == n and np.array_equal(payload, want[2])
    # one damaged byte inside the second packed block of the velocity field
    bad = image.copy()
    hdr_bytes = 16 + 48 + 8 * hdr[1]
    vel_start = hdr_bytes + fields[0][2].size
    tab = _oracle.parse_table(fields[1][2])
    pos_in_field = 64 + 48 * nb + int(tab["Offset16"][1]) * 16 + 5
    bad[vel_start + pos_in_field] ^= 0x10
    back, _ = orc.seg_from_bytes(bad)
    assert [b[3] for b in back] == [True, True, True]
    assert np.array_equal(back[0][2], fields[0][2]) and np.array_equal(back[2][2], fields[2][2])
    t2 = _oracle.parse_table(back[1][2])
    assert t2["Width"][1] == _oracle.WIDTH_DAMAGED and (np.delete(t2["Width"], 1) == np.delete(tab["Width"], 1)).all()
    good = orc.decode_field(fields[1][2])
    hurt = orc.decode_field(back[1][2])
    lost = np.isnan(hurt)
    assert lost.sum() == 1024 and lost[1024:2048].all()
    assert np.array_equal(good[~lost].view(np.uint32), hurt[~lost].view(np.uint32))
    # a damaged field header invalidates that field only; a damaged segment header everything
    bad = image.copy()
    bad[vel_start + 20] ^= 1
    back, _ = orc.seg_from_bytes(bad)
    assert [b[3] for b in back] == [True, False, True]
    bad = image.copy()
    bad[9] ^= 1
    with pytest.raises(ValueError):
        orc.seg_from_bytes(bad)
