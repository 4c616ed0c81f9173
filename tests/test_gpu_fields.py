"""GPU: the callers either side of the hot path -- the quantiser
alone, particle-ID and unsigned-int fields, per-block checksums and the segment
image -- against the CPU oracle, byte for byte / bit for bit."""
import ctypes as C

import numpy as np
import pytest

import _oracle
from _data import zeldovich_host

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
import minnow_c_b200 as mc  # noqa: E402

FLAVOURS = _oracle.flavours()


@pytest.fixture(scope="module")
def codec():
    c = mc.Codec(0)
    yield c
    c.close()


@pytest.fixture(params=FLAVOURS, ids=[o.flavour for o in FLAVOURS])
def orc(request):
    return request.param


def float_cases():
    pos, vel = zeldovich_host(32)
    yield "pos", mc.FIELD_POSN, pos.reshape(-1), 1e-3, 100.0
    yield "vel", mc.FIELD_VELC, vel.reshape(-1), 1.0, 0.0
    pos, vel = zeldovich_host(16, start=100, count=2500)
    yield "ragged_pos", mc.FIELD_POSN, pos.reshape(-1), 1e-2, 100.0
    yield "single", mc.FIELD_VELC, np.array([1.0, 2.0, 3.0], np.float32), 1e-3, 0.0


FCASES = list(float_cases())


@pytest.mark.parametrize("case", FCASES, ids=[c[0] for c in FCASES])
def test_quantizer_alone_matches_oracle(codec, orc, case):
    """quant_QField as intended: 3N values and the Quantization attached"""
    _, code, data, delta, width = case
    n = data.size // 3
    hd, q = orc.quantize_field(code, data, delta, width)
    got, info = codec.quantize_field(code, data, delta, width)
    assert np.array_equal(got, q.astype(np.uint64))
    assert info["Depth"] == int(hd["Depth"])
    assert np.array_equal(np.array(info["X0"], np.float32).view(np.uint32), hd["X0"].view(np.uint32))
    assert np.array_equal(np.array(info["X1"], np.float32).view(np.uint32), hd["X1"].view(np.uint32))
    # quant_Field: the decoder's dither on given integers
    codec.set_dither_seed(4242)
    out = codec.dequantize_field(code, n, got, info)
    hd2 = hd.copy()
    hd2["DitherSeed"] = 4242
    want = orc.dequantize_field(hd2, got)
    assert np.array_equal(out.view(np.uint32), want.view(np.uint32))


def id_cases():
    w = 64
    ix, iy, iz = np.meshgrid(np.arange(50, 80) % w, np.arange(10, 20), np.arange(60, 68) % w, indexing="ij")
    yield "wrapping_slab", (ix + w * iy + w * w * iz).astype(np.uint64).T.reshape(-1), w
    yield "run", np.arange(5 * 128 * 128 + 17, 5 * 128 * 128 + 17 + 40000, dtype=np.uint64), 128
    rng = np.random.default_rng(3)
    yield "shuffled", np.arange(1000, 1000 + 30000, dtype=np.uint64)[rng.permutation(30000)], 256
    yield "single", np.array([12345], np.uint64), 100
    yield "n1025", np.arange(1025, dtype=np.uint64) * 3, 4096


ICASES = list(id_cases())


@pytest.mark.parametrize("case", ICASES, ids=[c[0] for c in ICASES])
def test_id_fields_match_oracle(codec, orc, case):
    _, ids, w = case
    x0, x1, q = orc.id_quantize(ids, w)
    got, info = codec.quantize_field(mc.FIELD_PTID, ids, width=w)
    assert np.array_equal(got, q) and info["X0"] == x0.tolist() and info["X1"] == x1.tolist() and info["Width"] == w
    assert np.array_equal(codec.dequantize_field(mc.FIELD_PTID, ids.size, got, info), ids)
    want = orc.encode_id_field(ids, w)
    payload = codec.encode_ints(mc.FIELD_PTID, ids, w)
    assert np.array_equal(payload, want)
    assert np.array_equal(codec.decode_ints(mc.FIELD_PTID, ids.size, payload), ids)
    # the literal plugin boundary carries the same bytes
    quant = mc.IDQuantization(w, (C.c_uint64 * 3)(*info["X0"]), (C.c_uint64 * 3)(*info["X1"]))
    qf = mc.QField(mc.FieldHeader(mc.FIELD_PTID, mc.ALGO_DIFF, mc.ALGO_VERSION, ids.size), 1, got.ctypes.data,
                   C.addressof(quant))
    cf = mc.lib().MnwCudaCompress_v0_9(qf, codec.ctx)
    lit = np.ctypeslib.as_array(C.cast(cf.Data, C.POINTER(C.c_uint8)), shape=(cf.DataLen,)).copy()
    mc._libc.free(cf.Data)
    assert np.array_equal(lit, want)


@pytest.mark.parametrize("n", [1, 2, 3, 1000, 3073, 50001])
def test_uint_fields_match_oracle(codec, orc, n):
    rng = np.random.default_rng(n)
    data = (np.uint64(1) << np.uint64(40)) + rng.integers(0, 1 << 20, n).astype(np.uint64)
    x0, x1, q = orc.uint_quantize(data)
    got, info = codec.quantize_field(mc.FIELD_UNSI, data)
    assert np.array_equal(got, q) and info == {"X0": x0, "X1": x1}
    assert np.array_equal(codec.dequantize_field(mc.FIELD_UNSI, n, got, info), data)
    want = orc.encode_uint_field(data)
    payload = codec.encode_ints(mc.FIELD_UNSI, data)
    assert np.array_equal(payload, want)
    assert np.array_equal(codec.decode_ints(mc.FIELD_UNSI, n, payload), data)


@pytest.mark.parametrize("grid,sizes", [(32, [4096, 5000, 1, 8192]),
                                        # a field whose block 0 (header + table, 28 KB) spans several staging
                                        # chunks of the per-field CTA, an empty field, fields of unequal size
                                        (96, [600000, 0, 3000, 70000]),
                                        (64, [65536, 65536, 65536])],  # equal sizes: no field search
                         ids=["small", "long_table_and_empty", "uniform"])
def test_block_headers_of_a_batch_match_oracle(codec, grid, sizes):
    """mnwcuda_BlockHeadersDev on the device-resident stream of a batch: every
    block's {Length, Checksum} equals the oracle's segment image table"""
    orc = _oracle.port()
    pos, vel = zeldovich_host(grid)
    fields, chunks, off, start = [], [], 0, 0
    for i, n in enumerate(sizes):
        src, code, delta, width = (pos, mc.FIELD_POSN, 1e-3, 100.0) if i % 2 == 0 else (vel, mc.FIELD_VELC, 0.5, 0.0)
        chunks.append(np.ascontiguousarray(src[:, start:start + n]).reshape(-1))
        fields.append((off, n, code, delta, width, 7 + i))
        off += 3 * n
        start += n
    d_in = torch.from_numpy(np.concatenate(chunks)).cuda()
    d_out = torch.zeros(sum(mc.field_bound(n) for n in sizes), dtype=torch.uint8, device="cuda")
    d_off = torch.zeros(len(sizes) + 1, dtype=torch.int64, device="cuda")
    descs = mc.make_descs(fields)
    codec.encode_dev(d_in, descs, len(sizes), d_out, d_off)
    hdr = codec.block_headers_dev(d_out, d_off, descs, len(sizes))
    assert codec.status() == 0
    offs = d_off.cpu().numpy()
    out = d_out.cpu().numpy()
    payloads = [(f[2], f[1], out[offs[i]:offs[i + 1]]) for i, f in enumerate(fields)]
    image = orc.seg_to_bytes(payloads, sum(sizes))
    nblk = hdr.shape[0]
    table = image[16 + 16 * len(sizes):16 + 16 * len(sizes) + 8 * nblk].view(np.int32).reshape(nblk, 2)
    assert np.array_equal(hdr.cpu().numpy(), table)


def _segment(h, orc, seed):
    pos, vel = zeldovich_host(16)
    n = pos.shape[1]
    pos, vel = pos.reshape(-1).copy(), vel.reshape(-1).copy()
    ids = np.arange(n, dtype=np.uint64)[::-1].copy()  # every id below Width^3
    pacc = mc.PositionAccuracy(None, 1e-3, 100.0, 0)
    vacc = mc.VelocityAccuracy(None, 0.5, 0, 0, 0.0)
    iacc = mc.IDAccuracy(16)
    fields = (mc.Field * 3)()
    fields[0] = mc.Field(mc.FieldHeader(mc.FIELD_POSN, mc.ALGO_DIFF, mc.ALGO_VERSION, n), 1, pos.ctypes.data, C.addressof(pacc))
    fields[1] = mc.Field(mc.FieldHeader(mc.FIELD_VELC, mc.ALGO_DIFF, mc.ALGO_VERSION, n), 1, vel.ctypes.data, C.addressof(vacc))
    fields[2] = mc.Field(mc.FieldHeader(mc.FIELD_PTID, mc.ALGO_DIFF, mc.ALGO_VERSION, n), 1, ids.ctypes.data, C.addressof(iacc))
    want = [(mc.FIELD_POSN, n, orc.encode_field(mc.FIELD_POSN, pos, 1e-3, 100.0, seed=seed)),
            (mc.FIELD_VELC, n, orc.encode_field(mc.FIELD_VELC, vel, 0.5, 0.0, seed=seed)),
            (mc.FIELD_PTID, n, orc.encode_id_field(ids, 16))]
    return fields, want, n, (pos, vel, ids, pacc, vacc, iacc)


def test_segment_image_roundtrip_and_block_damage(orc):
    """a pos + vel + id segment like getSegment() (src/funcs.c:212-266): encode,
    image bytes equal to the oracle's, back through the image with one damaged
    block: only that block's particles are lost"""
    h = mc.host_lib()
    fields, want, n, keep = _segment(h, orc, 31)
    h.MnwSeg_SetDitherSeed(31)
    cs = h.MnwSeg_Encode(mc.Seg(fields, 3))
    for i in range(3):
        cf = cs.Fields[i]
        got = np.ctypeslib.as_array(C.cast(cf.Data, C.POINTER(C.c_uint8)), shape=(cf.DataLen,))
        assert np.array_equal(got, want[i][2]), "field %d" % i
        assert cf.Checksum == orc.checksum(want[i][2])
    image = h.MnwSeg_ToBytes(cs, n)
    got = np.ctypeslib.as_array(C.cast(image.Data, C.POINTER(C.c_uint8)), shape=(image.Len,)).copy()
    ref_image = orc.seg_to_bytes(want, n)
    assert np.array_equal(got, ref_image)
    h.MnwCSeg_Free(cs)
    # clean round trip
    valid = (C.c_int32 * 3)()
    pn = C.c_int32()
    back = h.MnwSeg_FromBytes(image, valid, C.byref(pn))
    assert back.FieldLen == 3 and list(valid) == [1, 1, 1] and pn.value == n
    s = h.MnwSeg_Decode(back)
    for i in range(3):
        assert s.Fields[i].Valid == 1
    f0 = np.ctypeslib.as_array(C.cast(s.Fields[0].Data, C.POINTER(C.c_float)), shape=(3 * n,))
    assert np.array_equal(f0.view(np.uint32), orc.decode_field(want[0][2]).view(np.uint32))
    ids = np.ctypeslib.as_array(C.cast(s.Fields[2].Data, C.POINTER(C.c_uint64)), shape=(n,))
    assert np.array_equal(ids, keep[2])
    h.MnwSeg_Free(s)
    h.MnwCSeg_Free(back)
    # damage one byte of the velocity field's third packed block and one of the ids' first
    nb = (n + 1023) // 1024
    hdr_bytes = 16 + 48 + 8 * 3 * (1 + 3 * nb)
    bad = got.copy()
    vtab = _oracle.parse_table(want[1][2])
    bad[hdr_bytes + want[0][2].size + 64 + 48 * nb + int(vtab["Offset16"][2]) * 16 + 3] ^= 0x80
    itab = _oracle.parse_table(want[2][2])
    bad[hdr_bytes + want[0][2].size + want[1][2].size + 64 + 48 * nb + int(itab["Offset16"][0]) * 16 + 1] ^= 0x01
    ref_back, _ = orc.seg_from_bytes(bad)
    back = h.MnwSeg_FromBytes(mc.Bytes(bad.ctypes.data, bad.size), valid, C.byref(pn))
    assert back.FieldLen == 3 and list(valid) == [1, 1, 1]
    for i in range(3):
        cf = back.Fields[i]
        payload = np.ctypeslib.as_array(C.cast(cf.Data, C.POINTER(C.c_uint8)), shape=(cf.DataLen,))
        assert np.array_equal(payload, ref_back[i][2]), "field %d after damage" % i
    s = h.MnwSeg_Decode(back)
    v = np.ctypeslib.as_array(C.cast(s.Fields[1].Data, C.POINTER(C.c_float)), shape=(3 * n,))
    ref_v = orc.decode_field(ref_back[1][2])
    assert np.isnan(v).sum() == 1024 and np.array_equal(v.view(np.uint32), ref_v.view(np.uint32))
    ids = np.ctypeslib.as_array(C.cast(s.Fields[2].Data, C.POINTER(C.c_uint64)), shape=(n,))
    assert np.array_equal(ids, orc.decode_id_field(ref_back[2][2]))
    assert (ids[:1024] == np.uint64(0xFFFFFFFFFFFFFFFF)).all() and np.array_equal(ids[1024:], keep[2][1024:])
    h.MnwSeg_Free(s)
    h.MnwCSeg_Free(back)
    # a damaged field header block invalidates that field; a damaged segment header everything
    bad = got.copy()
    bad[hdr_bytes + 20] ^= 1
    back = h.MnwSeg_FromBytes(mc.Bytes(bad.ctypes.data, bad.size), valid, C.byref(pn))
    assert list(valid) == [0, 1, 1]
    h.MnwCSeg_Free(back)
    bad = got.copy()
    bad[9] ^= 1
    back = h.MnwSeg_FromBytes(mc.Bytes(bad.ctypes.data, bad.size), valid, C.byref(pn))
    assert back.FieldLen == -1
    mc._libc.free(image.Data)
