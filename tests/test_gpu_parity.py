"""GPU parity: libminnow_cuda (through its C ABI) against the CPU oracle on the
same seeded inputs. Bit-exact for quantised integers, headers, tables, packed
bytes and decoded floats."""
import numpy as np
import pytest

import _oracle
from _data import uniform_component, zeldovich_host

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
import minnow_c_b200 as mc  # noqa: E402


@pytest.fixture(scope="module")
def codec():
    c = mc.Codec(0)
    yield c
    c.close()


orc = _oracle.port()
# every oracle flavour present: the restated port and, where oracle/_ref was built
# (it travels to the GPU box), the same composition on the reference's own objects
FLAVOURS = _oracle.flavours()


@pytest.fixture(params=FLAVOURS, ids=[o.flavour for o in FLAVOURS])
def checker(request):
    return request.param


def cases():
    pos, vel = zeldovich_host(32)
    yield "zel_pos", mc.FIELD_POSN, pos.reshape(-1), 1e-3, 100.0
    yield "zel_vel", mc.FIELD_VELC, vel.reshape(-1), 1.0, 0.0
    yield "uniform_2^20", mc.FIELD_POSN, uniform_component(1 << 20, L=1.0), 1.0 / 2**14 * 1.0001, 1.0
    pos, vel = zeldovich_host(16, start=100, count=2500)
    yield "ragged_pos", mc.FIELD_POSN, pos.reshape(-1), 1e-2, 100.0
    yield "ragged_vel", mc.FIELD_VELC, vel.reshape(-1), 0.37, 0.0
    yield "single", mc.FIELD_POSN, np.array([1.0, 2.0, 3.0], np.float32), 1e-3, 10.0
    yield "const", mc.FIELD_VELC, np.full(3 * 1500, 2.5, np.float32), 1e-3, 0.0
    yield "n1025", mc.FIELD_VELC, np.arange(3 * 1025, dtype=np.float32), 0.5, 0.0
    yield "n8193_unaligned", mc.FIELD_POSN, uniform_component(8193, seed=3, L=7.0), 1e-4, 7.0
    rng = np.random.default_rng(5)
    yield "wide_vel", mc.FIELD_VELC, (rng.standard_normal(3 * 70000) * 500).astype(np.float32), 1e-3, 0.0
    yield "negzero", mc.FIELD_VELC, np.array([-0.0, 0.0, 1.0] * 3, np.float32), 0.1, 0.0
    # the decoder skips util_Periodic for components that cannot leave [0, L): one
    # component hugging the top of the box (dither carries values past L), one at the
    # bottom, one well inside, planar x | y | z
    n = 5000
    r = np.random.default_rng(9)
    L = np.float32(10.0)
    x = (L - np.float32(1e-6) - r.uniform(0, 0.02, n).astype(np.float32)).astype(np.float32)
    y = r.uniform(0, 0.02, n).astype(np.float32)
    y[0] = 0.0
    z = r.uniform(4.0, 4.5, n).astype(np.float32)
    yield "box_edges", mc.FIELD_POSN, np.concatenate([x, y, z]), 1e-3, float(L)
    yield "box_edges_coarse", mc.FIELD_POSN, np.concatenate([x, y, z]), 0.3, float(L)
    # alternating extremes at depth 24: deltas near +-2^24, width 25, minDelta < -2^23
    # (the decoder's integer scan instead of the FP32 one)
    v = np.where(np.arange(3 * 3000) % 2 == 0, 0.0, 1000.0).astype(np.float32)
    v[7::11] = 333.25
    yield "alternating_depth24", mc.FIELD_VELC, v, 1000.0 / 2**24 * 1.0001, 0.0
    # width 0 with minDelta < -2^23 (two values, top of the range then bottom): FP32
    # unpack, integer scan
    yield "deep_drop_depth24", mc.FIELD_VELC, np.array([900.0, 0.0] * 3, np.float32), 900.0 / 2**24 * 1.0001, 0.0


CASES = list(cases())


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_encode_field_bytes_match_oracle(codec, case, checker):
    _, code, data, delta, width = case
    codec.set_dither_seed(77)
    got = codec.encode_field(code, data, delta, width)
    want = checker.encode_field(code, data, delta, width, seed=77)
    assert got.size == want.size
    assert np.array_equal(got[:64], want[:64]), "field header differs"
    assert np.array_equal(got, want)


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_decode_field_floats_match_oracle(codec, case, checker):
    _, code, data, delta, width = case
    n = data.size // 3
    payload = checker.encode_field(code, data, delta, width, seed=123456789)
    want = checker.decode_field(payload)
    got, rep_delta = codec.decode_field(code, n, payload)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    hd = _oracle.parse_header(payload)
    assert rep_delta == np.float32((hd["X1"][0] - hd["X0"][0]) / np.float32(1 << int(hd["Depth"])))


def test_literal_plugin_boundary(codec):
    pos, vel = zeldovich_host(32)
    for code, data, delta, width in ((mc.FIELD_POSN, pos.reshape(-1), 1e-3, 100.0),
                                     (mc.FIELD_VELC, vel.reshape(-1), 1.0, 0.0)):
        hd, q = orc.quantize_field(code, data, delta, width)
        codec.set_dither_seed(0)
        got = codec.compress(code, q.astype(np.uint64), hd["X0"], hd["X1"], int(hd["Depth"]), width)
        want = orc.encode_q(hd, q.astype(np.uint64))
        assert np.array_equal(got, want)
        q2, info = codec.decompress(code, data.size // 3, got)
        assert np.array_equal(q2, q.astype(np.uint64))
        assert info["Depth"] == int(hd["Depth"]) and info["X0"] == [float(v) for v in hd["X0"]]


def test_batch_dev_matches_single_fields(codec):
    """ragged batch of mixed fields through the device-resident entry points"""
    ngrid = 32
    pos, vel = zeldovich_host(ngrid)
    sizes = [4096, 5000, 1, 0, 8192, 15479]
    fields, chunks, off, start = [], [], 0, 0
    for i, n in enumerate(sizes):
        src, code, delta, width = (pos, mc.FIELD_POSN, 1e-3, 100.0) if i % 2 == 0 else (vel, mc.FIELD_VELC, 0.5, 0.0)
        data = np.ascontiguousarray(src[:, start:start + n]).reshape(-1)
        fields.append((off, n, code, delta, width, 1000 + i % 2))
        chunks.append(data)
        off += 3 * n
        start += n
    flat = np.concatenate(chunks) if off else np.zeros(0, np.float32)
    d_in = torch.from_numpy(flat).cuda()
    cap = sum(mc.field_bound(n) for n in sizes)
    d_out = torch.zeros(cap, dtype=torch.uint8, device="cuda")
    d_off = torch.zeros(len(sizes) + 1, dtype=torch.int64, device="cuda")
    descs = mc.make_descs(fields)
    codec.encode_dev(d_in, descs, len(sizes), d_out, d_off)
    assert codec.status() == 0
    offs = d_off.cpu().numpy()
    out = d_out.cpu().numpy()
    payloads = []
    for i, (o, n, code, delta, width, seed) in enumerate(fields):
        want = orc.encode_field(code, chunks[i], delta, width, seed=seed)
        got = out[offs[i]:offs[i + 1]]
        assert np.array_equal(got, want), "field %d" % i
        payloads.append(want)
    d_dec = torch.zeros(max(off, 1), dtype=torch.float32, device="cuda")
    codec.decode_dev(d_out, d_off, descs, len(sizes), d_dec)
    assert codec.status() == 0
    dec = d_dec.cpu().numpy()
    for i, (o, n, code, delta, width, seed) in enumerate(fields):
        want = orc.decode_field(payloads[i])
        assert np.array_equal(dec[o:o + 3 * n].view(np.uint32), want.view(np.uint32)), "field %d" % i


def test_host_batch_entry_points(codec):
    pos, vel = zeldovich_host(32)
    n = 32 ** 3 // 4
    fields, off = [], 0
    for i in range(4):
        fields.append((off, n, mc.FIELD_POSN, 1e-3, 100.0, 5))
        off += 3 * n
    flat = np.concatenate([np.ascontiguousarray(pos[:, i * n:(i + 1) * n]).reshape(-1) for i in range(4)])
    descs = mc.make_descs(fields)
    h_out = np.zeros(sum(mc.field_bound(n) for _ in range(4)), np.uint8)
    h_off = np.zeros(5, np.uint64)
    total = codec.encode_host(flat, descs, 4, h_out, h_off)
    assert total == int(h_off[4])
    for i in range(4):
        want = orc.encode_field(mc.FIELD_POSN, flat[3 * n * i:3 * n * (i + 1)], 1e-3, 100.0, seed=5)
        assert np.array_equal(h_out[int(h_off[i]):int(h_off[i + 1])], want)
    back = np.zeros_like(flat)
    codec.decode_host(h_out, h_off, descs, 4, back)
    for i in range(4):
        want = orc.decode_field(h_out[int(h_off[i]):int(h_off[i + 1])])
        assert np.array_equal(back[3 * n * i:3 * n * (i + 1)].view(np.uint32), want.view(np.uint32))


def test_host_pipeline_many_chunks(monkeypatch):
    """The host-buffer entry points as a multi-chunk pipeline (two staging sets,
    three streams): tiny chunks force 7 chunks of ragged size, with empty fields."""
    monkeypatch.setenv("MNWCUDA_HOST_CHUNK_FLOATS", str(3 * 5000))
    c = mc.Codec(0)
    try:
        rng = np.random.default_rng(77)
        sizes = [4097, 100, 0, 2500, 2500, 1, 6000, 1023, 0, 3072, 2048, 777]
        fields, chunks, off = [], [], 0
        for i, n in enumerate(sizes):
            code = mc.FIELD_POSN if i % 2 == 0 else mc.FIELD_VELC
            if code == mc.FIELD_POSN:
                x = (rng.uniform(0, 100, (3, 1)) + np.cumsum(rng.normal(0, 0.02, (3, n)), axis=1)) % 100.0
                delta, width = 1e-3, 100.0
            else:
                x = np.cumsum(rng.normal(0, 3.0, (3, n)), axis=1)
                delta, width = 0.5, 0.0
            x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1)
            fields.append((off, n, code, delta, width, 100 + i))
            chunks.append(x)
            off += 3 * n
        flat = np.concatenate(chunks) if off else np.zeros(0, np.float32)
        descs = mc.make_descs(fields)
        nf = len(sizes)
        h_out = np.zeros(sum(mc.field_bound(n) for n in sizes), np.uint8)
        h_off = np.zeros(nf + 1, np.uint64)
        total = c.encode_host(flat, descs, nf, h_out, h_off)
        assert total == int(h_off[nf])
        for i, (o, n, code, delta, width, seed) in enumerate(fields):
            want = orc.encode_field(code, chunks[i], delta, width, seed=seed)
            assert np.array_equal(h_out[int(h_off[i]):int(h_off[i + 1])], want), "field %d" % i
        back = np.full_like(flat, np.nan)
        c.decode_host(h_out, h_off, descs, nf, back)
        for i, (o, n, code, delta, width, seed) in enumerate(fields):
            want = orc.decode_field(h_out[int(h_off[i]):int(h_off[i + 1])])
            assert np.array_equal(back[o:o + 3 * n].view(np.uint32), want.view(np.uint32)), "field %d" % i
    finally:
        c.close()


def test_depth_overflow_is_reported(codec):
    d_in = torch.arange(30, dtype=torch.float32, device="cuda")
    d_out = torch.zeros(mc.field_bound(10), dtype=torch.uint8, device="cuda")
    d_off = torch.zeros(2, dtype=torch.int64, device="cuda")
    codec.encode_dev(d_in, mc.make_descs([(0, 10, mc.FIELD_VELC, 1e-9, 0.0, 0)]), 1, d_out, d_off)
    assert codec.status() == 1
    assert codec.status() == 0


def test_generator_roundtrip_accuracy(codec):
    """device generator -> encode -> decode: error bound, and encode == oracle on the device's floats"""
    ngrid, n = 64, 64 ** 3
    pos = torch.empty(3 * n, dtype=torch.float32, device="cuda")
    vel = torch.empty(3 * n, dtype=torch.float32, device="cuda")
    codec.generate(ngrid, 100.0, 99, 0, n, pos, vel)
    torch.cuda.synchronize()
    hp = pos.cpu().numpy()
    assert hp.min() >= 0 and hp.max() < 100.0
    assert abs(float(vel.std()) - 300.0 / np.sqrt(3)) < 40
    codec.set_dither_seed(3)
    payload = codec.encode_field(mc.FIELD_POSN, hp, 1e-3, 100.0)
    assert np.array_equal(payload, orc.encode_field(mc.FIELD_POSN, hp, 1e-3, 100.0, seed=3))
    out, _ = codec.decode_field(mc.FIELD_POSN, n, payload)
    err = np.abs(out.astype(np.float64) - hp)
    err = np.minimum(err, 100.0 - err)
    assert err.max() <= 1e-3 + 2e-5
    assert payload.size < 0.6 * hp.nbytes


def test_bench_sized_slab_matches_oracle(codec, checker):
    orc = checker  # noqa: F841 (shadows the module-level port on purpose)
    _bench_sized_slab(codec, checker)


def _bench_sized_slab(codec, orc):
    """Eight segments of 2^22 particles, positions and velocities (the bench's
    field size: 4096 blocks per component, every fast path of the persistent
    encoder) in one batch: every field's bytes against the oracle, twice (the
    schedule differs from run to run, the bytes must not), and the decoded
    floats of two fields against the oracle's decoder."""
    seg, nseg, ngrid = 1 << 22, 8, 512
    nf = 2 * nseg
    d_in = torch.empty(nf * 3 * seg, dtype=torch.float32, device="cuda")
    fields = []
    for s_ in range(nseg):
        po, vo = (2 * s_) * 3 * seg, (2 * s_ + 1) * 3 * seg
        codec.generate(ngrid, 100.0, 4242, s_ * seg, seg, d_in[po:po + 3 * seg], d_in[vo:vo + 3 * seg])
        fields.append((po, seg, mc.FIELD_POSN, 1e-3, 100.0, 1000 + s_))
        fields.append((vo, seg, mc.FIELD_VELC, 1.0, 0.0, 2000 + s_))
    torch.cuda.synchronize()
    descs = mc.make_descs(fields)
    d_out = torch.zeros(nf * mc.field_bound(seg), dtype=torch.uint8, device="cuda")
    d_off = torch.zeros(nf + 1, dtype=torch.int64, device="cuda")
    codec.encode_dev(d_in, descs, nf, d_out, d_off)
    assert codec.status() == 0
    offs = d_off.cpu().numpy()
    first = d_out[:int(offs[nf])].cpu().numpy().copy()
    h_in = d_in.cpu().numpy()
    for i, (o, n, code, delta, width, seed) in enumerate(fields):
        want = orc.encode_field(code, h_in[o:o + 3 * n], delta, width, seed=seed)
        assert np.array_equal(first[offs[i]:offs[i + 1]], want), "field %d" % i
    d_out.zero_()
    codec.encode_dev(d_in, descs, nf, d_out, d_off)
    assert codec.status() == 0
    assert np.array_equal(d_off.cpu().numpy(), offs)
    assert np.array_equal(d_out[:int(offs[nf])].cpu().numpy(), first)
    d_dec = torch.zeros_like(d_in)
    codec.decode_dev(d_out, d_off, descs, nf, d_dec)
    assert codec.status() == 0
    for i in (4, 11):
        o, n = fields[i][0], fields[i][1]
        want = orc.decode_field(first[offs[i]:offs[i + 1]])
        assert np.array_equal(d_dec[o:o + 3 * n].cpu().numpy().view(np.uint32), want.view(np.uint32)), "field %d" % i


@pytest.mark.parametrize("depth", list(range(0, 25)))
def test_depth_sweep_matches_oracle(codec, depth, checker):
    orc = checker
    """BASELINE config 4: the same data at every quantisation depth 0..24 (Delta
    chosen so that deltaToDepth, src/quant.c:684-688, lands on `depth`), encode
    bytes and decoded floats against the oracle. Depth 24 takes the encoder's
    unbiased truncation path, depths <= 23 the biased one."""
    rng = np.random.default_rng(900 + depth)
    n = 20000 + 37 * depth
    x = (50.0 + np.cumsum(rng.normal(0, 0.01, (3, n)), axis=1)) % 100.0
    x = np.ascontiguousarray(x, dtype=np.float32)
    # unwrapped range as the encoder sees it (first element anchors the unwrap)
    span = 0.0
    for c in range(3):
        u = x[c].astype(np.float32).copy()
        rel = u - u[0]
        u = np.where(rel >= 50.0, u - np.float32(100.0), np.where(rel < -50.0, u + np.float32(100.0), u))
        span = max(span, float(np.float32(u.max()) - np.float32(u.min())))
    delta = np.float32(span / 2.0 ** depth * 1.0001) if depth > 0 else np.float32(span * 2)
    for code, width in ((mc.FIELD_POSN, 100.0), (mc.FIELD_VELC, 0.0)):
        data = x.reshape(-1)
        want = orc.encode_field(code, data, float(delta), width, seed=depth)
        codec.set_dither_seed(depth)
        got = codec.encode_field(code, data, float(delta), width)
        assert np.array_equal(got, want), "encode depth %d code %x" % (depth, code)
        out, _ = codec.decode_field(code, n, got)
        ref = orc.decode_field(want)
        assert np.array_equal(out.view(np.uint32), ref.view(np.uint32)), "decode depth %d code %x" % (depth, code)


def test_tiny_range_field_decodes_like_the_oracle(codec):
    """A field whose bin width is below 2^-100: the decoder's folded dither scale
    (bw / 2^24) would underflow, so it must take the reference's two-step product."""
    rng = np.random.default_rng(31)
    n = 5000
    x = (rng.uniform(0, 1, 3 * n) * 1e-33).astype(np.float32)
    delta = float(np.float32(1e-36))
    want = orc.encode_field(mc.FIELD_VELC, x, delta, 0.0, seed=9)
    codec.set_dither_seed(9)
    got = codec.encode_field(mc.FIELD_VELC, x, delta, 0.0)
    assert np.array_equal(got, want)
    out, _ = codec.decode_field(mc.FIELD_VELC, n, got)
    ref = orc.decode_field(want)
    assert np.array_equal(out.view(np.uint32), ref.view(np.uint32))
