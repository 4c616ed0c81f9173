"""GPU tests added in round 2: shard invariance, per-field dither seeds through
the tabulated jump matrices, malformed payloads, the rough generator."""
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


def _slab(codec, nseg, seg, ngrid, seed0):
    nf = 2 * nseg
    d_in = torch.empty(nf * 3 * seg, dtype=torch.float32, device="cuda")
    fields = []
    for s_ in range(nseg):
        po, vo = (2 * s_) * 3 * seg, (2 * s_ + 1) * 3 * seg
        codec.generate(ngrid, 100.0, 777, s_ * seg, seg, d_in[po:po + 3 * seg], d_in[vo:vo + 3 * seg])
        fields.append((po, seg, mc.FIELD_POSN, 1e-3, 100.0, seed0 + 2 * s_))
        fields.append((vo, seg, mc.FIELD_VELC, 1.0, 0.0, seed0 + 2 * s_ + 1))
    torch.cuda.synchronize()
    return d_in, fields


def _encode(codec, d_in, fields):
    nf = len(fields)
    d_out = torch.zeros(sum(mc.field_bound(f[1]) for f in fields), dtype=torch.uint8, device="cuda")
    d_off = torch.zeros(nf + 1, dtype=torch.int64, device="cuda")
    codec.encode_dev(d_in, mc.make_descs(fields), nf, d_out, d_off)
    assert codec.status() == 0
    offs = d_off.cpu().numpy()
    return d_out[:int(offs[nf])].cpu().numpy(), offs


def test_shard_invariance(codec):
    """segments [0,8) encoded in one batch and as two batches
    [0,4) + [4,8) give the same per-field bytes, and the second shard's offsets
    are the single batch's offsets minus the first shard's size (what
    exchange_sizes' prefix adds back)."""
    seg, nseg = 1 << 18, 8
    d_in, fields = _slab(codec, nseg, seg, 256, 5000)
    whole, offs = _encode(codec, d_in, fields)
    half = len(fields) // 2
    cut = fields[half][0]
    a, offs_a = _encode(codec, d_in[:cut], fields[:half])
    fb = [(f[0] - cut,) + f[1:] for f in fields[half:]]
    b, offs_b = _encode(codec, d_in[cut:], fb)
    assert np.array_equal(np.concatenate([a, b]), whole)
    assert np.array_equal(offs_a, offs[:half + 1])
    assert np.array_equal(offs_b + offs_a[half], offs[half:])


@pytest.mark.parametrize("checker", FLAVOURS, ids=[o.flavour for o in FLAVOURS])
def test_per_field_seeds_use_the_jump_table(codec, checker):
    """20 fields, every one with its own dither seed: the decoder takes the
    tabulated-J^k path (16 fields or more); floats against the oracle's decoder,
    which walks rand_Seed's jumps one by one."""
    rng = np.random.default_rng(12)
    sizes = [33000, 1024, 70001, 5, 40960] * 4
    fields, chunks, off = [], [], 0
    for i, n in enumerate(sizes):
        code = mc.FIELD_POSN if i % 2 == 0 else mc.FIELD_VELC
        if code == mc.FIELD_POSN:
            x = (rng.uniform(0, 100, (3, 1)) + np.cumsum(rng.normal(0, 0.02, (3, n)), axis=1)) % 100.0
            delta, width = 1e-3, 100.0
        else:
            x = np.cumsum(rng.normal(0, 3.0, (3, n)), axis=1)
            delta, width = 0.5, 0.0
        fields.append((off, n, code, delta, width, 0xabc000 + 7919 * i))
        chunks.append(np.ascontiguousarray(x, dtype=np.float32).reshape(-1))
        off += 3 * n
    flat = np.concatenate(chunks)
    d_in = torch.from_numpy(flat).cuda()
    payload, offs = _encode(codec, d_in, fields)
    nf = len(fields)
    d_out = torch.from_numpy(payload).cuda()
    d_off = torch.from_numpy(offs.astype(np.int64)).cuda()
    d_dec = torch.zeros_like(d_in)
    codec.decode_dev(d_out, d_off, mc.make_descs(fields), nf, d_dec)
    assert codec.status() == 0
    dec = d_dec.cpu().numpy()
    for i, (o, n, code, delta, width, seed) in enumerate(fields):
        want_bytes = checker.encode_field(code, chunks[i], delta, width, seed=seed)
        assert np.array_equal(payload[offs[i]:offs[i + 1]], want_bytes), "field %d bytes" % i
        want = checker.decode_field(want_bytes)
        assert np.array_equal(dec[o:o + 3 * n].view(np.uint32), want.view(np.uint32)), "field %d floats" % i
    # the same batch again with a longest field twice as long: the table is rebuilt
    fields2 = fields + [(off, 150000, mc.FIELD_VELC, 0.5, 0.0, 99)]
    extra = np.cumsum(rng.normal(0, 3.0, (3, 150000)), axis=1).astype(np.float32).reshape(-1)
    d_in2 = torch.from_numpy(np.concatenate([flat, extra])).cuda()
    payload2, offs2 = _encode(codec, d_in2, fields2)
    d_dec2 = torch.zeros_like(d_in2)
    codec.decode_dev(torch.from_numpy(payload2).cuda(), torch.from_numpy(offs2.astype(np.int64)).cuda(),
                     mc.make_descs(fields2), nf + 1, d_dec2)
    assert codec.status() == 0
    want = checker.decode_field(payload2[offs2[nf]:offs2[nf + 1]])
    assert np.array_equal(d_dec2[off:].cpu().numpy().view(np.uint32), want.view(np.uint32))


def test_malformed_streams_raise_payload_error(codec):
    """A truncated region, a misaligned offset, offsets out of order and a table
    entry pointing past the region all end in status 3, never in a device fault."""
    pos, _ = zeldovich_host(32)
    n = 32 ** 3
    orc = _oracle.port()
    good = orc.encode_field(mc.FIELD_POSN, pos.reshape(-1), 1e-3, 100.0, seed=1)
    fields = [(0, n, mc.FIELD_POSN, 1e-3, 100.0, 1)]
    descs = mc.make_descs(fields)
    d_dec = torch.zeros(3 * n, dtype=torch.float32, device="cuda")

    def status_of(buf, offs):
        d_in = torch.from_numpy(buf).cuda()
        d_off = torch.tensor(offs, dtype=torch.int64, device="cuda")
        codec.decode_dev(d_in, d_off, descs, 1, d_dec)
        return codec.status()

    assert status_of(good, [0, good.size]) == 0
    assert status_of(good, [0, 64 + 48 * 10]) == 3                    # shorter than header + table
    assert status_of(good, [0, 63]) == 3
    pad = np.concatenate([np.zeros(8, np.uint8), good])
    assert status_of(pad, [8, 8 + good.size]) == 3                     # not 16-byte aligned
    assert status_of(good, [good.size, 0]) == 3                        # offsets out of order
    bad = good.copy()
    tab = bad[64:64 + 48 * 32].view(_oracle.ENTRY_DTYPE)
    tab["Offset16"][5] = 0x0fffffff                                    # a block past the end
    assert status_of(bad, [0, bad.size]) == 3
    cut = good[:good.size - 4096].copy()
    assert status_of(cut, [0, cut.size]) == 3                          # payload truncated
    assert status_of(good, [0, good.size]) == 0                        # the context is still usable


def test_offset_must_not_be_negative(codec):
    d = torch.zeros(64, dtype=torch.float32, device="cuda")
    o = torch.zeros(mc.field_bound(4), dtype=torch.uint8, device="cuda")
    off = torch.zeros(2, dtype=torch.int64, device="cuda")
    with pytest.raises(ValueError):
        codec.encode_dev(d, mc.make_descs([(-4, 4, mc.FIELD_VELC, 1.0, 0.0, 0)]), 1, o, off)


def test_cfunc_rejects_values_beyond_the_depth(codec, capfd):
    """QField integers that do not fit the declared depth raise kErrRange
    (status 4 -> Panic in the plugin function; checked here through the batch
    status by encoding in a subprocess-free way: the device flag)."""
    import ctypes as C
    import subprocess
    import sys
    code = (
        "import numpy as np, minnow_c_b200 as mc\n"
        "c = mc.Codec(0)\n"
        "q = np.arange(3 * 2000, dtype=np.uint64) % 1000\n"
        "q[4321] = 1 << 40\n"
        "c.compress(mc.FIELD_VELC, q, [0, 0, 0], [1, 1, 1], 10)\n"
    )
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=mc.ROOT)
    assert r.returncode == 1 and "do not fit the declared depth" in r.stderr


def test_rough_generator_is_deterministic_and_wider(codec):
    n, ngrid = 1 << 16, 64
    a = torch.empty(3 * n, dtype=torch.float32, device="cuda")
    b = torch.empty(3 * n, dtype=torch.float32, device="cuda")
    c = torch.empty(3 * n, dtype=torch.float32, device="cuda")
    codec.generate(ngrid, 100.0, 5, 0, n, a, None, noise_cells=0.3)
    codec.generate(ngrid, 100.0, 5, 0, n // 2, b[:3 * (n // 2)], None, noise_cells=0.3)
    codec.generate(ngrid, 100.0, 5, 0, n, c, None)
    torch.cuda.synchronize()
    ha, hb, hc = a.cpu().numpy(), b.cpu().numpy(), c.cpu().numpy()
    # the noise of a particle does not depend on how the ids are cut into calls
    assert np.array_equal(ha[:n // 2], hb[:n // 2])
    assert ha.min() >= 0 and ha.max() < 100.0
    d = np.abs(ha - hc)
    d = np.minimum(d, 100.0 - d)
    assert 0.2 * 100.0 / ngrid < d.max() <= 0.3 * 100.0 / ngrid * 1.001
    orc = _oracle.port()
    codec.set_dither_seed(1)
    rough = codec.encode_field(mc.FIELD_POSN, ha, 1e-3, 100.0)
    smooth = codec.encode_field(mc.FIELD_POSN, hc, 1e-3, 100.0)
    assert np.array_equal(rough, orc.encode_field(mc.FIELD_POSN, ha, 1e-3, 100.0, seed=1))
    assert rough.size > smooth.size


def _small_batch(codec_, seed0):
    d_in, fields = _slab(codec_, 4, 1 << 18, 256, seed0)
    return d_in, fields


def test_encode_while_other_kernels_hold_sms(codec):
    """k_encode's CTAs wait for each other, so the launch is cooperative: with
    long-running kernels of other streams sitting on SMs it starts when the whole
    grid fits, and ends with the right bytes -- never with a timeout (status 5)."""
    d_in, fields = _small_batch(codec, 100)
    want, offs_want = _encode(codec, d_in, fields)
    streams = [torch.cuda.Stream() for _ in range(12)]
    for st in streams:  # ~50 ms of spinning single-CTA kernels, one per stream
        with torch.cuda.stream(st):
            torch.cuda._sleep(100_000_000)
    enc_stream = torch.cuda.Stream()
    nf = len(fields)
    d_out = torch.zeros(sum(mc.field_bound(f[1]) for f in fields), dtype=torch.uint8, device="cuda")
    d_off = torch.zeros(nf + 1, dtype=torch.int64, device="cuda")
    with torch.cuda.stream(enc_stream):
        codec.encode_dev(d_in, mc.make_descs(fields), nf, d_out, d_off, stream=enc_stream.cuda_stream)
        assert codec.status(enc_stream.cuda_stream) == 0
    torch.cuda.synchronize()
    offs = d_off.cpu().numpy()
    assert np.array_equal(offs, offs_want)
    assert np.array_equal(d_out[:int(offs[nf])].cpu().numpy(), want)


def test_two_contexts_encode_concurrently():
    """The reference makes one Compressor per field (src/funcs.c:145-158): two
    contexts, two host threads, each encoding its own batch on its own stream at the
    same time. Both finish with status 0 and the bytes of a solo run."""
    import threading
    ca, cb = mc.Codec(0), mc.Codec(0)
    try:
        jobs = []
        for c, seed0 in ((ca, 300), (cb, 400)):
            d_in, fields = _small_batch(c, seed0)
            want, offs = _encode(c, d_in, fields)
            jobs.append((c, d_in, fields, want, offs))
        results = [None, None]

        def run(i):
            c, d_in, fields, _, _ = jobs[i]
            torch.cuda.set_device(0)
            st = torch.cuda.Stream()
            nf = len(fields)
            d_out = torch.zeros(sum(mc.field_bound(f[1]) for f in fields), dtype=torch.uint8, device="cuda")
            d_off = torch.zeros(nf + 1, dtype=torch.int64, device="cuda")
            torch.cuda.synchronize()
            ok = True
            for _ in range(20):
                c.encode_dev(d_in, mc.make_descs(fields), nf, d_out, d_off, stream=st.cuda_stream)
                ok = ok and c.status(st.cuda_stream) == 0
            offs = d_off.cpu().numpy()
            results[i] = (ok, offs, d_out[:int(offs[nf])].cpu().numpy())

        th = [threading.Thread(target=run, args=(i,)) for i in range(2)]
        for t in th:
            t.start()
        for t in th:
            t.join()
        for i in range(2):
            ok, offs, got = results[i]
            assert ok
            assert np.array_equal(offs, jobs[i][4])
            assert np.array_equal(got, jobs[i][3])
    finally:
        ca.close()
        cb.close()


def test_c_nccl_driver_world_of_one(codec):
    """libminnow_nccl.so through its C ABI with a communicator of one rank: the
    sharded encode is the plain encode, sizes[0] = the packed bytes, base = 0,
    gather and scatter move the slab. (N >= 2 runs through bench.py --gpus N.)"""
    d_in, fields = _slab(codec, 2, 1 << 16, 128, 900)
    want, offs_want = _encode(codec, d_in, fields)
    nf = len(fields)
    sh = mc.Sharded(codec, 0, 1, mc.Sharded.unique_id())
    try:
        d_out = torch.zeros(sum(mc.field_bound(f[1]) for f in fields), dtype=torch.uint8, device="cuda")
        d_off = torch.zeros(nf + 1, dtype=torch.int64, device="cuda")
        d_sizes = torch.full((1,), -1, dtype=torch.int64, device="cuda")
        d_base = torch.full((1,), -1, dtype=torch.int64, device="cuda")
        sh.encode(d_in, mc.make_descs(fields), nf, d_out, d_off, d_sizes, d_base)
        assert codec.status() == 0
        assert int(d_sizes[0]) == want.size and int(d_base[0]) == 0
        assert np.array_equal(d_off.cpu().numpy(), offs_want)
        d_glob = torch.zeros(want.size, dtype=torch.uint8, device="cuda")
        sh.gather(d_out, [want.size], 0, d_glob)
        torch.cuda.synchronize()
        assert np.array_equal(d_glob.cpu().numpy(), want)
        d_loc = torch.zeros(want.size, dtype=torch.uint8, device="cuda")
        sh.scatter(d_glob, [want.size], 0, d_loc)
        d_dec = torch.zeros_like(d_in)
        sh.decode(d_loc, d_off, mc.make_descs(fields), nf, d_dec)
        assert codec.status() == 0
        err = (d_dec - d_in).abs()
        assert float(err[3 * (1 << 16):6 * (1 << 16)].max()) <= 1.0 + 1e-3  # velocities of segment 0: Delta = 1
    finally:
        sh.close()


def test_many_fields_of_unequal_size_match_oracle(codec):
    """the general parallel reduce (k_encode<.., kGeneral>): thousands of fields of random
    sizes in one batch -- empty ones, single particles, fields of one block, two long ones
    whose components span many reduce windows -- so that a window of 512 blocks holds
    hundreds of component boundaries; every field byte for byte, then the round trip"""
    orc = _oracle.reference() or _oracle.port()
    rng = np.random.default_rng(5)
    nf = 2500
    sizes = rng.integers(0, 3000, nf)
    sizes[rng.integers(0, nf, 40)] = 0
    sizes[rng.integers(0, nf, 40)] = 1
    sizes[[7, 1900]] = [70000, 131072]
    sizes[-3:] = [0, 5, 0]  # the batch ends in empty fields
    total = int(sizes.sum())
    walk = np.cumsum(rng.normal(0.0, 0.01, 3 * total)).astype(np.float32)
    fields, off = [], 0
    data = np.empty(3 * total, np.float32)
    for i, n in enumerate(sizes.tolist()):
        pos = i % 2 == 0
        chunk = walk[off:off + 3 * n]
        data[off:off + 3 * n] = np.mod(chunk + 50.0, 100.0).astype(np.float32) if pos else chunk
        fields.append((off, n, mc.FIELD_POSN if pos else mc.FIELD_VELC, 1e-3 if pos else 1e-4, 100.0 if pos else 0.0, 900 + i))
        off += 3 * n
    data[data >= np.float32(100.0)] = 0.0
    d_in = torch.from_numpy(data).cuda()
    d_out = torch.zeros(sum(mc.field_bound(int(n)) for n in sizes), dtype=torch.uint8, device="cuda")
    d_off = torch.zeros(nf + 1, dtype=torch.int64, device="cuda")
    descs = mc.make_descs(fields)
    codec.encode_dev(d_in, descs, nf, d_out, d_off)
    assert codec.status() == 0
    offs, out = d_off.cpu().numpy(), d_out.cpu().numpy()
    for i, (o, n, code, delta, width, seed) in enumerate(fields):
        want = orc.encode_field(code, data[o:o + 3 * n], delta, width, seed=seed)
        assert np.array_equal(out[offs[i]:offs[i + 1]], want), "field %d (n = %d)" % (i, n)
    d_dec = torch.empty_like(d_in)
    codec.decode_dev(d_out, d_off, descs, nf, d_dec)
    assert codec.status() == 0
    err = (d_dec - d_in).abs().cpu().numpy()
    err = np.minimum(err, np.abs(100.0 - err))
    tol = np.concatenate([np.full(3 * n, 1e-3 if i % 2 == 0 else 1e-4) for i, n in enumerate(sizes.tolist())])
    assert (err <= tol * 1.0001).all()
