"""GPU: mnwcuda_DecodeFieldNoAlloc (include/minnow_cuda.h) -- the decode of one
float field whose scratch is a static device array -- against the existing
mnwcuda_DecodeFieldsDev on the same payload, bit for bit (the tolerance of every
decode test here), on two calls in a row. Collected last on purpose."""
import numpy as np
import pytest

from _data import zeldovich_host

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
import minnow_c_b200 as mc  # noqa: E402


@pytest.fixture(scope="module")
def codec():
    c = mc.Codec(0)
    yield c
    c.close()


def _payload(codec, code, data, n, delta, width, seed):
    """one field through the existing encoder; payload on the device, 16-byte aligned"""
    d_in = torch.from_numpy(data).cuda()
    d_out = torch.zeros(mc.field_bound(n), dtype=torch.uint8, device="cuda")
    d_off = torch.zeros(2, dtype=torch.int64, device="cuda")
    descs = mc.make_descs([(0, n, code, delta, width, seed)])
    codec.encode_dev(d_in, descs, 1, d_out, d_off)
    assert codec.status() == 0
    length = int(d_off[1].item())
    return d_out, length, descs


def _existing(codec, d_payload, length, descs, n):
    d_off = torch.tensor([0, length], dtype=torch.int64, device="cuda")
    out = torch.full((3 * n,), -7.0, dtype=torch.float32, device="cuda")
    codec.decode_dev(d_payload, d_off, descs, 1, out)
    assert codec.status() == 0
    return out.cpu().numpy()


def _no_alloc(d_payload, length, n, code, status):
    out = torch.full((max(3 * n, 1),), -7.0, dtype=torch.float32, device="cuda")
    rc = mc.lib().mnwcuda_DecodeFieldNoAlloc(0, d_payload.data_ptr(), length, n, code, out.data_ptr(),
                                             status.data_ptr(), None)
    assert rc == 0
    torch.cuda.synchronize()
    return out.cpu().numpy()[:3 * n]


def _cases():
    pos, vel = zeldovich_host(32)
    yield "pos", mc.FIELD_POSN, pos.reshape(-1).copy(), pos.shape[1], 1e-3, 100.0, 77
    yield "vel", mc.FIELD_VELC, vel.reshape(-1).copy(), vel.shape[1], 0.5, 0.0, 78
    pos, _ = zeldovich_host(16, start=100, count=2500)
    yield "ragged", mc.FIELD_POSN, pos.reshape(-1).copy(), 2500, 1e-2, 100.0, 79
    yield "one", mc.FIELD_VELC, np.array([1.0, 2.0, 3.0], np.float32), 1, 1e-3, 0.0, 80


CASES = list(_cases())


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_no_alloc_decode_equals_existing_decode_twice(codec, case):
    _, code, data, n, delta, width, seed = case
    d_payload, length, descs = _payload(codec, code, data, n, delta, width, seed)
    want = _existing(codec, d_payload, length, descs, n)
    status = torch.full((1,), -1, dtype=torch.int32, device="cuda")
    for _ in range(2):  # the array keeps the last call's contents: the second call must not see them
        got = _no_alloc(d_payload, length, n, code, status)
        assert int(status.item()) == 0
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_no_alloc_calls_of_different_fields_alternate(codec):
    """a large field, a small one, the large one again: leftovers of a longer
    call (stream states, field record) never leak into the next"""
    made = []
    for case in (CASES[0], CASES[3], CASES[1]):
        _, code, data, n, delta, width, seed = case
        d_payload, length, descs = _payload(codec, code, data, n, delta, width, seed)
        made.append((d_payload, length, n, code, _existing(codec, d_payload, length, descs, n)))
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    for d_payload, length, n, code, want in made + made[::-1]:
        got = _no_alloc(d_payload, length, n, code, status)
        assert int(status.item()) == 0
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_no_alloc_allocates_nothing(codec):
    """after the first call on the device (set-up), a call leaves both the driver's
    and torch's view of device memory unchanged"""
    _, code, data, n, delta, width, seed = CASES[0]
    d_payload, length, descs = _payload(codec, code, data, n, delta, width, seed)
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    out = torch.empty(3 * n, dtype=torch.float32, device="cuda")
    args = (0, d_payload.data_ptr(), length, n, code, out.data_ptr(), status.data_ptr(), None)
    assert mc.lib().mnwcuda_DecodeFieldNoAlloc(*args) == 0
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    reserved0 = torch.cuda.memory_reserved()
    for _ in range(3):
        assert mc.lib().mnwcuda_DecodeFieldNoAlloc(*args) == 0
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0 and torch.cuda.memory_reserved() == reserved0


def test_no_alloc_rejects_what_it_cannot_hold(codec, capfd):
    _, code, data, n, delta, width, seed = CASES[3]
    d_payload, length, _ = _payload(codec, code, data, n, delta, width, seed)
    out = torch.empty(16, dtype=torch.float32, device="cuda")
    f = mc.lib().mnwcuda_DecodeFieldNoAlloc
    big = mc.NOALLOC_MAX_PARTICLES + 1
    assert f(0, d_payload.data_ptr(), mc.field_bound(big), big, code, out.data_ptr(), None, None) == -7
    assert "MNWCUDA_NOALLOC_MAX_PARTICLES" in capfd.readouterr().err
    assert f(0, d_payload.data_ptr(), length, n, mc.FIELD_PTID, out.data_ptr(), None, None) == -3
    assert f(0, d_payload.data_ptr() + 4, length, n, code, out.data_ptr(), None, None) == -6
    assert f(0, d_payload.data_ptr(), 10, n, code, out.data_ptr(), None, None) == -1  # shorter than header + table
    assert f(0, d_payload.data_ptr(), length, -1, code, out.data_ptr(), None, None) == -1


def test_no_alloc_reports_a_malformed_payload(codec):
    """the wrong ParticleLen for the payload: status 3 in the caller's word, and the
    next good call starts from a clean status"""
    _, code, data, n, delta, width, seed = CASES[2]
    d_payload, length, descs = _payload(codec, code, data, n, delta, width, seed)
    want = _existing(codec, d_payload, length, descs, n)
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    out = torch.empty(3 * n, dtype=torch.float32, device="cuda")
    rc = mc.lib().mnwcuda_DecodeFieldNoAlloc(0, d_payload.data_ptr(), length, n - 1, code, out.data_ptr(),
                                             status.data_ptr(), None)
    torch.cuda.synchronize()
    assert rc == 0 and int(status.item()) == 3
    got = _no_alloc(d_payload, length, n, code, status)
    assert int(status.item()) == 0 and np.array_equal(got.view(np.uint32), want.view(np.uint32))
