"""The Diff v0.9 composition: restated oracle vs the same composition on the
reference's own primitives, round trips, accuracy bound, edge cases. CPU only."""
import numpy as np
import pytest

import _oracle
from _data import uniform_component, zeldovich_host

port = _oracle.port()
ref = _oracle.reference()


def field_cases():
    pos, vel = zeldovich_host(32)  # 32768 particles
    yield "zel_pos", _oracle.FIELD_POSN, pos.reshape(-1), 1e-3, 100.0
    yield "zel_vel", _oracle.FIELD_VELC, vel.reshape(-1), 1.0, 0.0
    yield "uniform", _oracle.FIELD_POSN, uniform_component(5000, L=1.0), 1.0 / 2**14 * 1.0001, 1.0
    pos, vel = zeldovich_host(16, start=100, count=2500)  # ragged: 2500 = 2*1024 + 452
    yield "ragged_pos", _oracle.FIELD_POSN, pos.reshape(-1), 1e-2, 100.0
    yield "ragged_vel", _oracle.FIELD_VELC, vel.reshape(-1), 0.37, 0.0
    yield "single", _oracle.FIELD_POSN, np.array([1.0, 2.0, 3.0], np.float32), 1e-3, 10.0
    yield "const", _oracle.FIELD_VELC, np.full(3 * 1500, 2.5, np.float32), 1e-3, 0.0
    yield "n1025", _oracle.FIELD_VELC, np.arange(3 * 1025, dtype=np.float32), 0.5, 0.0


CASES = list(field_cases())


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_roundtrip_and_bound(case):
    _, code, data, delta, width = case
    n = data.size // 3
    payload, q = port.encode_field(code, data, delta, width, seed=42, want_q=True)
    hd = _oracle.parse_header(payload)
    assert hd["Magic"] == _oracle.ALGO_DIFF and hd["ParticleLen"] == n
    out, q2 = port.decode_field(payload, want_q=True)
    assert np.array_equal(q, q2)
    md = max(float(hd["X1"][c]) - float(hd["X0"][c]) for c in range(3))
    binw = md / (1 << int(hd["Depth"]))
    assert binw <= delta * 1.0000001 or md == 0
    err = np.abs(out.astype(np.float64) - data)
    if code == _oracle.FIELD_POSN:
        err = np.minimum(err, np.abs(width - err))
    scale = max(1.0, float(np.abs(data).max()))
    assert err.max() <= binw + 4e-7 * scale * 4
    # the table is consistent with the payload length
    tab = _oracle.parse_table(payload)
    assert tab["Width"].max() <= 25
    assert np.all(np.diff(tab["Offset16"].astype(np.int64)) >= 0)


@pytest.mark.skipif(ref is None, reason="oracle/_ref not built")
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_port_equals_reference_composition(case):
    _, code, data, delta, width = case
    a = port.encode_field(code, data, delta, width, seed=9)
    b = ref.encode_field(code, data, delta, width, seed=9)
    assert np.array_equal(a, b)
    fa = port.decode_field(a)
    fb = ref.decode_field(a)
    assert np.array_equal(fa.view(np.uint32), fb.view(np.uint32))


def test_literal_q_path_roundtrip():
    pos, _ = zeldovich_host(16)
    hd, q = port.quantize_field(_oracle.FIELD_POSN, pos.reshape(-1), 1e-3, 100.0)
    payload = port.encode_q(hd, q.astype(np.uint64))
    assert np.array_equal(port.decode_q(payload), q.astype(np.uint64))
    # and it is the same payload the float entry point makes (seed 0)
    assert np.array_equal(payload, port.encode_field(_oracle.FIELD_POSN, pos.reshape(-1), 1e-3, 100.0, seed=0))


def test_depth_overflow_is_an_error():
    with pytest.raises(ValueError):
        port.encode_field(_oracle.FIELD_VELC, np.arange(30, dtype=np.float32), 1e-9, 0.0)


def test_empty_field():
    payload = port.encode_field(_oracle.FIELD_POSN, np.zeros(0, np.float32), 1e-3, 1.0)
    assert payload.size == 64
    assert port.decode_field(payload).size == 0


def test_dither_streams_follow_rand_seed():
    # value j of block g uses draw (g%32)*32 + j%32 of stream (g/32)*32 + j/32
    n = 40 * 1024
    data = np.tile(np.linspace(0, 50, n, dtype=np.float32), 3)
    payload = port.encode_field(_oracle.FIELD_VELC, data, 1e-2, 0.0, seed=5)
    out, q = port.decode_field(payload, want_q=True)
    hd = _oracle.parse_header(payload)
    md = np.float32(max(hd["X1"][c] - hd["X0"][c] for c in range(3)))
    ns = port.lib.orc_diff_nstreams(n)
    assert ns == ((3 * 40 + 31) // 32) * 32
    st = port.rand_seed(5, ns)
    g, j = 37, 700  # block 37 of component 0 -> task 1, lane 21, draw 5*32+28
    s = st[(g // 32) * 32 + j // 32].copy()
    for _ in range((g % 32) * 32 + j % 32):
        port.lib.orc_rand_u64(s)
    bw = np.float32(md / np.float32(1 << int(hd["Depth"])))
    r = np.float32(port.lib.orc_rand_float(s))
    expect = np.float32(np.float32(hd["X0"][0] + np.float32(bw * np.float32(q[g * 1024 + j]))) + np.float32(r * bw))
    assert out[g * 1024 + j] == expect
