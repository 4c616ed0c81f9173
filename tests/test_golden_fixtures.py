"""Committed golden vectors made by the reference's own compiled primitives
(tests/golden/make_golden.py): the restated oracle (CPU) and the CUDA path (GPU)
must both reproduce them bit for bit."""
import os

import numpy as np
import pytest

import _oracle

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "diff_v0_9.npz"))
NAMES = [str(n) for n in G["names"]]


def _case(name):
    code, seed = (int(v) for v in G[name + "/meta"])
    delta, width = (float(v) for v in G[name + "/acc"])
    return code, seed, np.float32(delta), np.float32(width), G[name + "/in"], G[name + "/payload"], G[name + "/decoded_bits"]


@pytest.mark.parametrize("name", NAMES)
def test_port_oracle_reproduces_reference_vectors(name):
    code, seed, delta, width, data, payload, bits = _case(name)
    orc = _oracle.port()
    assert np.array_equal(orc.encode_field(code, data, delta, width, seed=seed), payload)
    assert np.array_equal(orc.decode_field(payload).view(np.uint32), bits)


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_cuda_reproduces_reference_vectors(name):
    import minnow_c_b200 as mc
    code, seed, delta, width, data, payload, bits = _case(name)
    codec = mc.Codec(0)
    try:
        codec.set_dither_seed(seed)
        got = codec.encode_field(code, data, float(delta), float(width))
        assert np.array_equal(got, payload)
        dec, _ = codec.decode_field(code, data.size // 3, payload)
        assert np.array_equal(dec.view(np.uint32), bits)
    finally:
        codec.close()
