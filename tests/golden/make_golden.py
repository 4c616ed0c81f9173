"""Generates tests/golden/diff_v0_9.npz from the REFERENCE's own compiled
primitives (oracle/_ref/liboracle_ref.so = the Diff v0.9 composition running on
/root/reference/src/{util,rand,quant}.c built by oracle/Makefile). Run in the
build container (needs /root/reference); the .npz travels with the repo so the
GPU box, which has no /root/reference, checks against reference-made vectors.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import _oracle  # noqa: E402
from _data import uniform_component, zeldovich_host  # noqa: E402

POSN, VELC = 0x506F736E, 0x56656C63


def cases():
    pos, vel = zeldovich_host(16)
    yield "zel16_pos", POSN, pos.reshape(-1), 1e-3, 100.0, 7
    yield "zel16_vel", VELC, vel.reshape(-1), 1.0, 0.0, 7
    yield "uniform_4099", POSN, uniform_component(4099, seed=11, L=10.0), 10.0 / 2**15 * 1.0001, 10.0, 1
    pos, vel = zeldovich_host(16, start=37, count=1500)
    yield "ragged_vel", VELC, vel.reshape(-1), 0.37, 0.0, 123456789
    rng = np.random.default_rng(2)
    yield "wide_vel", VELC, (rng.standard_normal(3 * 2500) * 300).astype(np.float32), 1e-3, 0.0, 5
    yield "const", VELC, np.full(3 * 1100, -1.25, np.float32), 1e-2, 0.0, 0


def main():
    ref = _oracle.reference()
    if ref is None:
        raise SystemExit("oracle/_ref is not built (needs /root/reference)")
    out = {}
    names = []
    for name, code, data, delta, width, seed in cases():
        payload = ref.encode_field(code, data, delta, width, seed=seed)
        dec = ref.decode_field(payload)
        names.append(name)
        out[name + "/in"] = data.astype(np.float32)
        out[name + "/meta"] = np.array([code, seed], np.uint64)
        out[name + "/acc"] = np.array([delta, width], np.float32)
        out[name + "/payload"] = payload
        out[name + "/decoded_bits"] = dec.view(np.uint32)
    out["names"] = np.array(names)
    path = os.path.join(HERE, "diff_v0_9.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes,", len(names), "cases")


if __name__ == "__main__":
    main()
