#!/usr/bin/env python
"""Many small fields of random sizes in one batch (the non-uniform paths at scale):
encode, decode, round-trip bound on everything, byte parity of a sample against the oracle."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
import minnow_c_b200 as mc
import _oracle
nf = int(sys.argv[1]) if len(sys.argv) > 1 else 40000
rng = np.random.default_rng(7)
sizes = rng.integers(0, 5000, nf)
sizes[rng.integers(0, nf, 50)] = 0
sizes[rng.integers(0, nf, 20)] = 300000
total = int(sizes.sum())
codec = mc.Codec(0)
L = 100.0
d_in = torch.empty(3 * total, dtype=torch.float32, device="cuda")
# smooth-ish synthetic data: positions in [0, L) with a wrap now and then, velocities signed
g = torch.Generator(device="cuda"); g.manual_seed(11)
base = torch.cumsum(torch.randn(3 * total, device="cuda", generator=g) * 0.01, 0)
fields, off = [], 0
for i, n in enumerate(sizes.tolist()):
    code = mc.FIELD_POSN if i % 2 == 0 else mc.FIELD_VELC
    fields.append((off, n, code, 1e-3 if code == mc.FIELD_POSN else 1e-4, L if code == mc.FIELD_POSN else 0.0, 1000 + i))
    off += 3 * n
pos_mask = torch.zeros(3 * total, dtype=torch.bool, device="cuda")
for (o, n, code, *_r) in fields[::2]:
    pos_mask[o:o + 3 * n] = True
d_in.copy_(torch.where(pos_mask, torch.remainder(base + 50.0, L), base))
d_in[d_in >= L] = 0.0
descs = mc.make_descs(fields)
cap = sum(mc.field_bound(int(n)) for n in sizes)
d_out = torch.empty(cap, dtype=torch.uint8, device="cuda")
d_off = torch.zeros(nf + 1, dtype=torch.int64, device="cuda")
d_dec = torch.empty_like(d_in)
codec.encode_dev(d_in, descs, nf, d_out, d_off)  # (first call: scratch allocations)
codec.status()
codec.enable_kernel_timing(True)
torch.cuda.synchronize(); t0 = time.perf_counter()
codec.encode_dev(d_in, descs, nf, d_out, d_off)
st_e = codec.status(); t1 = time.perf_counter()
codec.decode_dev(d_out, d_off, descs, nf, d_dec)
st_d = codec.status(); t2 = time.perf_counter()
print("fields %d, particles %d | encode %.1f ms status %d | decode %.1f ms status %d | packed %.1f MB of %.1f MB"
      % (nf, total, 1e3 * (t1 - t0), st_e, 1e3 * (t2 - t1), st_d, int(d_off[-1]) / 1e6, 12 * total / 1e6))
print("kernel times (ms):", {k: round(v, 3) for k, v in codec.kernel_times().items()})
err = (d_dec - d_in).abs()
err = torch.minimum(err, (L - err).abs())
tol = torch.where(pos_mask, torch.tensor(1e-3, device="cuda"), torch.tensor(1e-4, device="cuda"))
worst = float((err / tol).max())
print("round trip: max |x - x'| / Delta = %.4f" % worst)
orc = _oracle.reference() or _oracle.port()
offs = d_off.cpu().numpy(); out = d_out.cpu().numpy(); h_in = d_in.cpu().numpy()
bad = 0
sample = list(rng.integers(0, nf, 200)) + [int(i) for i in np.nonzero(sizes == 300000)[0][:3]] + [int(i) for i in np.nonzero(sizes == 0)[0][:3]]
for i in sample:
    o, n, code, delta, width, seed = fields[i]
    want = orc.encode_field(code, h_in[o:o + 3 * n], delta, width, seed=seed)
    got = out[offs[i]:offs[i + 1]]
    bad += not np.array_equal(got, want)
print("byte parity against the oracle (%s): %d of %d fields differ" % (orc.flavour, bad, len(sample)))
codec.close()
sys.exit(0 if (st_e == 0 and st_d == 0 and worst <= 1.0 and bad == 0) else 1)
