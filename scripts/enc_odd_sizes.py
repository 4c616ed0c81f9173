#!/usr/bin/env python
"""k_encode / k_unpack_dequant on segments whose size is not 2^22: equal-sized fields whose
components are not whole reduce windows (the general reduce), next to the bench's shape.
usage: python scripts/enc_odd_sizes.py [particles per field] [fields]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
import minnow_c_b200 as mc

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4000000
nf = int(sys.argv[2]) if len(sys.argv) > 2 else 128
codec = mc.Codec(0)
d_in = torch.empty(nf * 3 * n, dtype=torch.float32, device="cuda")
fields = []
for i in range(0, nf, 2):
    po, vo = i * 3 * n, (i + 1) * 3 * n
    codec.generate(1024, bench.BOX, bench.GEN_SEED, (i // 2) * n, n, d_in[po:po + 3 * n], d_in[vo:vo + 3 * n])
    fields.append((po, n, mc.FIELD_POSN, bench.POS_DELTA, bench.BOX, 100 + i))
    fields.append((vo, n, mc.FIELD_VELC, bench.VEL_DELTA, 0.0, 101 + i))
descs = mc.make_descs(fields)
d_out = torch.empty(nf * mc.field_bound(n), dtype=torch.uint8, device="cuda")
d_off = torch.zeros(nf + 1, dtype=torch.int64, device="cuda")
d_dec = torch.empty_like(d_in)
codec.enable_kernel_timing(True)
for rep in range(3):
    codec.encode_dev(d_in, descs, nf, d_out, d_off)
    torch.cuda.synchronize()
    te = codec.kernel_times()["k_encode"]
    codec.decode_dev(d_out, d_off, descs, nf, d_dec)
    torch.cuda.synchronize()
    td = codec.kernel_times()["k_unpack_dequant"]
st = codec.status()
vals = nf * 3 * n
c = int(d_off[-1]) / vals
peak = bench.peaks()[0]
err = float(((d_dec - d_in).abs()[: 3 * n]).max())
print("n = %d x %d fields (%.2f GB raw, %d blocks per component): k_encode %.3f ms (%.3f of peak), k_unpack_dequant %.3f ms (%.3f), "
      "c = %.4f B/value, status %d, max err of field 0 %.2e"
      % (n, nf, vals * 4 / 1e9, (n + 1023) // 1024, te, (4 + c) * vals / (te * 1e-3) / 1e9 / peak, td,
         (4 + c) * vals / (td * 1e-3) / 1e9 / peak, c, st, err))
codec.close()
