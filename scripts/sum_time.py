#!/usr/bin/env python
"""k_block_headers on the bench slab: time and parity of a sample of blocks against a host checksum."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, bench
import minnow_c_b200 as mc
n_seg = int(sys.argv[1]) if len(sys.argv) > 1 else 256
codec = mc.Codec(0)
fields = bench.field_list(n_seg, bench.SEG); nf = len(fields); descs = mc.make_descs(fields)
d_in = torch.empty(nf * 3 * bench.SEG, dtype=torch.float32, device="cuda")
bench.generate(codec, d_in, 1024, 0, n_seg, bench.SEG)
d_out = torch.empty(nf * mc.field_bound(bench.SEG), dtype=torch.uint8, device="cuda")
d_off = torch.zeros(nf + 1, dtype=torch.int64, device="cuda")
codec.encode_dev(d_in, descs, nf, d_out, d_off); torch.cuda.synchronize()
comp = int(d_off[-1].item())
r = bench.run_checksums(torch, mc, codec, d_out, d_off, descs, nf, comp, bench.peaks()[0])
print(r)
