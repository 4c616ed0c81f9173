#!/usr/bin/env python
"""k_encode's two halves on their own (timing experiment): A work alone (min/max
through the ring), B work alone (quantise .. place, with the parameters of the
previous full call), next to the fused launch. usage: python scripts/enc_only.py [segments]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
import minnow_c_b200 as mc

n_seg = int(sys.argv[1]) if len(sys.argv) > 1 else 256
codec = mc.Codec(0)
fields = bench.field_list(n_seg, bench.SEG)
nf = len(fields)
descs = mc.make_descs(fields)
d_in = torch.empty(nf * 3 * bench.SEG, dtype=torch.float32, device="cuda")
bench.generate(codec, d_in, 1024, 0, n_seg, bench.SEG)
d_out = torch.empty(nf * mc.field_bound(bench.SEG), dtype=torch.uint8, device="cuda")
d_off = torch.zeros(nf + 1, dtype=torch.int64, device="cuda")
codec.enable_kernel_timing(True)


def run(tag, reps=4):
    ms = []
    for _ in range(reps):
        codec.encode_dev(d_in, descs, nf, d_out, d_off)
        torch.cuda.synchronize()
        ms.append(codec.kernel_times()["k_encode"])
    st = codec.status()
    print("%-8s k_encode %s ms (status %d)" % (tag, " ".join("%.3f" % m for m in ms), st), flush=True)


run("fused")
os.environ["MNWCUDA_ENCODE_ONLY"] = "B"
run("B only")
del os.environ["MNWCUDA_ENCODE_ONLY"]
run("fused")
os.environ["MNWCUDA_DBG"] = "64"
run("A only")
