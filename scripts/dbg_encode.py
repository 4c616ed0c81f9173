import faulthandler, sys, os
faulthandler.dump_traceback_later(25, exit=True)
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
print("import", flush=True)
import torch
import minnow_c_b200 as mc
import _oracle
from _data import zeldovich_host
print("create", flush=True)
c = mc.Codec(0)
print("created", flush=True)
pos, vel = zeldovich_host(int(os.environ.get("NG", "32")))
orc = _oracle.port()
for code, data, delta, width in ((mc.FIELD_VELC, vel.reshape(-1), 1.0, 0.0), (mc.FIELD_POSN, pos.reshape(-1), 1e-3, 100.0)):
    print("encode", hex(code), flush=True)
    got = c.encode_field(code, data, delta, width)
    print("encoded", got.size, flush=True)
    want = orc.encode_field(code, data, delta, width, seed=0)
    print("match", np.array_equal(got, want), got.size, want.size, flush=True)
    if not np.array_equal(got, want):
        n = min(got.size, want.size)
        bad = np.nonzero(got[:n] != want[:n])[0]
        print("first diffs", bad[:10], "count", bad.size, flush=True)
        tg = _oracle.parse_table(got); tw = _oracle.parse_table(want)
        for k in list(range(0, min(6, tg.size))) + list(range(max(0, tg.size - 3), tg.size)):
            print(k, "got", tg[k], "want", tw[k], flush=True)
        nbad = sum(1 for k in range(tg.size) if tg[k] != tw[k])
        print("table entries differing:", nbad, "of", tg.size, flush=True)
print("done", flush=True)
