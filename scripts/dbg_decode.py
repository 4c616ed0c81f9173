import sys, os, numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import minnow_c_b200 as mc
z = np.load("tests/golden/diff_v0_9.npz")
for name in ("wide_vel", "const", "ragged_vel"):
    meta = z[name + "/meta"]; payload = z[name + "/payload"]; bits = z[name + "/decoded_bits"]; data = z[name + "/in"]
    codec = mc.Codec(0)
    dec, _ = codec.decode_field(int(meta[0]), data.size // 3, payload)
    d = dec.view(np.uint32)
    bad = np.nonzero(d != bits)[0]
    print(name, os.environ.get("MNWCUDA_DEC_RNG"), "mismatches", bad.size, bad[:10], (bad % (data.size // 3))[:10])
    if bad.size:
        i = bad[0]
        print("  got", dec[i], "want", bits[i:i+1].view(np.float32)[0], "diff", dec[i] - bits[i:i+1].view(np.float32)[0])
    codec.close()
