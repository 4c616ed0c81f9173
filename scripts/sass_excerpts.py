#!/usr/bin/env python
"""SASS excerpts of the two hot kernels for profiles/ (static view of the built library).
  python scripts/sass_excerpts.py r02   ->  profiles/r02_sass_k_encode.txt, profiles/r02_sass_k_unpack_dequant.txt"""
import collections, os, re, subprocess, sys, tempfile
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rnd = sys.argv[1] if len(sys.argv) > 1 else "r02"
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(root, "minnow_c_b200", "libminnow_cuda.so")], cwd=tmp, capture_output=True)
cubin = [os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-c", cubin], capture_output=True, text=True).stdout.splitlines()


def func(sym):
    start = next(i for i, l in enumerate(dis) if l.startswith(".text.") and sym in l)
    out = []
    for l in dis[start + 1:]:
        if l.startswith(".text.") or l.startswith("//--------------------- ."):
            if out:
                break
        m = re.match(r"\s*/\*([0-9a-f]+)\*/\s+(.*?);", l)
        if m:
            out.append("        /*%s*/  %s ;" % (m.group(1), m.group(2).rstrip()))
    return out


def write(sym, title, path, specs):
    L = func(sym)
    ops = collections.Counter()
    for l in L:
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_]+)", l)
        if m:
            ops[m.group(2)] += 1
    out = ["%s (sm_100a), %d SASS instructions = %d bytes of code; nvdisasm of minnow_c_b200/libminnow_cuda.so" % (title, len(L), 16 * len(L)),
           "opcode histogram (static): " + ", ".join("%s %d" % (k, v) for k, v in ops.most_common(45)), ""]
    for pat, t, ctx, maxn in specs:
        out.append("== " + t)
        n, last = 0, -100
        for i, l in enumerate(L):
            if re.search(pat, l) and i - last > ctx * 2:
                out.extend(L[max(0, i - ctx):i + ctx + 1])
                out.append("   ...")
                last = i
                n += 1
                if n >= maxn:
                    break
        out.append("")
    open(path, "w").write("\n".join(out) + "\n")
    print(out[0])


write("k_encodeILb0ELb0ELb0", "k_encode<false,false>", os.path.join(root, "profiles", rnd + "_sass_k_encode.txt"), [
    (r"UTMALDG", "TMA loads of the issuer warp (cp.async.bulk.tensor.2d -> UTMALDG.2D; the second form carries the evict_first L2 policy of the B reads), each after the stage's mbarrier arrive.expect_tx (SYNCS.ARRIVE.TRANS64)", 7, 2),
    (r"SYNCS\.ARRIVE", "mbarrier arrive (stage handed back / expect_tx)", 3, 2),
    (r"SYNCS\.PHASECHK", "mbarrier try_wait (SYNCS.PHASECHK.TRANS64.TRYWAIT): issuer on `empty`, consumers on `full`", 3, 2),
    (r"ATOMS", "stage claim: elect.sync + atom.shared.add", 6, 1),
    (r"FMNMX3", "three-input min/max: A work extrema, lane-class unwrap, delta extrema on the FMA pipe", 4, 2),
    (r"F2F\.F64\.F32", "the exact quotient: F2F.F64.F32 / DMUL / F2F.F32.F64, then FMNMX (clamp) + FADD.RZ (biasing add, round toward zero)", 10, 1),
    (r"REDUX", "warp reductions of the block's extrema", 3, 1),
])
write("k_unpack_dequantILb0", "k_unpack_dequant<false>", os.path.join(root, "profiles", rnd + "_sass_k_unpack_dequant.txt"), [
    (r"LDG\.E\.128", "128-bit coalesced loads of the packed block", 3, 1),
    (r"SHF\.[LR]\.W", "xorshift128+ step on 32-bit halves (funnel shifts + LOP3), then the dither product and the de-quantisation on the FMA pipe", 12, 1),
    (r"I2FP\.F32\.U32", "dither: I2FP of the draw, FMUL by bw * 2^-32, FADD to the bin edge", 6, 1),
    (r"STG\.E\.128", "128-bit coalesced stores of the decoded floats", 3, 1),
    (r"SHFL\.UP", "warp scan of the lane totals (FP32, exact below 2^24)", 4, 1),
])
