#!/usr/bin/env python
"""Per-source-line instruction counts of one kernel from an ncu report.

  python scripts/ncu_lines.py REPORT.ncu-rep CUBIN MANGLED_SUBSTR [top]

ncu's csv source page lists per-SASS-instruction metrics without line numbers;
nvdisasm -g lists the same instructions with //## File/line markers. The two are
joined by instruction offset. Prints executed warp instructions and stall samples
per (file, line), innermost inlined location."""
import csv, re, subprocess, sys, collections

rep, cubin, sym = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 60
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
start = next(i for i, l in enumerate(dis) if l.startswith(".text.") and sym in l and l.rstrip().endswith(":"))
loc = {}
cur = ("?", 0)
for l in dis[start + 1:]:
    if l.startswith("//--------------------- .") or l.startswith("\t.section"):
        if loc: break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]+)\*/\s+(.*?);", l)
    if m:
        loc[int(m.group(1), 16)] = (cur, m.group(2))
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
ix = {n: hdr.index(n) for n in ("Address", "Source", "# Samples", "Instructions Executed")}
body = []
for r in rows[hi + 1:]:  # the first kernel of the report only
    if r and r[0] == "Address":
        break
    if len(r) == len(hdr):
        body.append(r)
base = int(body[0][ix["Address"]], 16)
agg = collections.defaultdict(lambda: [0, 0])
tot = [0, 0]
ops = collections.Counter()
for r in body:
    off = int(r[ix["Address"]], 16) - base
    (f, ln), text = loc.get(off, (("?", 0), ""))
    n, s = int(r[ix["Instructions Executed"]] or 0), int(r[ix["# Samples"]] or 0)
    agg[(f, ln)][0] += n; agg[(f, ln)][1] += s
    tot[0] += n; tot[1] += s
    ops[r[ix["Source"]].split()[0] if r[ix["Source"]].split()[0][0] != "@" else r[ix["Source"]].split()[1]] += n
print("total warp instructions %d, samples %d" % tuple(tot))
byfile = collections.defaultdict(lambda: [0, 0])
for (f, ln), (n, s) in agg.items():
    byfile[f][0] += n; byfile[f][1] += s
for f, (n, s) in sorted(byfile.items(), key=lambda kv: -kv[1][0]):
    print("%-24s inst %5.1f%%  samples %5.1f%%" % (f, 100.0 * n / tot[0], 100.0 * s / max(tot[1], 1)))
print("--- top lines")
for (f, ln), (n, s) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%-22s %4d  inst %5.2f%%  samples %5.2f%%" % (f, ln, 100.0 * n / tot[0], 100.0 * s / max(tot[1], 1)))
print("--- top opcodes")
for o, n in ops.most_common(25):
    print("%-20s %5.2f%%" % (o, 100.0 * n / tot[0]))
