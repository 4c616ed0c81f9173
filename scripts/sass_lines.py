#!/usr/bin/env python
"""SASS of one kernel grouped by source line (static view, no GPU needed).
  python scripts/sass_lines.py MANGLED_SUBSTR [file.cuh:line ...]   (no lines: per-file static counts)"""
import re, subprocess, sys, collections, os, tempfile
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(root, "minnow_c_b200", "libminnow_cuda.so")
sym = sys.argv[1]
want = set(sys.argv[2:])
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=tmp, capture_output=True)
cubin = [os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
start = next(i for i, l in enumerate(dis) if l.startswith(".text.") and sym in l and l.rstrip().endswith(":"))
cur = ("?", 0); n = 0
counts = collections.Counter(); ops = collections.Counter()
for l in dis[start + 1:]:
    if (l.startswith("//--------------------- .") or l.startswith("\t.section")) and n: break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    m = re.match(r"\s*/\*([0-9a-f]+)\*/\s+(.*?);", l)
    if m:
        n += 1; counts[cur[0]] += 1
        key = "%s:%d" % cur
        if key in want: print(key, m.group(1), m.group(2))
if not want:
    print("static instructions:", n)
    for f, c in counts.most_common(): print("  %-28s %d" % (f, c))
