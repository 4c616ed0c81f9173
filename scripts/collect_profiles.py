#!/usr/bin/env python
"""Copy one gpu_round.sh run from gpurun_out/ into profiles/ (tracked):
bench lines, ncu launch list, full-size DRAM traffic (-> profiles/traffic.json,
which bench.py reads for roofline.traffic), ncu --set full summary.

usage: python scripts/collect_profiles.py r01c c     # TAG, suffix of the files under profiles/
"""
import collections
import csv
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag, suf = sys.argv[1], sys.argv[2]
rnd = tag[:3]
go = os.path.join(ROOT, "gpurun_out")
pr = os.path.join(ROOT, "profiles")


def strip(src, dst):
    with open(src) as fh, open(dst, "w") as out:
        out.writelines(l for l in fh if not l.startswith("=="))


shutil.copy(os.path.join(go, "bench_%s.json" % tag), os.path.join(pr, "%s_bench_%s.json" % (rnd, suf)))
if os.path.exists(os.path.join(go, "bench_ref_%s.json" % tag)):
    shutil.copy(os.path.join(go, "bench_ref_%s.json" % tag), os.path.join(pr, "%s_bench_ref_%s.json" % (rnd, suf)))
strip(os.path.join(go, "launches_%s.csv" % tag), os.path.join(pr, "%s_launches_%s.csv" % (rnd, suf)))
tcsv = os.path.join(pr, "%s_traffic_%s.csv" % (rnd, suf))
strip(os.path.join(go, "traffic_%s.csv" % tag), tcsv)

rows = list(csv.reader(open(tcsv)))
h = rows[0]
ik, im, iv, iid = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("ID")
byid = collections.defaultdict(dict)
for r in rows[1:]:
    byid[(r[iid], r[ik])][r[im]] = float(r[iv])
per = collections.defaultdict(lambda: collections.defaultdict(list))
for (_, k), m in byid.items():
    name = "k_encode" if "k_encode" in k else "k_unpack_dequant"
    per[name]["read"].append(m["dram__bytes_read.sum"])
    per[name]["write"].append(m["dram__bytes_write.sum"])
    per[name]["ns"].append(m["gpu__time_duration.sum"])
out, detail = {}, {}
for k, v in per.items():
    detail[k] = {kk: sum(vv) / len(vv) for kk, vv in v.items()}
    out[k] = detail[k]["read"] + detail[k]["write"]
out["_source"] = ("profiles/%s_traffic_%s.csv: ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum on bench.py "
                  "--steps 2 --warmup 1 --no-cpu --no-e2e (full 1024^3 pos+vel workload, 6442450944 values per "
                  "launch), mean over the captured launches" % (rnd, suf))
out["_detail"] = detail
json.dump(out, open(os.path.join(pr, "traffic.json"), "w"), indent=1)

rep = os.path.join(go, "prof_%s.ncu-rep" % tag)
if os.path.exists(rep):
    note = ("ncu --set full --clock-control none, bench.py --segments 16 --steps 1 --warmup 1 --no-cpu --no-e2e "
            "(16 segments = 402653184 values per launch)")
    with open(os.path.join(pr, "%s_ncu_full_%s.json" % (rnd, suf)), "w") as fh:
        subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "ncu_summary.py"), rep, note], stdout=fh)
print(json.dumps(out, indent=1))
