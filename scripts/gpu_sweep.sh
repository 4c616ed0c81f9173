#!/bin/bash
# bench.py encode/decode times for a list of environment settings: VARS="A=1 B=2;A=3"
mkdir -p gpurun_out
IFS=';' read -ra SETS <<< "$VARS"
for s in "${SETS[@]}"; do
  env $s timeout 120 python bench.py --steps ${STEPS:-5} --warmup 2 --no-cpu --no-e2e ${EXTRA} > gpurun_out/sweep.json 2> gpurun_out/sweep.err || tail -3 gpurun_out/sweep.err
  python - "$s" <<'PY'
import json,sys
try:
    d=json.loads(open("gpurun_out/sweep.json").read().strip().splitlines()[-1])
    print(sys.argv[1], "| enc %.2f ms dec %.2f ms | frac" % (d["encode_ms"], d["decode_ms"]), {k: v["frac"] for k, v in d["roofline"]["kernels"].items()})
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
grep "encode stats" gpurun_out/sweep.err | tail -1
done
