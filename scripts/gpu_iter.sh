#!/bin/bash
# parity tests, then the bench with encode scheduling variants given as "interleave lag" pairs
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
short() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], "value %.0f enc %.0f dec %.0f GB/s | enc %.2f ms dec %.2f ms |" % (d["value"], d["encode_gbs"], d["decode_gbs"], d["encode_ms"], d["decode_ms"]), "frac", {k: v["frac"] for k, v in d["roofline"]["kernels"].items()})
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
}
for v in ${VARIANTS:-"1024:-1"}; do
  il=${v%%:*}; lag=${v##*:}
  tag="il${il}_lag${lag}"
  MNWCUDA_ENCODE_LAG2=$il MNWCUDA_ENCODE_LAG=$lag timeout 600 python bench.py --steps 5 --warmup 2 --no-cpu --no-e2e > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err || tail -3 gpurun_out/bench_$tag.err
  short gpurun_out/bench_$tag.json
done
if [ -n "$NOLOOK" ]; then
  MNWCUDA_DEBUG_NOLOOKBACK=1 timeout 600 python bench.py --steps 5 --warmup 2 --no-cpu --no-e2e > gpurun_out/bench_nolook.json 2> gpurun_out/bench_nolook.err || tail -3 gpurun_out/bench_nolook.err
  short gpurun_out/bench_nolook.json
fi
