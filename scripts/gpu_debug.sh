#!/bin/bash
mkdir -p gpurun_out
timeout 60 python scripts/dbg_encode.py > gpurun_out/dbg1.log 2>&1; echo "plain exit $?" >> gpurun_out/dbg1.log
grep -v "  tile" gpurun_out/dbg1.log | tail -20
if [ -n "$TESTS" ]; then
timeout 200 python -m pytest tests -m gpu -x -q > gpurun_out/dbg_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/dbg_pytest.log
tail -30 gpurun_out/dbg_pytest.log
fi
if [ -n "$BENCH" ]; then
timeout 200 python bench.py --steps 5 --warmup 2 --no-cpu --no-e2e > gpurun_out/dbg_bench.json 2> gpurun_out/dbg_bench.err; echo "bench exit $?"
python - <<'PY'
import json
try:
    d=json.loads(open("gpurun_out/dbg_bench.json").read().strip().splitlines()[-1])
    print("value %.0f enc %.0f dec %.0f GB/s | enc %.2f ms dec %.2f ms" % (d["value"], d["encode_gbs"], d["decode_gbs"], d["encode_ms"], d["decode_ms"]), d["kernel_ms"], {k: v["frac"] for k, v in d["roofline"]["kernels"].items()})
except Exception as e:
    print("bench unreadable", e)
PY
tail -5 gpurun_out/dbg_bench.err
fi
