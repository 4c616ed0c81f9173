#!/bin/bash
# round 2 GPU call: TESTS=1 pytest -m gpu; BENCH=1 full bench; QUICK=1 kernel-only bench;
# NCU=regex source-level capture on a 16-segment slab; REF=1 reference arm
mkdir -p gpurun_out
TAG=${TAG:-r02}
if [ -n "$TESTS" ]; then
  timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_${TAG}.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_${TAG}.log
  tail -8 gpurun_out/pytest_${TAG}.log
fi
if [ -n "$QUICK" ]; then
  timeout 300 python bench.py --steps 5 --warmup 2 --no-cpu --no-e2e --no-sweep $QARGS > gpurun_out/quick_${TAG}.json 2> gpurun_out/quick_${TAG}.err; echo "quick exit $?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/quick_${TAG}.json").read().strip().splitlines()[-1])
    print("enc %.3f ms dec %.3f ms" % (d["encode_ms"], d["decode_ms"]), d["kernel_ms"], {k: v["frac"] for k, v in d["roofline"]["kernels"].items()}, d["check"])
except Exception as e:
    print("unreadable", e); print(open("gpurun_out/quick_${TAG}.err").read()[-2000:])
PY
fi
if [ -n "$STATS" ]; then
  MNWCUDA_ENCODE_STATS=1 timeout 300 python bench.py --steps 3 --warmup 1 --no-cpu --no-e2e --no-sweep > /dev/null 2> gpurun_out/stats_${TAG}.err; grep "encode stats" gpurun_out/stats_${TAG}.err | tail -7
fi
if [ -n "$BENCH" ]; then
  timeout 1500 python bench.py > gpurun_out/bench_${TAG}.json 2> gpurun_out/bench_${TAG}.err; echo "bench exit $?"
  head -c 6000 gpurun_out/bench_${TAG}.json; tail -7 gpurun_out/bench_${TAG}.err
fi
if [ -n "$REF" ]; then
  timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_${TAG}.json 2> gpurun_out/bench_ref_${TAG}.err; echo "ref exit $?"
  cat gpurun_out/bench_ref_${TAG}.json; tail -3 gpurun_out/bench_ref_${TAG}.err
fi
if [ -n "$NCU" ]; then
  CMD2="python bench.py --segments 16 --steps 1 --warmup 1 --no-cpu --no-e2e --no-sweep"
  timeout 300 $CMD2 > gpurun_out/plain_full.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$NCU" -s ${SKIP:-1} -c ${COUNT:-1} -f \
    -o gpurun_out/prof_${TAG} $CMD2 > gpurun_out/ncu_full.log 2>&1
  echo "ncu full exit $?"; tail -2 gpurun_out/ncu_full.log
fi
