#!/bin/bash
# one GPU call: parity tests, smoke, full bench, reference arm, ncu launch list,
# ncu --set full of the hot kernels (each ncu pass only after the same command
# exited 0 without ncu). Outputs land in gpurun_out/.
mkdir -p gpurun_out
TAG=${TAG:-r01}
nvidia-smi > gpurun_out/nvidia_smi.txt 2>&1
nproc > gpurun_out/nproc.txt; free -g >> gpurun_out/nproc.txt
if [ -z "$SKIP_TESTS" ]; then
  timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "alone exit $?" >> gpurun_out/pytest_noalloc_alone.log
  tail -4 gpurun_out/pytest_noalloc_alone.log
  timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
  tail -6 gpurun_out/pytest_gpu.log
  timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
  tail -2 gpurun_out/smoke.log
fi
if [ -z "$SKIP_BENCH" ]; then
  timeout 1200 python bench.py > gpurun_out/bench_${TAG}.json 2> gpurun_out/bench_${TAG}.err; echo "bench exit $?"
  cat gpurun_out/bench_${TAG}.json; tail -5 gpurun_out/bench_${TAG}.err
fi
if [ -n "$REF_ARM" ]; then
  timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_${TAG}.json 2> gpurun_out/bench_ref_${TAG}.err; echo "ref exit $?"
  cat gpurun_out/bench_ref_${TAG}.json
fi
if [ -z "$SKIP_NCU" ]; then
  # launch list of the full-size bench command
  CMD="python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-sweep"
  timeout 600 $CMD > gpurun_out/plain_launch.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"^k_" -c 80 --csv \
    --log-file gpurun_out/launches_${TAG}.csv $CMD > gpurun_out/ncu_launch.log 2>&1
  echo "ncu launch list exit $?"; tail -3 gpurun_out/ncu_launch.log
  # DRAM traffic of the two hot kernels at full size (one light pass)
  timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
    -k regex:"^k_(encode|unpack_dequant)" -s 2 -c 4 --csv --log-file gpurun_out/traffic_${TAG}.csv $CMD > gpurun_out/ncu_traffic.log 2>&1
  echo "ncu traffic exit $?"; tail -2 gpurun_out/ncu_traffic.log
  # full capture of the two hot kernels on a 16-segment slab
  CMD2="python bench.py --segments 16 --steps 1 --warmup 1 --no-cpu --no-e2e --no-sweep"
  timeout 300 $CMD2 > gpurun_out/plain_full.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_encode|k_unpack_dequant" -s 2 -c 2 -f \
    -o gpurun_out/prof_${TAG} $CMD2 > gpurun_out/ncu_full.log 2>&1
  echo "ncu full exit $?"; tail -3 gpurun_out/ncu_full.log
fi
