#!/bin/bash
# round 2, first call: GPU tests, baseline bench, source-level capture of k_encode
mkdir -p gpurun_out
nvidia-smi > gpurun_out/nvidia_smi.txt 2>&1
nproc > gpurun_out/nproc.txt; free -g >> gpurun_out/nproc.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --no-cpu --no-e2e > gpurun_out/bench_r02a.json 2> gpurun_out/bench_r02a.err; echo "bench exit $?"
cat gpurun_out/bench_r02a.json | head -c 3000
MNWCUDA_ENCODE_STATS=1 timeout 300 python bench.py --steps 3 --warmup 1 --no-cpu --no-e2e > /dev/null 2> gpurun_out/stats.err; grep "encode stats" gpurun_out/stats.err | tail -2
CMD2="python bench.py --segments 16 --steps 1 --warmup 1 --no-cpu --no-e2e"
timeout 300 $CMD2 > gpurun_out/plain_full.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_encode" -s 1 -c 1 -f \
  -o gpurun_out/prof_r02a $CMD2 > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?"; tail -3 gpurun_out/ncu_full.log; ls -la gpurun_out/
