#!/bin/bash
# first GPU call: parity tests, smoke, small then full bench
mkdir -p gpurun_out
nvidia-smi > gpurun_out/nvidia_smi.txt 2>&1
nproc > gpurun_out/nproc.txt; free -g >> gpurun_out/nproc.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
tail -3 gpurun_out/smoke.log
timeout 600 python bench.py --segments 16 --e2e-segments 4 --steps 3 --warmup 1 --no-cpu > gpurun_out/bench_small.json 2> gpurun_out/bench_small.err; echo "bench small exit $?"
cat gpurun_out/bench_small.json; tail -5 gpurun_out/bench_small.err
timeout 1200 python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench full exit $?"
cat gpurun_out/bench_full.json; tail -5 gpurun_out/bench_full.err
