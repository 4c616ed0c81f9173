#!/bin/bash
# ncu --set full of the two hot kernels on a 16-segment slab (after a clean plain run of the same command)
mkdir -p gpurun_out
CMD="python bench.py --segments 16 --steps 1 --warmup 1 --no-cpu --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"k_encode|k_unpack_dequant" -s 2 -c 2 -f -o gpurun_out/prof $CMD > gpurun_out/ncu.log 2>&1
echo "ncu exit $?"; tail -5 gpurun_out/ncu.log
