#!/bin/bash
# ncu --set full of the hot kernels on a 16-segment slab (after a clean plain run of the same command)
mkdir -p gpurun_out
TAG=${TAG:-x}
KRE=${KRE:-k_encode|k_unpack_dequant}
CMD="python bench.py --segments ${SEGS:-16} --steps 1 --warmup 1 --no-cpu --no-e2e"
timeout 200 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"$KRE" -s ${SKIP:-2} -c ${COUNT:-2} -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/ncu.log
