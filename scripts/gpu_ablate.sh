#!/bin/bash
mkdir -p gpurun_out
for m in ${MASKS:-0 1 2 3 4 7 15}; do
  echo -n "mask $m: "
  MNWCUDA_DEBUG_MASK=$m timeout 300 python bench.py --steps 5 --warmup 2 --encode-only 2>&1 | grep encode_only_ms
done
