#!/bin/bash
# decoder variants: short bench per MNWCUDA_DEC_RNG value, then the GPU parity suite under each
mkdir -p gpurun_out
VARS="${VARS:-MNWCUDA_DEC_RNG=1;MNWCUDA_DEC_RNG=2}" STEPS=${STEPS:-5} bash scripts/gpu_sweep.sh 2>&1 | tee gpurun_out/dec_sweep.log
for v in ${TEST_RNG:-1 2}; do
  echo "== pytest -m gpu with MNWCUDA_DEC_RNG=$v"
  MNWCUDA_DEC_RNG=$v timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee -a gpurun_out/dec_tests_$v.log
done
