#!/bin/bash
# DRAM traffic and L2 hit rate of k_encode under a list of environment settings (full-size bench, light ncu pass)
mkdir -p gpurun_out
IFS=';' read -ra SETS <<< "${VARS:-X=0}"
for v in "${SETS[@]}"; do
  env $v timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none \
    -k regex:"^k_encode" -s 1 -c 1 --csv --log-file gpurun_out/l2.csv python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e > gpurun_out/l2.log 2>&1
  echo "[$v]"; grep -v "^==" gpurun_out/l2.csv | awk -F'","' 'NR>1{print "   " $(NF-2), $(NF)}' | tr -d '"'
done
