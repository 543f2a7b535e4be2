#!/bin/bash
# DRAM bytes + L2 hit rate of the bench launch for a set of environment settings:
#   gpu_ncu_dram.sh "ENV=.. ENV=.." ...    (an empty string = defaults)
mkdir -p gpurun_out
BCMD="python bench.py --steps 3 --warmup 3 --no-cpu --no-extras"
i=0
for envs in "$@"; do
  i=$((i+1))
  echo "== [$i] $envs" | tee -a gpurun_out/ncu_dram.log
  env $envs timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,gpu__time_duration.sum --clock-control none -k regex:nmpc_solve -s 4 -c 1 --csv --log-file gpurun_out/ncu_dram_$i.csv $BCMD > /dev/null 2>&1
  grep -E "dram__|lts__|gpu__time" gpurun_out/ncu_dram_$i.csv | awk -F'","' '{print $(NF-2), $(NF-1), $NF}' | tee -a gpurun_out/ncu_dram.log
done
