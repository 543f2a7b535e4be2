#!/bin/bash
mkdir -p gpurun_out
export NMPC_B200_LIB=$PWD/variants/lib_m3.so
for cap in 1 2; do
  NMPC_B200_MAX_CTAS_PER_SM=$cap timeout 300 python scripts/prof_rti.py 16384 > gpurun_out/prof_plain_cap$cap.log 2>&1 || exit 1
done
for cap in 1 2; do
  NMPC_B200_MAX_CTAS_PER_SM=$cap timeout 600 ncu --set full --clock-control none --import-source on -k regex:nmpc_solve -s 1 -c 1 -o gpurun_out/prof_cap$cap -f python scripts/prof_rti.py 16384 > gpurun_out/ncu_cap$cap.log 2>&1
  echo "cap $cap rc=$?"; cat gpurun_out/prof_plain_cap$cap.log
done
