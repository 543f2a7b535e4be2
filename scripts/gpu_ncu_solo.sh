#!/bin/bash
# ncu full capture (source counters) of the solve kernel with one warp per scheduler busy
mkdir -p gpurun_out
B=${1:-148}
timeout 300 python scripts/prof_rti.py $B > gpurun_out/prof_solo_plain.log 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:nmpc_solve -s 1 -c 1 -o gpurun_out/prof_solo -f python scripts/prof_rti.py $B > gpurun_out/ncu_solo.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/prof_solo_plain.log
