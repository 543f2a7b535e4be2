#!/bin/bash
# ncu full capture (with source counters) of the solve kernel on a small RTI batch: teams on / off
mkdir -p gpurun_out
B=${1:-148}
timeout 300 python scripts/prof_rti.py $B > gpurun_out/prof_small_plain.log 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:nmpc_solve -s 1 -c 1 -o gpurun_out/prof_small_teams -f python scripts/prof_rti.py $B > gpurun_out/ncu_small_teams.log 2>&1
echo "ncu teams rc=$?"
NMPC_B200_NO_TEAMS=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:nmpc_solve -s 1 -c 1 -o gpurun_out/prof_small_solo -f python scripts/prof_rti.py $B > gpurun_out/ncu_small_solo.log 2>&1
echo "ncu solo rc=$?"
cat gpurun_out/prof_small_plain.log
