#!/bin/bash
# one full ncu capture of the solve kernel on the RTI aggressive batch (after a plain run)
mkdir -p gpurun_out
timeout 300 python scripts/prof_rti.py 16384 > gpurun_out/prof_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:nmpc_solve -s 1 -c 1 -o gpurun_out/prof_rti -f python scripts/prof_rti.py 16384 > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/prof_plain.log
