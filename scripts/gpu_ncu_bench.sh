#!/bin/bash
# one full ncu capture of the solve kernel inside the bench command (after a plain run of it)
mkdir -p gpurun_out
BCMD="python bench.py --steps 3 --warmup 3 --no-cpu --no-extras"
timeout 300 $BCMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:nmpc_solve -s 4 -c 1 -o gpurun_out/prof_bench -f $BCMD > gpurun_out/ncu_full_bench.log 2>&1
echo "ncu full (bench kernel) rc=$?"; tail -2 gpurun_out/ncu_full_bench.log
