#!/bin/bash
# One gpurun call: GPU parity tests, smoke, a short bench, launch list + full ncu capture.
set -o pipefail
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py --steps 20 --warmup 3 --extras > gpurun_out/bench.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench.log
timeout 300 python scripts/quick_time.py > gpurun_out/quick_time.log 2>&1; echo "quick rc=$?"; cat gpurun_out/quick_time.log
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
timeout 300 python scripts/prof_rti.py 16384 > gpurun_out/prof_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:nmpc_solve -s 1 -c 1 -o gpurun_out/prof_rti -f python scripts/prof_rti.py 16384 > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"; cat gpurun_out/prof_plain.log
