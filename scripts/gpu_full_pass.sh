#!/bin/bash
# Full 1-GPU measurement pass: tests, smoke, both bench arms, FP32 line, launch list, the two full
# ncu captures, N = 600, agreement scripts.  Outputs under gpurun_out/.
bash scripts/gpu_check.sh
timeout 600 python bench.py --dtype f32 --no-cpu > gpurun_out/bench_f32.log 2>&1; echo "bench f32 rc=$?"; grep "^{" gpurun_out/bench_f32.log | tail -1 | cut -c1-250
bash scripts/gpu_ncu_bench.sh
bash scripts/gpu_ncu.sh
timeout 300 python scripts/long_horizon.py > gpurun_out/long_horizon.log 2>&1; echo "n600 rc=$?"; tail -4 gpurun_out/long_horizon.log
timeout 300 python scripts/team_time.py > gpurun_out/small_batches.log 2>&1; tail -9 gpurun_out/small_batches.log
timeout 600 python scripts/monte_carlo.py --trajectories 131072 --steps 200 > gpurun_out/mc1.log 2>&1; grep "^{" gpurun_out/mc1.log | tail -1 | cut -c1-400
timeout 600 python scripts/monte_carlo.py --trajectories 131072 --steps 200 --precision f32 > gpurun_out/mc1_f32.log 2>&1; grep "^{" gpurun_out/mc1_f32.log | tail -1 | cut -c1-300
timeout 600 python scripts/monte_carlo.py --trajectories 65536 --steps 200 --rotors-plant > gpurun_out/mc1_rotors.log 2>&1; grep "^{" gpurun_out/mc1_rotors.log | tail -1 | cut -c1-500
timeout 600 python scripts/monte_carlo.py --trajectories 65536 --steps 100 --figure8 --motor-lag > gpurun_out/mc1_f8.log 2>&1; grep "^{" gpurun_out/mc1_f8.log | tail -1 | cut -c1-300
