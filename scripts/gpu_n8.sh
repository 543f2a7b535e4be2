#!/bin/bash
# One multi-GPU gpurun call (gpurun --gpus N): the bench at N GPUs (both arms), configs[2] at full
# size (1M trajectories x 200 steps) and the RotorS-plant Monte Carlo.
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus $N > gpurun_out/bench_n$N.log 2>&1; echo "bench n$N rc=$?"; grep "^{" gpurun_out/bench_n$N.log | tail -1 | cut -c1-300
timeout 600 $TR bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/bench_ref_n$N.log 2>&1; echo "ref rc=$?"; grep "^{" gpurun_out/bench_ref_n$N.log | tail -1 | cut -c1-200
timeout 600 $TR scripts/monte_carlo.py --trajectories 1000000 --steps 200 > gpurun_out/mc_n$N.log 2>&1; echo "mc rc=$?"; grep "^{" gpurun_out/mc_n$N.log | tail -1 | cut -c1-700
timeout 600 $TR scripts/monte_carlo.py --trajectories 524288 --steps 200 --rotors-plant > gpurun_out/mc_rotors_n$N.log 2>&1; echo "mc rotors rc=$?"; grep "^{" gpurun_out/mc_rotors_n$N.log | tail -1 | cut -c1-700
timeout 300 python -m pytest tests/test_gpu_sharding.py -q 2>&1 | tail -2
