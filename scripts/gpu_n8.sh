#!/bin/bash
# One 8-GPU gpurun call: the bench at N=8 and configs[2] at full size (1M trajectories x 200 steps).
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus 8 --no-extras > gpurun_out/bench_n8.log 2>&1; echo "bench n8 rc=$?"; tail -1 gpurun_out/bench_n8.log | cut -c1-400
timeout 600 $TR scripts/monte_carlo.py --trajectories 1000000 --steps 200 > gpurun_out/mc_n8.log 2>&1; echo "mc rc=$?"; tail -1 gpurun_out/mc_n8.log | cut -c1-600
