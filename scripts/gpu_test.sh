#!/bin/bash
# usage: gpu_test.sh [pytest args...]
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q "$@" 2>&1 | tail -30
