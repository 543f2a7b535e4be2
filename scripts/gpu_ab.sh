#!/bin/bash
# A/B timing of library variants: gpu_ab.sh "ENV=.. ENV=.. lib.so" ...   (each argument: optional env
# assignments followed by a file under variants/); prints scripts/variant_time.py's lines
mkdir -p gpurun_out
for spec in "$@"; do
  lib=${spec##* }; envs=${spec% *}; [ "$envs" == "$spec" ] && envs=""
  echo "== $spec" | tee -a gpurun_out/ab.log
  env $envs NMPC_B200_LIB=$PWD/variants/$lib timeout 300 python scripts/variant_time.py 2>&1 | grep -v Warning | tee -a gpurun_out/ab.log
done
