#!/bin/bash
mkdir -p gpurun_out
if [ "$1" == "test" ]; then shift; timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15; fi
for v in "$@"; do
  lib=${v%%:*}; cap=${v##*:}
  NMPC_B200_MAX_CTAS_PER_SM=$cap NMPC_B200_LIB=$PWD/variants/$lib timeout 300 python scripts/variant_time.py 2>&1 | grep -v Warning | sed "s/^/cap=$cap /" | tee -a gpurun_out/variants.log
done
