#!/bin/bash
mkdir -p gpurun_out
for cfg in "8 0.047" "8 0.065" "12 0.065" "12 0.10"; do set -- $cfg; DOOR=$1 OCC=$2 timeout 300 python scripts/try_queries2.py 32 > gpurun_out/r2c_queries_$1_$2.log 2>&1; tail -5 gpurun_out/r2c_queries_$1_$2.log; done
