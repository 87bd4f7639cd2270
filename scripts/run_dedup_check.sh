#!/bin/bash
# planner parity incl. the dedup flag, then queries-to-goal with and without it
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_planner.py -q -m gpu -x > gpurun_out/dedup_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/dedup_pytest.log
EXTRA_FLAGS=0 timeout 300 python scripts/try_queries.py > gpurun_out/dedup_q0.log 2>&1; cat gpurun_out/dedup_q0.log
EXTRA_FLAGS=16 timeout 300 python scripts/try_queries.py > gpurun_out/dedup_q16.log 2>&1; cat gpurun_out/dedup_q16.log
