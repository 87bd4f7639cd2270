#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_collide.py tests/test_gpu_planner.py -q > gpurun_out/r2h_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2h_pytest.log
timeout 600 python scripts/tune_gate.py > gpurun_out/r2h_tune.log 2>&1; cat gpurun_out/r2h_tune.log
