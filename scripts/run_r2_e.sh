#!/bin/bash
# round 2, call E: tile kNN — parity tests, tuning sweep
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tile_knn.py tests/test_gpu_knn.py -q -x > gpurun_out/r2e_pytest.log 2>&1; echo "pytest rc=$?"; tail -30 gpurun_out/r2e_pytest.log
timeout 600 python scripts/tune_tile.py > gpurun_out/r2e_tune.log 2>&1; cat gpurun_out/r2e_tune.log
