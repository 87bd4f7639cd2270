#!/bin/bash
# round 2, call A: full GPU test suite, L2 microbenchmark, configs[4](i) experiment
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r2a_smi.txt 2>&1
nvidia-smi topo -m > gpurun_out/r2a_topo.txt 2>&1
lscpu | head -30 > gpurun_out/r2a_lscpu.txt 2>&1
numactl -H >> gpurun_out/r2a_lscpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2a_pytest.log
timeout 120 ./scripts/micro/l2_read_bench > gpurun_out/r2a_l2_bench.txt 2>&1; cat gpurun_out/r2a_l2_bench.txt
timeout 600 python scripts/try_queries2.py 32 > gpurun_out/r2a_queries.log 2>&1; cat gpurun_out/r2a_queries.log
