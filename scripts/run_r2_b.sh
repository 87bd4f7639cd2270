#!/bin/bash
# round 2, call B: full GPU test suite (no -x), configs[4](i) experiment at lower clutter
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2b_pytest.log
for o in 0.02 0.04; do OCC=$o timeout 300 python scripts/try_queries2.py 32 > gpurun_out/r2b_queries_$o.log 2>&1; cat gpurun_out/r2b_queries_$o.log; done
