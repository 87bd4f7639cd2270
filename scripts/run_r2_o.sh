#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_planner.py tests/test_gpu_knn.py -q -x > gpurun_out/r2o_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2o_pytest.log
timeout 300 python scripts/exp_query_step.py 2>&1 | tail -20
timeout 600 python bench.py --skip-edges --skip-cpu --skip-sharded --skip-brute --steps 10 > gpurun_out/r2o_bench.json 2> gpurun_out/r2o_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r2o_bench.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r2o_bench.json').read().strip().splitlines()[-1])
for k,v in d["planner"].items():
    if isinstance(v, dict): print(k, {kk: v[kk] for kk in ("nodes_per_s","seconds","batches","seconds_all_runs","queries_per_s","success_rate_rank0","one_after_the_other") if kk in v})
P
