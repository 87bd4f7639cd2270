#!/bin/bash
# round 2, after the group kNN / lane sphere kernels: N=2 bench + the multi-GPU tests, as the driver launches them
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -q > gpurun_out/r2t_n2_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2t_n2_pytest.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2t_bench_n2.json 2> gpurun_out/r2t_bench_n2.err; echo "bench rc=$?"; tail -5 gpurun_out/r2t_bench_n2.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r2t_bench_n2.json').read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"], d["e2e"]["legs_alone_ms"])
print("edges", d["edge_checks"]["value"], "spheres", d["state_checks"]["value"], d["state_checks"]["e2e"]["value"])
print("sharded", {k: d["sharded_knn"].get(k) for k in ("nccl","peer","parity","exchanges_identical")})
print("planner q", d["planner"]["config4_independent_queries_256"]["queries_per_s"], d["planner"]["config4_independent_queries_256"]["success_rate_rank0"])
P
