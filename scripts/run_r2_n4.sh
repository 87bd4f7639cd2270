#!/bin/bash
# round 2: N=4 bench (both arms) + the multi-GPU tests, as the driver launches them
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2_topo_n4.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_multi.py -q > gpurun_out/r2_n4_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_n4_pytest.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/r2_bench_n4.json 2> gpurun_out/r2_bench_n4.err; echo "bench rc=$?"; tail -5 gpurun_out/r2_bench_n4.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 4 --steps 3 --warmup 3 > gpurun_out/r2_bench_ref_n4.json 2>> gpurun_out/r2_bench_n4.err; echo "ref rc=$?"
python - <<'P'
import json
d=json.loads(open('gpurun_out/r2_bench_n4.json').read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"], d["e2e"]["legs_alone_ms"])
print("edges", d["edge_checks"]["value"], d["edge_checks"]["roofline"]["kernel_ms"])
print("sharded", {k: d["sharded_knn"].get(k) for k in ("nccl","peer","parity","exchanges_identical")})
print("planner q", d["planner"]["config4_independent_queries_256"]["queries_per_s"], d["planner"]["config4_independent_queries_256"]["success_rate_rank0"])
P
