#!/bin/bash
# sphere-kernel change: collision + planner parity tests, then the bench without CPU arms / planner
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_collide.py tests/test_gpu_planner.py -m gpu -x -q > gpurun_out/pytest_sph.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_sph.log
timeout 600 python bench.py --skip-cpu --skip-planner --steps 20 > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
python - <<'P'
import json
d=json.loads(open('gpurun_out/bench_quick.json').read().strip().splitlines()[-1])
s=d['state_checks']; print('spheres', s['value'], s['ms_per_step'], s['config'])
e=d['edge_checks']; print('edges', e['ms_per_step'])
P
