# all GPU tests + default bench (no CPU baselines, no planner)
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python bench.py --skip-cpu --skip-planner --steps 20 > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?" >> gpurun_out/bench_quick.err
tail -4 gpurun_out/pytest_gpu.log; tail -3 gpurun_out/bench_quick.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/bench_quick.json').read().strip().splitlines()[-1])
print('knn', d['ms_per_step'], d['e2e'])
e=d['edge_checks']; print('edges', e['ms_per_step'], e['roofline'].get('kernel_ms'), e['e2e'])
P
