# edge-path check: collide tests + planner/path tests, bench edges only, per-kernel times
python -m pytest tests/test_gpu_collide.py tests/test_gpu_planner.py tests/test_gpu_path.py -x -q > gpurun_out/pytest_edges.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_edges.log
python bench.py --skip-cpu --skip-planner --steps 20 > gpurun_out/bench_edges.json 2> gpurun_out/bench_edges.err; echo "bench rc=$?" >> gpurun_out/bench_edges.err
tail -5 gpurun_out/pytest_edges.log; tail -3 gpurun_out/bench_edges.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/bench_edges.json').read().strip().splitlines()[-1])
e=d['edge_checks']; print(e['ms_per_step'], e['roofline'].get('kernel_ms'), e['config'].get('sweep'), e['e2e'])
P
