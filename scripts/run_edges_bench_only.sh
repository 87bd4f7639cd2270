python bench.py --skip-cpu --skip-planner --skip-sharded --steps 20 > gpurun_out/bench_edges.json 2> gpurun_out/bench_edges.err; echo "bench rc=$?"
python - <<'P'
import json
d=json.loads(open('gpurun_out/bench_edges.json').read().strip().splitlines()[-1])
e=d['edge_checks']; print(e['ms_per_step'], e['roofline'].get('kernel_ms'), e['config'].get('sweep'))
P
