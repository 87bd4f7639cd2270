python bench.py --skip-cpu --skip-planner --skip-sharded --skip-edges --steps 30 > gpurun_out/bench_knn.json 2> gpurun_out/bench_knn.err; echo "bench rc=$?"
python - <<'P'
import json
d=json.loads(open('gpurun_out/bench_knn.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['roofline']['fp32']['pairs_evaluated_per_step'], d['config']['index_build_ms'])
P
