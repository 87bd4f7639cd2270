#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_planner.py -q -x > gpurun_out/r2u_pytest.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/r2u_pytest.log
for g in 0 1; do
echo "== BGR_PLAN_NO_GRAPH=$g"
if [ $g = 1 ]; then export BGR_PLAN_NO_GRAPH=1; fi
timeout 600 python bench.py --skip-edges --skip-cpu --skip-sharded --skip-brute --steps 5 2>gpurun_out/r2u_err_$g.log | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
p=d['planner']
print('cfg0 b1 %.4g b256 %.4g nodes/s' % (p['config0_rrtstar_64_batch1']['nodes_per_s'], p['config0_rrtstar_64_batch256']['nodes_per_s']), 'cfg3', p['config3_bottleneck_256_batch4096']['seconds_all_runs'], 'queries %.1f /s' % p['config4_independent_queries_256']['queries_per_s'], 'seq %.1f /s' % p['config4_independent_queries_256']['one_after_the_other']['queries_per_s'], p['config4_independent_queries_256']['success_rate_rank0'])"
tail -2 gpurun_out/r2u_err_$g.log
done
