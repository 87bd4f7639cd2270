#!/bin/bash
for mt in 4 2 1; do
echo "== BGR_SCAN_MIN_TILES=$mt"
BGR_SCAN_MIN_TILES=$mt python bench.py --skip-edges --skip-cpu --skip-sharded --skip-brute --steps 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
p=d['planner']
print('cfg0 b256 %.4g nodes/s' % p['config0_rrtstar_64_batch256']['nodes_per_s'], 'cfg3', [round(v,3) for v in p['config3_bottleneck_256_batch4096']['seconds_all_runs']], 'queries %.1f /s' % p['config4_independent_queries_256']['queries_per_s'], 'seq %.1f /s' % p['config4_independent_queries_256']['one_after_the_other']['queries_per_s'])"
done
