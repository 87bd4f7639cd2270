#!/bin/bash
cat > /tmp/cfg3.py <<'PY'
import os, sys, time, torch, numpy as np
sys.path.insert(0, os.getcwd())
import bottleneck_guided_rrt_b200 as B
from bottleneck_guided_rrt_b200 import synth
import bench
occ = synth.make_map(256, 5, n_walls=3, wall_thick=4, door_cells=5)
boxes = synth.grid_to_boxes(occ, 0.05, (0, 0, 0), 3)
ms = bench.door_milestones(occ, 256)
table = synth.points(9, 200_000, 12.8)
for rep in range(3):
    m = B.OccupancyMap(0.05, (0, 0, 0), (256,) * 3); m.set_boxes(boxes)
    P = B.PlannerParams((0.2, 0.2, 0.2), (12.6, 12.6, 12.6), mode=B.PLAN_MODE_MS, k=16, batch=4096, flags=B.PLAN_EDGE_CHECK | B.PLAN_REWIRE | B.PLAN_NO_GOAL, seed=5489)
    pl = B.Planner(P, m, table, ms)
    pl.run(64 * 4096 + 64, 3); torch.cuda.synchronize()
    t0 = time.perf_counter(); pl.run(200_000); torch.cuda.synchronize(); el = time.perf_counter() - t0
    print("cfg3 growth: %.3f s, %d batches, %.3f ms/step" % (el, pl.stats()["batches"], 1e3 * el / pl.stats()["batches"]))
PY
for r in default 1 2 4; do echo "== BGR_SCAN_R=$r"; if [ $r = default ]; then python /tmp/cfg3.py 2>&1 | tail -3; else BGR_SCAN_R=$r python /tmp/cfg3.py 2>&1 | tail -3; fi; done
