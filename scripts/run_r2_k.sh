#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_planner.py tests/test_gpu_path.py -q -x > gpurun_out/r2k_pytest.log 2>&1; echo "pytest rc=$?"; tail -30 gpurun_out/r2k_pytest.log
cat > /tmp/pl.py <<'PY'
import os, sys, time, torch
sys.path.insert(0, os.getcwd())
import bottleneck_guided_rrt_b200 as B
from bottleneck_guided_rrt_b200 import synth
occ = synth.make_map(64, 1, n_walls=3, wall_thick=2, door_cells=6)
boxes = synth.grid_to_boxes(occ, 0.05, (0, 0, 0), 3)
for extra, name in ((0, "persistent"), (B.PLAN_NO_PERSIST, "host-driven")):
    for batch in (1, 4):
        m = B.OccupancyMap(0.05, (0, 0, 0), (64,) * 3); m.set_boxes(boxes)
        table = synth.points(9, 200_000, 3.2)
        P = B.PlannerParams((0.2, 0.2, 0.2), (3.0, 3.0, 3.0), mode=B.PLAN_MODE_RS, k=10, batch=batch, flags=B.PLAN_QUIRKS | B.PLAN_NO_GOAL | extra, seed=5489)
        pl = B.Planner(P, m, table, None)
        pl.run(200, 3); torch.cuda.synchronize()
        n0, t0 = len(pl), time.perf_counter()
        pl.run(5000); torch.cuda.synchronize()
        el = time.perf_counter() - t0
        print("%-12s batch %d: %d nodes in %.3f s = %.0f nodes/s, %d steps, %.1f us/step %s" % (name, batch, len(pl), el, (len(pl) - n0) / el, pl.stats()["batches"], 1e6 * el / pl.stats()["batches"], pl.persist_stats()))
PY
timeout 300 python /tmp/pl.py 2>&1 | tail -6
