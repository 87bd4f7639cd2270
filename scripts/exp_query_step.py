"""Per-step wall time of ONE host-driven planning query of the configs[4](i) shape (256^3 query map, MS sampler, batch 1024,
k = 16, edge checks + rewiring + dedup) as the tree grows; and the same with the library's index statistics.
usage: python scripts/exp_query_step.py"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bottleneck_guided_rrt_b200 as B  # noqa: E402
from bottleneck_guided_rrt_b200 import synth  # noqa: E402
import bench  # noqa: E402

occ = synth.make_map(256, 5, n_walls=3, wall_thick=4, door_cells=bench.QUERY_DOOR_CELLS, occupancy=bench.QUERY_OCCUPANCY)
ms = bench.door_milestones(occ, 256)
boxes = synth.grid_to_boxes(occ, 0.05, (0, 0, 0), 3)
m = B.OccupancyMap(0.05, (0, 0, 0), (256,) * 3)
m.set_boxes(boxes)
table = synth.points(9, 200_000, 12.8)
for batch in (1024, 4096):
    P = B.PlannerParams((0.3, 0.3, 0.3), (12.5, 12.5, 12.5), mode=B.PLAN_MODE_MS, k=16, batch=batch,
                        flags=B.PLAN_EDGE_CHECK | B.PLAN_REWIRE | B.PLAN_DEDUP | B.PLAN_NO_GOAL, seed=101)
    pl = B.Planner(P, m, table, ms)
    pl.run(2000, 5)
    torch.cuda.synchronize()
    last_n, last_b, t0 = len(pl), pl.stats()["batches"], time.perf_counter()
    while len(pl) < 60000:
        pl.run(len(pl) + 6000)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        nb = pl.stats()["batches"]
        dbg = pl.nns_debug_stats()
        print("batch %d: tree %6d -> %6d nodes, %3d steps, %.3f ms/step, index builds so far %d, brute kNN launches %d, grid kNN piled %d / lean %d, scan nearest %d, grid nearest %d"
              % (batch, last_n, len(pl), nb - last_b, 1e3 * (t1 - t0) / max(nb - last_b, 1), dbg["index_builds"], dbg["brute_knn_launches"],
                 dbg["grid_knn_piled_launches"], dbg["grid_knn_launches"], dbg["scan_nearest_launches"], dbg["grid_nearest_launches"]))
        last_n, last_b, t0 = len(pl), nb, time.perf_counter()
