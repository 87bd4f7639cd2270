"""Planner step under ncu: grow the configs[3] tree to ~150k nodes unprofiled, then profile a
few batches (cudaProfilerStart/Stop; run ncu with --profile-from-start off)."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bottleneck_guided_rrt_b200 as B
from bottleneck_guided_rrt_b200 import synth
n_map = 256
occ = synth.make_map(n_map, 5, n_walls=3, wall_thick=4, door_cells=5)
boxes = synth.grid_to_boxes(occ, 0.05, (0, 0, 0), 3)
omap = B.OccupancyMap(0.05, (0, 0, 0), (n_map,) * 3)
omap.set_boxes(boxes)
ms = []
for w in range(3):
    x0 = (w + 1) * 256 // 4
    zz, yy = np.nonzero(occ[:, :, x0] == 0)
    ms.append([(x0 + 0.5) * 0.05, (yy.mean() + 0.5) * 0.05, (zz.mean() + 0.5) * 0.05])
ms = np.asarray(ms, dtype=np.float32)
table = synth.points(9, 200_000, n_map * 0.05)
P = B.PlannerParams((0.2, 0.2, 0.2), (12.6, 12.6, 12.6), mode=B.PLAN_MODE_MS, k=16, batch=int(os.environ.get("BATCH", "4096")),
                    flags=B.PLAN_EDGE_CHECK | B.PLAN_REWIRE | B.PLAN_NO_GOAL, seed=5489)
pl = B.Planner(P, omap, table, ms)
target = int(os.environ.get("GROW", "150000"))
t = time.perf_counter()
pl.run(target)
torch.cuda.synchronize()
print("grown to", len(pl), "in %.3f s" % (time.perf_counter() - t), pl.stats())
torch.cuda.profiler.start()
t = time.perf_counter()
pl.run(1 << 30, 6)
torch.cuda.synchronize()
print("6 batches: %.3f ms each" % ((time.perf_counter() - t) / 6 * 1e3))
torch.cuda.profiler.stop()
