"""independent planning queries (configs[4](i) shape): start/goal pairs on a 256^3 map, each run to the goal."""
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
# free start / goal points with clearance: candidates that pass the state check
cand = synth.points(77, 4000, n_map * 0.05)
hit, _ = omap.collide_spheres(cand, 0.1)
free = cand[hit == 0]
print("free candidates", len(free))
EXTRA = int(os.environ.get("EXTRA_FLAGS", "0"))
for batch in (64, 256, 1024):
    t0 = time.perf_counter(); ok = 0; nodes = 0; plen = 0.0
    NQ = 8
    for qi in range(NQ):
        s, g = free[2 * qi], free[2 * qi + 1]
        P = B.PlannerParams(tuple(s), tuple(g), mode=B.PLAN_MODE_MS, k=16, batch=batch,
                            flags=B.PLAN_EDGE_CHECK | B.PLAN_REWIRE | EXTRA, seed=100 + qi)
        pl = B.Planner(P, omap, table, ms)
        reached = pl.run(60000)
        ok += int(reached); nodes += len(pl)
        if reached:
            pts = pl.tree()[0][pl.path()].astype(np.float64)
            plen += float(np.linalg.norm(np.diff(pts, axis=0), axis=1).sum())
    torch.cuda.synchronize()
    el = time.perf_counter() - t0
    print("flags+%d batch %d: %d/%d reached, %.1f nodes avg, %.3f s per query, mean path length %.3f m" % (EXTRA, batch, ok, NQ, nodes / NQ, el / NQ, plen / max(ok, 1)))
