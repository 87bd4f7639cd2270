"""configs[4](i) experiment: independent start/goal queries on the 256^3 narrow-passage map through
bgr_planner_run_many; success rate / time / nodes for a few planner settings.
usage: python scripts/try_queries2.py NQ"""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bottleneck_guided_rrt_b200 as B
from bottleneck_guided_rrt_b200 import synth
NQ = int(sys.argv[1]) if len(sys.argv) > 1 else 32
OCC = float(os.environ.get("OCC", "0.10"))
DOOR = int(os.environ.get("DOOR", "5"))
occ = synth.make_map(256, 5, n_walls=3, wall_thick=4, door_cells=DOOR, occupancy=OCC)
print("door cells", DOOR)
print("clutter occupancy target", OCC, "actual", float(occ.mean()))
boxes = synth.grid_to_boxes(occ, 0.05, (0, 0, 0), 3)
ms = []
for w in range(3):
    x0 = (w + 1) * 256 // 4
    zz, yy = np.nonzero(occ[:, :, x0] == 0)
    ms.append([(x0 + 0.5) * 0.05, (yy.mean() + 0.5) * 0.05, (zz.mean() + 0.5) * 0.05])
ms = np.asarray(ms, dtype=np.float32)
print("milestones", ms.tolist())
table = synth.points(9, 200_000, 12.8)
nconc = 16
streams = [torch.cuda.Stream() for _ in range(nconc)]
maps = []
for s in streams:
    m = B.OccupancyMap(0.05, (0, 0, 0), (256,) * 3, stream=s.cuda_stream)
    m.set_boxes(boxes)
    maps.append(m)
cand = synth.points(77, 8 * NQ + 4000, 12.8)
hit, _ = maps[0].collide_spheres(cand, 0.1)
free = cand[np.asarray(hit) == 0]
print("free", len(free))
for flags, batch, cap in ((3, 1024, 60000), (3 | 16, 1024, 60000), (3, 256, 60000)):
    ok = 0; nodes = []; t0 = time.perf_counter(); okn = []
    for g0 in range(0, NQ, nconc):
        pls = []
        for j, qi in enumerate(range(g0, min(NQ, g0 + nconc))):
            P = B.PlannerParams(tuple(free[2 * qi]), tuple(free[2 * qi + 1]), mode=B.PLAN_MODE_MS, k=16, batch=batch, flags=flags, seed=100 + qi)
            pls.append(B.Planner(P, maps[j], table, ms, stream=streams[j].cuda_stream))
        fl = B.Planner.run_many(pls, cap)
        torch.cuda.synchronize()
        ok += sum(fl); nodes += [len(p) for p in pls]; okn += [len(p) for p, f in zip(pls, fl) if f]
        del pls
    el = time.perf_counter() - t0
    print("flags %2d batch %4d cap %6d: %d/%d reached, %.3f s total (%.1f q/s), median nodes %d, median nodes of successes %d"
          % (flags, batch, cap, ok, NQ, el, NQ / el, int(np.median(nodes)), int(np.median(okn)) if okn else -1), flush=True)
