"""Short driver for ncu: each hot kernel twice (one warm launch, one to capture).
usage: python scripts/prof_run.py [knn_grid] [knn_brute] [edges] [nearest] [spheres]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bottleneck_guided_rrt_b200 as B  # noqa: E402
from bottleneck_guided_rrt_b200 import synth  # noqa: E402

what = sys.argv[1:] or ["knn_grid", "knn_brute", "edges"]
dev = torch.device("cuda", 0)
if any(w.startswith("knn") or w == "nearest" for w in what):
    pts = torch.from_numpy(synth.points(2, 1_000_000, 25.6)).to(dev)
    q = torch.from_numpy(synth.points(3, 65_536, 25.6)).to(dev)
    nns = B.NNSearchF.createKDTreeLinearHeap(pts)
    nns.build_index()
    for w in what:
        if w == "knn_grid":
            nns.set_method(B.METHOD_GRID)
        elif w == "knn_brute":
            nns.set_method(B.METHOD_BRUTE)
        elif w == "nearest":
            for _ in range(2):
                nns.nearest(q[:4096])
            continue
        else:
            continue
        for _ in range(2):
            nns.knn(q, 16)
        torch.cuda.synchronize()
if "edges" in what or "spheres" in what:
    n = int(os.environ.get("MAP_N", "512"))
    occ = synth.make_map(n, 4, n_walls=8, wall_thick=4, door_cells=5)
    boxes = synth.grid_to_boxes(occ, 0.05, (0, 0, 0), 3)
    m = B.OccupancyMap(0.05, (0, 0, 0), (n, n, n))
    m.set_boxes(boxes)
    p1, p2 = synth.edges(4, 1_000_000, occ)
    d1, d2 = torch.from_numpy(p1).to(dev), torch.from_numpy(p2).to(dev)
    for _ in range(2):
        if "edges" in what:
            m.collide_edges(d1, d2, 0.03, 0)
        if "spheres" in what:
            m.collide_spheres(d1, 0.1)
    torch.cuda.synchronize()
    print(m.last_stats())
print("done", B.lib().bgr_launch_count())
