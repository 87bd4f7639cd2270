"""Sphere state checks on configs[2]'s map (1 M spheres = the edges' first endpoints, 512^3 map at 0.05 m): the
lane-per-sphere kernel (BGR_SPHERES=lane) against the warp forms (warp32 / warp1); CUDA events, L2 flushed between
launches, hit / band bytes compared.  usage: python scripts/exp_spheres.py [radius ...]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bottleneck_guided_rrt_b200 as B  # noqa: E402
from bottleneck_guided_rrt_b200 import synth  # noqa: E402

radii = [float(a) for a in sys.argv[1:]] or [0.1, 0.05, 0.14, 0.03]
dev = torch.device("cuda", 0)
occ = synth.make_map(512, 4, n_walls=8, wall_thick=4, door_cells=5)
boxes = synth.grid_to_boxes(occ, 0.05, (0, 0, 0), 3)
m = B.OccupancyMap(0.05, (0, 0, 0), (512,) * 3)
m.set_boxes(boxes)
p1, _ = synth.edges(4, 1_000_000, occ)
p1 = p1[:int(os.environ.get('NSPH', '1000000'))]
# a tenth of the centres moved outside / onto the border of the map
p1 = p1.copy()
p1[::10] = (p1[::10] - np.float32(12.8)) * np.float32(1.05) + np.float32(12.8)
d1 = torch.from_numpy(p1).to(dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for r in radii:
    res = {}
    for form in ("warp32", "lane", "warp1"):
        os.environ["BGR_SPHERES"] = form
        hit = torch.full((len(p1),), 7, dtype=torch.uint8, device=dev)
        band = torch.full((len(p1),), 7, dtype=torch.uint8, device=dev)
        for _ in range(3):
            m.collide_spheres(d1, r, hit=hit, band=band)
        torch.cuda.synchronize()
        ts = []
        for _ in range(10):
            flush.fill_(1)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            m.collide_spheres(d1, r, hit=hit, band=band)
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        res[form] = (hit.cpu().numpy(), band.cpu().numpy())
        print("r %.3f %-7s mean %.4f ms  min %.4f ms  hit %.3f band %d" % (r, form, np.mean(ts), np.min(ts), res[form][0].mean(), int(res[form][1].sum())))
    same = all(np.array_equal(res["warp32"][0], res[f][0]) and np.array_equal(res["warp32"][1], res[f][1]) for f in ("lane", "warp1"))
    print("r %.3f bytes identical across forms: %s" % (r, same))
