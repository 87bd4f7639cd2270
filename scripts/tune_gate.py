"""edges_sweep_kernel gate tuning on configs[2]: per-kernel ms (library events) for BGR_GATE=seg,rounds settings.
usage: python scripts/tune_gate.py"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bottleneck_guided_rrt_b200 as B  # noqa: E402
from bottleneck_guided_rrt_b200 import synth  # noqa: E402

dev = torch.device("cuda", 0)
occ = synth.make_map(512, 4, n_walls=8, wall_thick=4, door_cells=5)
boxes = synth.grid_to_boxes(occ, 0.05, (0, 0, 0), 3)
m = B.OccupancyMap(0.05, (0, 0, 0), (512,) * 3)
m.set_boxes(boxes)
p1, p2 = synth.edges(4, 1_000_000, occ)
d1, d2 = torch.from_numpy(p1).to(dev), torch.from_numpy(p2).to(dev)
hit = torch.empty(len(p1), dtype=torch.uint8, device=dev)
band = torch.empty(len(p1), dtype=torch.uint8, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
ref = None
CFGS = os.environ.get("BGR_TUNE_CFGS", "8,0 4,4 4,2 4,3 4,6 4,8 8,1 8,2 8,3 2,4 2,8 16,1 4,32 8,16").split()
for cfg in CFGS:
    os.environ["BGR_GATE"] = cfg
    for _ in range(3):
        m.collide_edges(d1, d2, 0.03, 0, hit=hit, band=band)
    torch.cuda.synchronize()
    m.set_timing(True)
    tot, kms = 0.0, {}
    for _ in range(10):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        m.collide_edges(d1, d2, 0.03, 0, hit=hit, band=band)
        b.record()
        torch.cuda.synchronize()
        tot += a.elapsed_time(b) / 10
        for k, v in m.last_kernel_ms().items():
            kms[k] = kms.get(k, 0.0) + v / 10
    m.set_timing(False)
    hb = (hit.cpu().numpy(), band.cpu().numpy())
    if ref is None:
        ref = hb
    same = np.array_equal(ref[0], hb[0]) and np.array_equal(ref[1], hb[1])
    print("gate seg,rounds %-5s step %.4f ms sweep %.4f narrow %.4f  identical %s  %s" % (cfg, tot, kms["edges_sweep_kernel"], kms["edges_narrow_kernel"], same, m.last_stats()))
