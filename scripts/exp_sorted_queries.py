"""Experiment: grid kNN time with the query batch in random order vs sorted by grid cell (z,y,x) vs by coarse tile.
usage: python scripts/exp_sorted_queries.py"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bottleneck_guided_rrt_b200 as B  # noqa: E402
from bottleneck_guided_rrt_b200 import synth  # noqa: E402

dev = torch.device("cuda", 0)
pts = torch.from_numpy(synth.points(2, 1_000_000, 25.6)).to(dev)
nns = B.NNSearchF.createKDTreeLinearHeap(pts)
nns.set_method(B.METHOD_GRID)
nns.build_index()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
idx = torch.empty((65536, 16), dtype=torch.int32, device=dev)
d2 = torch.empty((65536, 16), dtype=torch.float32, device=dev)
qh = synth.points(3, 65_536, 25.6)
cs = 25.6 / 50
c = np.floor(qh / cs).astype(np.int64)
orders = {"random": np.arange(len(qh)),
          "cell zyx": np.lexsort((c[:, 0], c[:, 1], c[:, 2])),
          "tile 4x4x4 then cell": np.lexsort((c[:, 0], c[:, 1], c[:, 2], c[:, 0] // 4, c[:, 1] // 4, c[:, 2] // 4)),
          "row (z,y) only": np.lexsort((c[:, 1], c[:, 2]))}
for name, o in orders.items():
    q = torch.from_numpy(np.ascontiguousarray(qh[o])).to(dev)
    for _ in range(10):
        nns.knn(q, 16, indices=idx, dists2=d2)
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(20):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        nns.knn(q, 16, indices=idx, dists2=d2)
        b.record()
        torch.cuda.synchronize()
        tot += a.elapsed_time(b)
    print("%-24s %.4f ms" % (name, tot / 20))
