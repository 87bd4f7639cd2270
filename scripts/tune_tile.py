"""Tile kNN tuning on the configs[1] workload: kernel time (CUDA events, L2 flushed between runs) per tile shape.
usage: python scripts/tune_tile.py [nodes] [queries]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bottleneck_guided_rrt_b200 as B  # noqa: E402
from bottleneck_guided_rrt_b200 import synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 65_536
ext = 25.6 * (n / 1e6) ** (1 / 3)
dev = torch.device("cuda", 0)
pts = torch.from_numpy(synth.points(2, n, ext)).to(dev)
q = torch.from_numpy(synth.points(3, nq, ext)).to(dev)
nns = B.NNSearchF.createKDTreeLinearHeap(pts)
nns.set_method(B.METHOD_GRID)
nns.build_index()
idx = torch.empty((nq, 16), dtype=torch.int32, device=dev)
d2 = torch.empty((nq, 16), dtype=torch.float32, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(reps=20):
    for _ in range(5):
        nns.knn(q, 16, indices=idx, dists2=d2)
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(reps):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        nns.knn(q, 16, indices=idx, dists2=d2)
        b.record()
        torch.cuda.synchronize()
        tot += a.elapsed_time(b)
    return tot / reps


t0 = time_ref = None
nns.set_tile(-1)
time_ref = timed()
ref_i, ref_d = idx.clone(), d2.clone()
print("warp-per-query grid search: %.4f ms" % time_ref)
shapes = [(0, 0, 0, 0), (4, 4, 3, 1), (4, 4, 3, 2), (4, 3, 4, 1), (3, 4, 4, 1), (4, 4, 4, 1), (4, 4, 4, 2), (6, 4, 2, 1), (8, 3, 2, 1),
          (5, 5, 4, 2), (6, 6, 6, 4), (6, 6, 3, 2), (8, 4, 3, 2), (8, 6, 6, 4), (3, 3, 3, 1), (4, 3, 3, 1), (2, 2, 2, 1)]
for sh in shapes:
    try:
        nns.set_tile(*sh)
        ms = timed()
        st = nns.tile_stats()
        ok = bool(torch.equal(idx, ref_i) and torch.equal(d2.view(torch.int32), ref_d.view(torch.int32)))
        print("tile %s warps %d tiles %6d: %.4f ms  fallback %5d  identical %s" % (st["tile"], st["warps"], st["tiles"], ms, st["fallback_queries_last"], ok))
    except Exception as ex:
        print("tile", sh, "failed:", repr(ex)[:200])
