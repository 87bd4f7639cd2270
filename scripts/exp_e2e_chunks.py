"""e2e kNN (host pinned buffers through bgr_nns_knn) vs the number of chunks the call is cut into (BGR_KNN_CHUNKS).
usage: python scripts/exp_e2e_chunks.py"""
import os
import sys
import time

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
q = torch.from_numpy(synth.points(3, 65_536, 25.6)).pin_memory()
ih = torch.empty((65536, 16), dtype=torch.int32).pin_memory()
dh = torch.empty((65536, 16), dtype=torch.float32).pin_memory()
qn, inn, dn = q.numpy(), ih.numpy(), dh.numpy()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for ch in ("1", "2", "3", "4", "6", "8"):
    os.environ["BGR_KNN_CHUNKS"] = ch
    for want_d2 in (True, False):
        for _ in range(5):
            nns.knn(qn, 16, indices=inn, dists2=dn if want_d2 else None, want_dists2=want_d2)
        tot = 0.0
        for _ in range(20):
            flush.fill_(1)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            nns.knn(qn, 16, indices=inn, dists2=dn if want_d2 else None, want_dists2=want_d2)
            tot += time.perf_counter() - t0
        print("chunks %s  %-12s %.4f ms (host wall clock around the blocking call)" % (ch, "idx + d2" if want_d2 else "idx only", tot / 20 * 1e3))
