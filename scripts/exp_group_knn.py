"""grid_knn_group_kernel<16 | 8> (a query per 16 / 8 lanes) against grid_knn_kernel (BGR_KNN_GROUP=0) on configs[1]
(1 M uniform nodes, 65,536 queries, k = 16): kernel ms by CUDA events with an L2 flush between iterations, rows
compared bit for bit.  usage: python scripts/exp_pair_knn.py [n] [nq] [k]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bottleneck_guided_rrt_b200 as B  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
k = int(sys.argv[3]) if len(sys.argv) > 3 else 16
rng = np.random.default_rng(1)
pts = (rng.random((n, 3), dtype=np.float32) * np.float32(25.6)).astype(np.float32)
q = (rng.random((nq, 3), dtype=np.float32) * np.float32(25.6)).astype(np.float32)
dev = torch.device("cuda", 0)
nns = B.NNSearchF.createKDTreeLinearHeap(pts)
nns.set_method(B.METHOD_GRID)
dq = torch.from_numpy(q).to(dev)
idx = torch.empty((nq, k), dtype=torch.int32, device=dev)
d2 = torch.empty((nq, k), dtype=torch.float32, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
res = {}
for rnd in range(2):
    for pair in ("0", "16", "8"):
        os.environ["BGR_KNN_GROUP"] = pair
        for _ in range(3):
            nns.knn(dq, k, 0.0, 0, indices=idx, dists2=d2)
        torch.cuda.synchronize()
        ts = []
        for _ in range(20):
            flush.fill_(1)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            nns.knn(dq, k, 0.0, 0, indices=idx, dists2=d2)
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        res[pair] = (idx.cpu().numpy().copy(), d2.cpu().numpy().copy())
        print("lanes per query=%s  n %d nq %d k %d: mean %.4f ms  min %.4f ms  (%.3g queries/s at the mean)" % (pair, n, nq, k, np.mean(ts), np.min(ts), nq / np.mean(ts) * 1e3))
same = all(np.array_equal(res["0"][0], res[v][0]) and np.array_equal(res["0"][1].view(np.uint32), res[v][1].view(np.uint32)) for v in ("16", "8"))
print("rows identical:", same)
