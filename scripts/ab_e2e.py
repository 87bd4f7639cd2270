"""A/B of the chunked copy/compute overlap of the host-pointer entry points, same process, same box.
BGR_NO_PIPE=1 disables the overlap (read per call)."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bottleneck_guided_rrt_b200 as B
from bottleneck_guided_rrt_b200 import synth
dev = torch.device("cuda", 0)
n = 512
occ = synth.make_map(n, 4, n_walls=8, wall_thick=4, door_cells=5)
boxes = synth.grid_to_boxes(occ, 0.05, (0, 0, 0), 3)
m = B.OccupancyMap(0.05, (0, 0, 0), (n, n, n))
m.set_boxes(boxes)
p1, p2 = synth.edges(4, 1_000_000, occ)
p1h, p2h = torch.from_numpy(p1).pin_memory(), torch.from_numpy(p2).pin_memory()
hh, bh = torch.empty(len(p1), dtype=torch.uint8).pin_memory(), torch.empty(len(p1), dtype=torch.uint8).pin_memory()
pts = synth.points(2, 1_000_000, 25.6); q = synth.points(3, 65_536, 25.6)
nns = B.NNSearchF.createKDTreeLinearHeap(pts); nns.build_index()
qh = torch.from_numpy(q).pin_memory()
ih, dh = torch.empty((65536, 16), dtype=torch.int32).pin_memory(), torch.empty((65536, 16), dtype=torch.float32).pin_memory()
ref = {}
for rnd in range(3):
    for mode in ("0", "1"):
        os.environ["BGR_NO_PIPE"] = mode
        for name, fn in (("edges", lambda: m.collide_edges(p1h.numpy(), p2h.numpy(), 0.03, 0, hit=hh.numpy(), band=bh.numpy())),
                         ("knn", lambda: nns.knn(qh.numpy(), 16, indices=ih.numpy(), dists2=dh.numpy()))):
            for _ in range(3): fn()
            torch.cuda.synchronize(); t = time.perf_counter()
            for _ in range(20): r = fn()
            torch.cuda.synchronize(); el = (time.perf_counter() - t) / 20 * 1e3
            print("round %d no_pipe=%s %-6s %.3f ms" % (rnd, mode, name, el))
            key = name
            out = (hh.numpy().copy(), bh.numpy().copy()) if name == "edges" else (r[0].copy(), r[1].copy())
            if key in ref: assert all(np.array_equal(a, b) for a, b in zip(ref[key], out)), "pipe changed results"
            else: ref[key] = out
print("identical results with and without the overlap")
