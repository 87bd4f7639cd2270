import os, sys, numpy as np
sys.path.insert(0, os.getcwd())
exec(open('scripts/prof_planner.py').read().split("torch.cuda.profiler.start()")[0])
xyz, par = pl.tree()
n = len(xyz)
mn, mx = xyz.min(0), xyz.max(0)
print("bbox", mn, mx)
for ppc in (8, 2):
    cs = ((mx - mn).prod() / (n / ppc)) ** (1 / 3)
    g = np.maximum(1, np.floor((mx - mn) / cs).astype(int) + 1)
    c = np.minimum(((xyz - mn) / cs).astype(int), g - 1)
    cid = (c[:, 2] * g[1] + c[:, 1]) * g[0] + c[:, 0]
    cnt = np.bincount(cid, minlength=g.prod())
    print("ppc", ppc, "grid", g, "cells", g.prod(), "nonempty", (cnt > 0).sum(), "max", cnt.max(), "p99 of nonempty", np.percentile(cnt[cnt > 0], [50, 90, 99, 99.9]))
    srt = np.sort(cnt)[::-1]
    print(" top10 cells", srt[:10], " nodes in top 100 cells", srt[:100].sum())
