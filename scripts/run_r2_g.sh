#!/bin/bash
mkdir -p gpurun_out
cat > /tmp/one.py <<'PY'
import os, sys, torch
sys.path.insert(0, os.getcwd())
import bottleneck_guided_rrt_b200 as B
from bottleneck_guided_rrt_b200 import synth
dev = torch.device("cuda", 0)
pts = torch.from_numpy(synth.points(2, 1_000_000, 25.6)).to(dev)
q = torch.from_numpy(synth.points(3, 65_536, 25.6)).to(dev)
nns = B.NNSearchF.createKDTreeLinearHeap(pts); nns.set_method(B.METHOD_GRID); nns.build_index()
for _ in range(3):
    nns.knn(q, 16)
torch.cuda.synchronize()
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"tile_knn_kernel|grid_knn_list" --launch-skip 2 -o gpurun_out/r2g_prof_tile -f python /tmp/one.py > gpurun_out/r2g_ncu.log 2>&1; tail -3 gpurun_out/r2g_ncu.log
