#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tile_knn.py tests/test_gpu_knn.py -q > gpurun_out/r2f_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r2f_pytest.log
cat > /tmp/one.py <<'PY'
import os, sys, torch
sys.path.insert(0, os.getcwd())
import bottleneck_guided_rrt_b200 as B
from bottleneck_guided_rrt_b200 import synth
dev = torch.device("cuda", 0)
pts = torch.from_numpy(synth.points(2, 1_000_000, 25.6)).to(dev)
q = torch.from_numpy(synth.points(3, 65_536, 25.6)).to(dev)
nns = B.NNSearchF.createKDTreeLinearHeap(pts); nns.set_method(B.METHOD_GRID); nns.build_index()
for sh in ((4,4,3,1),(4,4,4,2),(6,6,6,4)):
    nns.set_tile(*sh)
    for _ in range(3):
        nns.knn(q, 16)
    torch.cuda.synchronize()
PY
BGR_TILE_PROF=1 timeout 300 python /tmp/one.py > gpurun_out/r2f_prof.log 2>&1; cat gpurun_out/r2f_prof.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2f_launches.csv python /tmp/one.py > /dev/null 2>&1
grep -E "tile_|grid_knn_list" gpurun_out/r2f_launches.csv | cut -d, -f5,13- | head -40
