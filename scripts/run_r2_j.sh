#!/bin/bash
mkdir -p gpurun_out
cat > /tmp/seeds.py <<'PY'
import os, sys, torch
sys.path.insert(0, os.getcwd())
import bottleneck_guided_rrt_b200 as B
from bottleneck_guided_rrt_b200 import synth
dev = torch.device("cuda", 0)
pts = torch.from_numpy(synth.points(2, 1_000_000, 25.6)).to(dev)
nns = B.NNSearchF.createKDTreeLinearHeap(pts); nns.set_method(B.METHOD_GRID); nns.build_index()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
idx = torch.empty((65536, 16), dtype=torch.int32, device=dev); d2 = torch.empty((65536, 16), dtype=torch.float32, device=dev)
for seed in (3, 4, 5, 6, 3):
    q = torch.from_numpy(synth.points(seed, 65_536, 25.6)).to(dev)
    for _ in range(10): nns.knn(q, 16, indices=idx, dists2=d2)
    torch.cuda.synchronize(); tot = 0
    for _ in range(20):
        flush.fill_(1); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); nns.knn(q, 16, indices=idx, dists2=d2); b.record(); torch.cuda.synchronize(); tot += a.elapsed_time(b)
    print("query seed", seed, "%.4f ms" % (tot / 20))
PY
timeout 900 python -m pytest tests/test_gpu_knn.py tests/test_gpu_tile_knn.py tests/test_gpu_planner.py -q > gpurun_out/r2j_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2j_pytest.log
echo "== launch bounds (256,4)"; timeout 300 python /tmp/seeds.py 2>&1 | tail -5
sed -i 's/__launch_bounds__(kGridWarps \* 32, 4) grid_knn_kernel/__launch_bounds__(kGridWarps * 32, 5) grid_knn_kernel/' bottleneck_guided_rrt_b200/csrc/grid.cuh
make -C bottleneck_guided_rrt_b200/csrc > /dev/null 2>&1
echo "== launch bounds (256,5)"; timeout 300 python /tmp/seeds.py 2>&1 | tail -5
