#!/bin/bash
# edges_sweep_kernel: gate probe loads batched kGateBatch slabs at a time (variants/libbgrrt_b*.so built with
# -DBGR_GATE_BATCH=n) against the unbatched loop (libbgrrt_base.so); per-kernel ms from scripts/tune_gate.py
cp bottleneck_guided_rrt_b200/libbgrrt.so /tmp/libbgrrt_keep.so
for v in base b1 b2 b4 b8 base b4; do
  cp variants/libbgrrt_$v.so bottleneck_guided_rrt_b200/libbgrrt.so
  echo "== variant $v"
  BGR_TUNE_CFGS="8,2 16,1 16,2 8,0" timeout 300 python scripts/tune_gate.py 2>&1 | tail -6
done
cp /tmp/libbgrrt_keep.so bottleneck_guided_rrt_b200/libbgrrt.so
