#!/bin/bash
# grid_knn_group_kernel build variants (variants/libbgrrt_r<rows in flight>w<warps per SM>.so): scripts/exp_group_knn.py on each
cp bottleneck_guided_rrt_b200/libbgrrt.so /tmp/libbgrrt_keep.so
for v in "$@"; do
  cp variants/libbgrrt_$v.so bottleneck_guided_rrt_b200/libbgrrt.so
  echo "== variant $v"
  timeout 300 python scripts/exp_group_knn.py 2>&1 | grep "query=16\|identical" | tail -2
done
cp /tmp/libbgrrt_keep.so bottleneck_guided_rrt_b200/libbgrrt.so
