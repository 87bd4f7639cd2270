#!/bin/bash
# grid_knn_pair_kernel at __launch_bounds__(256, 4 | 5 | 6) (variants/libbgrrt_c*.so): scripts/exp_pair_knn.py on each
cp bottleneck_guided_rrt_b200/libbgrrt.so /tmp/libbgrrt_keep.so
for v in "$@"; do
  cp variants/libbgrrt_$v.so bottleneck_guided_rrt_b200/libbgrrt.so
  echo "== variant $v"
  timeout 300 python scripts/exp_pair_knn.py 2>&1 | grep -v "pair=0" | tail -3
done
cp /tmp/libbgrrt_keep.so bottleneck_guided_rrt_b200/libbgrrt.so
