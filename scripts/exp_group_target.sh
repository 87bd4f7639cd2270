#!/bin/bash
# grid_knn_group_kernel: survivors the a-priori threshold aims at = k + f sqrt(k) + 1 (BGR_FAST_TARGET=f; default 2.5)
for f in "$@"; do
  echo "== f $f"
  BGR_FAST_TARGET=$f timeout 300 python scripts/exp_group_knn.py 2>&1 | grep -v "query=0" | tail -3
done
