#!/bin/bash
for v in timing abl2 abl32; do
  echo "== $v"; WDMPNN_LIB=/root/repo/polymer_chemprop_b200/libwdmpnn_$v.so timeout 300 python scripts/t2_timing.py 2>&1 | grep -A5 "^dW"
done
