#!/bin/bash
# Benchmarks several build variants of the library (WDMPNN_LIB) back to back.
set -u
mkdir -p gpurun_out
for n in "" _v1 _v2 _v3 _v4; do
  lib=/root/repo/polymer_chemprop_b200/libwdmpnn$n.so
  [ -f "$lib" ] || continue
  echo "== variant '$n'"
  WDMPNN_LIB=$lib timeout 300 python -m pytest tests/test_gpu_tf32.py -q -m gpu -k "golden or dw_matches or gemm_matches" 2>&1 | tail -1
  WDMPNN_LIB=$lib timeout 600 python bench.py --steps 10 --warmup 3 --math tf32 --no-cpu-baseline --no-e2e > gpurun_out/bench_var$n.json 2> gpurun_out/bench_var$n.err; echo "rc=$?"; tail -2 gpurun_out/bench_var$n.err
  python - <<PY
import json
d=json.load(open('gpurun_out/bench_var$n.json'))
print('variant $n', round(d['value']), 'graphs/s', round(d['ms_per_step'],3), 'ms', ' '.join(f"{k}={v['ms_per_launch']:.3f}" for k,v in d['kernels'].items()))
PY
done
