#!/bin/bash
# Timing ablations of the tcgen05 GEMM kernel (WD_ABL build variants; results are wrong by design).
set -u
mkdir -p gpurun_out
for v in "" _abl1 _abl2 _abl3 _abl4 _abl8 _abl16 _abl32 _abl63; do
  lib=/root/repo/polymer_chemprop_b200/libwdmpnn$v.so
  [ -f "$lib" ] || continue
  WDMPNN_LIB=$lib timeout 300 python bench.py --steps 10 --warmup 3 --math tf32 --no-cpu-baseline --no-e2e > gpurun_out/bench$v.json 2> gpurun_out/bench$v.err; rc=$?
  python - <<PY
import json
try:
    d=json.load(open('gpurun_out/bench$v.json'))
    print('variant "$v"', round(d['ms_per_step'],3), 'ms', ' '.join(f"{k}={v['ms_per_launch']:.3f}" for k,v in d['kernels'].items() if k in ('input_proj','depth_step','atom_update','dx')))
except Exception as e:
    print('variant "$v" failed rc=$rc', e)
PY
done
