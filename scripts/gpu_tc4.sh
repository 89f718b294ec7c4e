#!/bin/bash
set -u
mkdir -p gpurun_out
echo "== probe"; timeout 300 python scripts/tc_probe.py > gpurun_out/tc_probe.log 2>&1; echo "probe rc=$?"; grep -E "dw |BAD|PROBE|EXC" gpurun_out/tc_probe.log | cut -c1-200
echo "== pytest tf32"; timeout 900 python -m pytest tests/test_gpu_tf32.py -q -m gpu > gpurun_out/pytest_tf32.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_tf32.log
for v in 3 2; do
  echo "== bench tf32 DW=$v"; WD_TC_DW=$v timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_dw$v.json 2> gpurun_out/bench_dw$v.err; echo "rc=$?"; tail -2 gpurun_out/bench_dw$v.err
  python - <<PY
import json
d=json.load(open('gpurun_out/bench_dw$v.json'))
print('DW=$v', round(d['value']), 'graphs/s', round(d['ms_per_step'],3), 'ms', ' '.join(f"{k}={x['ms_per_launch']:.3f}x{x['launches_per_step']:.0f}" for k,x in d['kernels'].items()))
PY
done
echo "== e2e probe"; timeout 300 python scripts/e2e_probe.py 4096 2>&1 | tail -7; timeout 300 python scripts/e2e_probe.py 50 2>&1 | tail -3
echo "== ncu launch list"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_tf32.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
