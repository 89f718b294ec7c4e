#!/bin/bash
set -u
mkdir -p gpurun_out
echo "== probe"; timeout 300 python scripts/tc_probe.py > gpurun_out/tc_probe.log 2>&1; echo "probe rc=$?"; tail -30 gpurun_out/tc_probe.log
echo "== pytest tf32"; timeout 900 python -m pytest tests/test_gpu_tf32.py -q -m gpu > gpurun_out/pytest_tf32.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_tf32.log
for v1 in 0; do
  echo "== bench tf32 V1=$v1"; WD_TC_V1=$v1 timeout 600 python bench.py --steps 10 --warmup 3 --math tf32 --no-cpu-baseline --no-e2e > gpurun_out/bench_tf32_v$v1.json 2> gpurun_out/bench_tf32_v$v1.err; echo "rc=$?"; tail -2 gpurun_out/bench_tf32_v$v1.err
  python - <<PY
import json
d=json.load(open('gpurun_out/bench_tf32_v$v1.json'))
print('V1=$v1', round(d['value']), 'graphs/s', round(d['ms_per_step'],3), 'ms')
for k,v in d['kernels'].items(): print('  ',k, round(v['ms_per_launch'],3), v['launches_per_step'])
PY
done
echo "== ncu launch list"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_tf32.csv python bench.py --steps 2 --warmup 3 --math tf32 --no-cpu-baseline --no-e2e > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
echo "== ncu full (tc kernels of one step)"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"k_tc_(gemm|dw)" -c 11 -f -o gpurun_out/tc_full python bench.py --steps 1 --warmup 3 --math tf32 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"; tail -3 gpurun_out/ncu_full.log
