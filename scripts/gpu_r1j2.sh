#!/bin/bash
# Session J run 2: GPU tests (dropout, inference), PDL variants A/B: early trigger / wait only / off.
set -u
mkdir -p gpurun_out
echo "== pytest -m gpu"; timeout 1200 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
run() {  # name lib pdl batch steps
  WDMPNN_LIB=$2 WD_PDL=$3 timeout 300 python bench.py --steps $5 --warmup 5 --batch $4 --no-cpu-baseline --no-e2e > gpurun_out/v_$1.json 2> gpurun_out/v_$1.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/v_$1.json")); print("$1", round(d["value"]), "graphs/s", round(d["ms_per_step"], 4), "ms; with events", round(d["ms_per_step_with_events"], 4))
except Exception as e:
    print("$1 failed", e)
PY
}
L=/root/repo/polymer_chemprop_b200
for b in 4096 512 50; do
  s=20; [ $b = 50 ] && s=100; [ $b = 512 ] && s=50
  run trig_b$b $L/libwdmpnn.so 1 $b $s
  run wait_b$b $L/libwdmpnn_nt.so 1 $b $s
  run off_b$b $L/libwdmpnn.so 0 $b $s
done
run off2_b4096 $L/libwdmpnn.so 0 4096 20
run trig2_b4096 $L/libwdmpnn.so 1 4096 20
