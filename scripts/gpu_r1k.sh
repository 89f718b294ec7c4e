#!/bin/bash
# Session K: TMA box for message[b2revb] in the gathered depth step -- tests, A/B against the gathered path, benches.
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_gpu.log | cut -c1-300
run() {  # name env batch steps extra
  env $2 timeout 400 python bench.py --steps $4 --warmup 5 --batch $3 --no-cpu-baseline $5 > gpurun_out/k_$1.json 2> gpurun_out/k_$1.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/k_$1.json")); e = d.get("e2e"); ef = d.get("e2e_full_tensors")
    print("$1", round(d["value"]), "graphs/s", round(d["ms_per_step"], 4), "ms; with events", round(d["ms_per_step_with_events"], 4), "; depth_step", round(d["kernels"]["depth_step"]["ms_per_launch"], 4),
          "; e2e", e and (round(e["value"]), round(e["ms_per_step"], 3), round(e["ms_per_step_median"], 3)), "; e2e full", ef and (round(ef["value"]), round(ef["ms_per_step"], 3)))
except Exception as ex:
    print("$1 failed", ex)
PY
}
run x1tma_b4096 WD_T4_X1_TMA=1 4096 20 ""
run gather_b4096 WD_T4_X1_TMA=0 4096 20 --no-e2e
run x1tma2_b4096 WD_T4_X1_TMA=1 4096 20 --no-e2e
run x1tma_b512 WD_T4_X1_TMA=1 512 50 ""
run gather_b512 WD_T4_X1_TMA=0 512 50 --no-e2e
run x1tma_b50 WD_T4_X1_TMA=1 50 100 ""
run gather_b50 WD_T4_X1_TMA=0 50 100 --no-e2e
