#!/bin/bash
# Session L: device-resident store -- tests and bench lines with e2e_device_store.
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_gpu.log | cut -c1-300
for b in 4096 512 50; do
  s=20; [ $b = 50 ] && s=100; [ $b = 512 ] && s=50
  timeout 400 python bench.py --steps $s --warmup 5 --batch $b --no-cpu-baseline > gpurun_out/l_b$b.json 2> gpurun_out/l_b$b.err; echo "bench b$b rc=$?"; tail -2 gpurun_out/l_b$b.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/l_b$b.json"))
    f = lambda e: e and (round(e["value"]), round(e["ms_per_step"], 3), round(e["ms_per_step_median"], 3), e["h2d_bytes_per_step"])
    print("b$b", round(d["value"]), "graphs/s", round(d["ms_per_step"], 4), "ms; e2e", f(d.get("e2e")), "; full", f(d.get("e2e_full_tensors")), "; store", f(d.get("e2e_device_store")))
except Exception as ex:
    print("b$b failed", ex)
PY
done
