#!/bin/bash
# Session O: message gradient as a block-local flat loop (no index division) -- tests + bench at three batch sizes.
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log | cut -c1-300
for b in 4096 512 50; do
  s=30; [ $b = 50 ] && s=200; [ $b = 512 ] && s=100
  timeout 400 python bench.py --steps $s --warmup 5 --batch $b --no-cpu-baseline > gpurun_out/o_b$b.json 2> gpurun_out/o_b$b.err; echo "bench b$b rc=$?"; tail -2 gpurun_out/o_b$b.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/o_b$b.json"))
    f = lambda e: e and (round(e["value"]), round(e["ms_per_step"], 3), round(e["ms_per_step_median"], 3))
    print("b$b", round(d["value"]), "graphs/s", round(d["ms_per_step"], 4), "ms; e2e", f(d["e2e"]), "full", f(d["e2e_full_tensors"]), "store", f(d["e2e_device_store"]), "step-sync", f(d["e2e_host_waits_for_backward"]), "roof", d["roofline"]["kernel"], round(d["roofline"]["frac"], 3))
    print("   kernels", {k: round(v["ms_per_launch"], 4) for k, v in d["kernels"].items()})
except Exception as ex:
    print("b$b failed", ex)
PY
done
