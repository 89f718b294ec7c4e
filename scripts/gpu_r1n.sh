#!/bin/bash
# Session N: sign bits for the message gradient -- parity (bitwise A/B), bench A/B at three batch sizes.
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu.log | cut -c1-300
for v in 1 0; do
for b in 4096 512 50; do
  s=30; [ $b = 50 ] && s=200; [ $b = 512 ] && s=100
  WD_SIGN_BITS=$v timeout 400 python bench.py --steps $s --warmup 5 --batch $b --no-cpu-baseline > gpurun_out/n_sign${v}_b$b.json 2> gpurun_out/n_sign${v}_b$b.err; echo "bench sign=$v b$b rc=$?"; tail -2 gpurun_out/n_sign${v}_b$b.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/n_sign${v}_b$b.json"))
    f = lambda e: e and (round(e["value"]), round(e["ms_per_step"], 3), round(e["ms_per_step_median"], 3))
    print("sign=$v b$b", round(d["value"]), "graphs/s", round(d["ms_per_step"], 4), "ms; e2e", f(d["e2e"]), "full", f(d["e2e_full_tensors"]), "store", f(d["e2e_device_store"]), "step-sync", f(d["e2e_host_waits_for_backward"]), "clocks", d["clocks"])
    print("   kernels", {k: round(v["ms_per_launch"], 4) for k, v in d["kernels"].items()})
except Exception as ex:
    print("b$b failed", ex)
PY
done
done
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/n_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/n_smoke.log | cut -c1-300
