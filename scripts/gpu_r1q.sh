#!/bin/bash
# Session Q: per-step wall times of the e2e legs (outliers), optimiser tests after the fused gradient pack.
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_optim.py -q -m gpu -x > gpurun_out/pytest_optim.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_optim.log | cut -c1-300
for i in 1 2; do
timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/q_bench$i.json 2> gpurun_out/q_bench$i.err; echo "bench rc=$?"; tail -3 gpurun_out/q_bench$i.err
python - <<PY
import json
d = json.load(open("gpurun_out/q_bench$i.json"))
print(round(d["value"]), "graphs/s", round(d["ms_per_step"], 4), "optimizer", {k: (round(v, 1) if isinstance(v, float) else v) for k, v in d["optimizer_step"].items() if k != "note"})
for k in ("e2e","e2e_full_tensors","e2e_device_store","e2e_host_waits_for_backward"):
    e=d[k]; print(" ", k, round(e["value"]), round(e["ms_per_step"],3), round(e["ms_per_step_median"],3), e["step_ms"])
PY
done
