#!/bin/bash
# Session P: fused optimiser step (wd_adam_step / FlatAdam) -- tests + one bench line with optimizer_step.
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log | cut -c1-300
timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/p_bench.json 2> gpurun_out/p_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/p_bench.err
python - <<PY
import json
d = json.load(open("gpurun_out/p_bench.json"))
print(round(d["value"]), "graphs/s", round(d["ms_per_step"], 4), "ms; e2e", round(d["e2e"]["value"]), "optimizer", d["optimizer_step"])
PY
