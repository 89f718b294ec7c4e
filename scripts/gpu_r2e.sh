#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -p no:cacheprovider -k "cuda_graph or host_buffer or golden_fp32" 2>&1 | tail -15
for g in 1 0; do
  WD_GRAPH=$g timeout 300 python bench.py --steps 200 --warmup 20 --batch 50 --no-cpu-baseline > gpurun_out/bench_b50_g$g.json 2> gpurun_out/bench_b50_g$g.err; echo "rc=$?"; tail -2 gpurun_out/bench_b50_g$g.err
done
python - <<'PY'
import json
for g in (1, 0):
    d = json.loads(open(f"gpurun_out/bench_b50_g{g}.json").read().strip().splitlines()[-1])
    print("graph", g, round(d["value"]), "graphs/s", round(d["ms_per_step"], 4), "ms; e2e", round(d["e2e"]["value"]), round(d["e2e"]["ms_per_step"], 3), "median", round(d["e2e"]["ms_per_step_median"], 3), "store", round(d["e2e_device_store"]["value"]), "infer", round(d["inference"]["value"]), "train", d["train_step"] and round(d["train_step"]["value"]))
PY
