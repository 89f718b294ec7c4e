#!/bin/bash
# Round 2: parity tests, bench at B=4096 / B=50, forward A/B after a kernel change.
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log | cut -c1-250
timeout 300 python bench.py --steps 200 --warmup 20 --batch 50 --no-cpu-baseline > gpurun_out/bench_b50.json 2> gpurun_out/bench_b50.err; echo "rc=$?"
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/bench_b4096.json 2> gpurun_out/bench_b4096.err; echo "rc=$?"
python - <<'PY'
import json
for f in ("bench_b50", "bench_b4096"):
    d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    print(f, round(d["value"]), "graphs/s", round(d["ms_per_step"], 4), "ms; infer", round(d["inference"]["value"]), "e2e", d["e2e"] and round(d["e2e"]["value"]))
    print("   ", " ".join(f"{k}={v['ms_per_launch']*1e3:.1f}x{v['launches_per_step']:.0f}" for k, v in d["kernels"].items()))
PY
timeout 300 python scripts/fwd_ab.py 2>&1 | tail -2 | tee -a gpurun_out/fwd_ab.log
