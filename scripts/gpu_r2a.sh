#!/bin/bash
# Round 2, call A: GPU parity tests (all modes), smoke, short benches of the tensor-core modes.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu.txt 2>&1
echo "== pytest -m gpu" ; timeout 1500 python -m pytest tests -q -m gpu -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" ; tail -40 gpurun_out/pytest_gpu.log | cut -c1-300
echo "== smoke" ; timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -5 gpurun_out/smoke.log
for m in tf32x3 tf32; do
  echo "== bench $m" ; timeout 600 python bench.py --steps 20 --warmup 5 --math $m --no-cpu-baseline --no-e2e > gpurun_out/bench_$m.json 2> gpurun_out/bench_$m.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_$m.err
done
python - <<'PY'
import json
for f in ("bench_tf32x3", "bench_tf32"):
    try:
        d = json.load(open(f"gpurun_out/{f}.json"))
    except Exception as e:
        print(f, "unreadable", e); continue
    print(f, round(d["value"]), "graphs/s", round(d["ms_per_step"], 3), "ms; with events", round(d["ms_per_step_with_events"], 3))
    print("   ", " ".join(f"{k}={v['ms_per_launch']:.3f}x{v['launches_per_step']:.0f}" for k, v in d["kernels"].items()))
PY
