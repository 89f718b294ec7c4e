#!/bin/bash
# One gpurun call: GPU parity tests, smoke, benches (B=4096 and B=50), ncu launch list.
# Usage (from the build container): gpurun --timeout 2400 -- 'bash scripts/gpu_check.sh'
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu.txt 2>&1
echo "== pytest -m gpu" ; timeout 1200 python -m pytest tests -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" ; tail -6 gpurun_out/pytest_gpu.log
echo "== smoke" ; timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/smoke.log
echo "== bench" ; timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
echo "== bench B=50" ; timeout 300 python bench.py --steps 100 --warmup 10 --batch 50 --no-cpu-baseline > gpurun_out/bench_b50.json 2> gpurun_out/bench_b50.err; echo "rc=$?"
echo "== bench fp32" ; timeout 600 python bench.py --steps 5 --warmup 3 --math fp32 --no-cpu-baseline --no-e2e > gpurun_out/bench_fp32.json 2> gpurun_out/bench_fp32.err; echo "rc=$?"
echo "== bench reference arm" ; timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "rc=$?"; cat gpurun_out/bench_ref.json | cut -c1-400
python - <<'PY'
import json
for f in ("bench", "bench_b50", "bench_fp32"):
    try:
        d = json.load(open(f"gpurun_out/{f}.json"))
    except Exception as e:
        print(f, "unreadable", e); continue
    print(f, round(d["value"]), "graphs/s", round(d["ms_per_step"], 3), "ms; e2e", d.get("e2e") and round(d["e2e"]["value"]),
          "; cpu", d.get("cpu_baseline") and round(d["cpu_baseline"]["value"]), "; roofline", d["roofline"]["kernel"], round(d["roofline"]["frac"], 3))
    print("   ", " ".join(f"{k}={v['ms_per_launch']:.3f}x{v['launches_per_step']:.0f}" for k, v in d["kernels"].items()))
PY
if [ "${WD_NCU:-1}" = "1" ]; then
  echo "== ncu launch list"
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
      --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu.log 2>&1
  echo "ncu rc=$?"; tail -2 gpurun_out/ncu.log | cut -c1-200
fi
