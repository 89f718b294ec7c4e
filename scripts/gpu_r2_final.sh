#!/bin/bash
# Round-2 record run: GPU parity tests, smoke, the default bench line (timed by the wall clock too), reference arm,
# B=50, and -- last, after the same command exited 0 without it -- ONE ncu launch list.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu.txt 2>&1
echo "== pytest -m gpu"; timeout 1500 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log | cut -c1-200
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log | cut -c1-300
echo "== bench (defaults)"; t0=$(date +%s); timeout 1200 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$? wall=$(( $(date +%s) - t0 ))s"; tail -2 gpurun_out/bench.err
echo "== reference arm"; t0=$(date +%s); timeout 900 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "rc=$? wall=$(( $(date +%s) - t0 ))s"
echo "== B=50"; timeout 300 python bench.py --steps 200 --warmup 20 --batch 50 --no-cpu-baseline > gpurun_out/bench_b50.json 2> gpurun_out/bench_b50.err; echo "rc=$?"
python - <<'PY'
import json
for f in ("bench", "bench_b50", "bench_ref"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "unreadable", e); continue
    print(f, round(d["value"]), "graphs/s", round(d["ms_per_step"], 4), "ms; e2e", d.get("e2e") and round(d["e2e"]["value"]),
          "; cpu", d.get("cpu_baseline") and round(d["cpu_baseline"]["value"]), "; roofline", d.get("roofline") and (d["roofline"]["kernel"], round(d["roofline"]["frac"], 3)))
    if d.get("small_batch"): print("    small_batch", {k: (round(v, 1) if isinstance(v, float) else v) for k, v in d["small_batch"].items() if not isinstance(v, dict)}, "e2e", d["small_batch"].get("e2e"))
    if d.get("kernels"): print("   ", " ".join(f"{k}={v['ms_per_launch']*1e3:.1f}x{v['launches_per_step']:.0f}" for k, v in d["kernels"].items()))
PY
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-small-batch --no-single-pass"
echo "== ncu launch list"; $CMD > gpurun_out/ncu_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_o_ncu_launches.csv $CMD > gpurun_out/ncu_list.log 2>&1; echo "list rc=$?"
