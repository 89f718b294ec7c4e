#!/bin/bash
# Round-1 session J: GPU parity tests, PDL on/off A/B at B=4096 and B=50, compact-upload e2e.
set -u
mkdir -p gpurun_out
echo "== pytest -m gpu"; timeout 1200 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
echo "== smoke"; timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
for pdl in 1 0; do
  WD_PDL=$pdl timeout 600 python bench.py --steps 20 --warmup 5 $( [ $pdl = 0 ] && echo --no-cpu-baseline --no-e2e ) > gpurun_out/bench_pdl$pdl.json 2> gpurun_out/bench_pdl$pdl.err; echo "bench pdl=$pdl rc=$?"; tail -2 gpurun_out/bench_pdl$pdl.err
  WD_PDL=$pdl timeout 300 python bench.py --steps 100 --warmup 10 --batch 50 --no-cpu-baseline $( [ $pdl = 0 ] && echo --no-e2e ) > gpurun_out/bench_b50_pdl$pdl.json 2> gpurun_out/bench_b50_pdl$pdl.err; echo "bench b50 pdl=$pdl rc=$?"
done
python - <<'PY'
import json
for f in ("bench_pdl1", "bench_pdl0", "bench_b50_pdl1", "bench_b50_pdl0"):
    try:
        d = json.load(open(f"gpurun_out/{f}.json"))
    except Exception as e:
        print(f, "unreadable", e); continue
    e, ef = d.get("e2e"), d.get("e2e_full_tensors")
    print(f, round(d["value"]), "graphs/s", round(d["ms_per_step"], 4), "ms; with events", round(d.get("ms_per_step_with_events", 0), 4), "; e2e", e and (round(e["value"]), round(e["ms_per_step"], 3), e["h2d_bytes_per_step"]),
          "; e2e full", ef and (round(ef["value"]), round(ef["ms_per_step"], 3), ef["h2d_bytes_per_step"]), "; cpu", d.get("cpu_baseline") and round(d["cpu_baseline"]["value"]))
    print("   ", " ".join(f"{k}={v['ms_per_launch']:.3f}x{v['launches_per_step']:.0f}" for k, v in d["kernels"].items()))
PY
