#!/bin/bash
# Session J run 3: GPU tests (incl. full-model parity), bench at B=4096 / 512 / 50, e2e probe.
set -u
mkdir -p gpurun_out
echo "== pytest -m gpu"; timeout 1200 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -2 gpurun_out/bench.err
timeout 300 python bench.py --steps 100 --warmup 10 --batch 50 --no-cpu-baseline > gpurun_out/bench_b50.json 2> gpurun_out/bench_b50.err; echo "bench b50 rc=$?"
timeout 300 python bench.py --steps 50 --warmup 10 --batch 512 --no-cpu-baseline > gpurun_out/bench_b512.json 2> gpurun_out/bench_b512.err; echo "bench b512 rc=$?"
python - <<'PY'
import json
for f in ("bench", "bench_b512", "bench_b50"):
    try:
        d = json.load(open(f"gpurun_out/{f}.json"))
    except Exception as e:
        print(f, "unreadable", e); continue
    e, ef = d.get("e2e"), d.get("e2e_full_tensors")
    print(f, round(d["value"]), "graphs/s", round(d["ms_per_step"], 4), "ms; with events", round(d.get("ms_per_step_with_events", 0), 4), "; e2e", e and (round(e["value"]), round(e["ms_per_step"], 3), e["h2d_bytes_per_step"]),
          "; e2e full", ef and (round(ef["value"]), round(ef["ms_per_step"], 3), ef["h2d_bytes_per_step"]), "; cpu", d.get("cpu_baseline") and round(d["cpu_baseline"]["value"]))
    print("   ", " ".join(f"{k}={v['ms_per_launch']:.3f}x{v['launches_per_step']:.0f}" for k, v in d["kernels"].items()))
PY
WD_AGG_CSR=1 timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/bench_csr.json 2>/dev/null; python -c "import json; d=json.load(open(\"gpurun_out/bench_csr.json\")); print(\"CSR lists:\", round(d[\"ms_per_step\"],4), {k: round(v[\"ms_per_launch\"],4) for k,v in d[\"kernels\"].items() if k in (\"aggregate\",\"scatter\")})"
timeout 300 python -m pytest tests/test_gpu_model.py -q -s 2>&1 | grep "prediction rel err"
