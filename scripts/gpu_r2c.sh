#!/bin/bash
# Round 2: the no-allocation entry point's test alone in a fresh process, then the whole GPU suite (it runs last), smoke, bench.
set -u
mkdir -p gpurun_out
echo "== zz alone"; timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu -p no:cacheprovider 2>&1 | tail -3
echo "== pytest -m gpu -x"; timeout 1500 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu.log | cut -c1-250
echo "== smoke"; timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -5 gpurun_out/smoke.log
echo "== bench"; timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
echo "== ref"; timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; tail -3 gpurun_out/bench_ref.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench.json"))
print(round(d["value"]), "graphs/s", round(d["ms_per_step"], 3), "ms; e2e", round(d["e2e"]["value"]), "full", round(d["e2e_full_tensors"]["value"]))
print("roofline", {k: d["roofline"][k] for k in ("kernel", "bound", "frac") }, d["roofline"].get("frac_contract"), "step", d["step_hbm"]["frac"], d["step_hbm"]["frac_contract"])
print(" ".join(f"{k}={v['ms_per_launch']:.3f}x{v['launches_per_step']:.0f}" for k, v in d["kernels"].items()))
cb = d["cpu_baseline"]; print("cpu", round(cb["value"]), {k: round(v["value"]) for k, v in cb["other_figures"].items()})
r = json.load(open("gpurun_out/bench_ref.json")); print("ref", round(r["value"]), r["steps"], r["warmup"], {k: round(v["value"]) for k, v in r["cpu_baseline"]["other_figures"].items()}, r["config"] == d["config"])
PY
