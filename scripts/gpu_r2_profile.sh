#!/bin/bash
# Round 2 evidence: config-4 bench (tensor roofline), ncu launch list and ncu --set full of one whole step at HEAD.
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e"
echo "== config 4"; timeout 900 python bench.py --hidden 2400 --depth 6 --batch 8192 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r2_bench_h2400_d6_b8192.json 2> gpurun_out/r2_bench_h2400.err; echo "rc=$?"; tail -2 gpurun_out/r2_bench_h2400.err
echo "== qm9"; timeout 600 python bench.py --kind qm9 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2_bench_qm9.json 2> gpurun_out/r2_bench_qm9.err; echo "rc=$?"
echo "== plain"; $CMD > gpurun_out/ncu_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2_ncu_launches.csv $CMD > gpurun_out/ncu_list.log 2>&1; echo "list rc=$?"
$CMD > gpurun_out/ncu_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -s 85 -c 26 -o gpurun_out/r2_step_full $CMD > gpurun_out/ncu_full.log 2>&1; echo "full rc=$?"; tail -2 gpurun_out/ncu_full.log | cut -c1-200
ls -la gpurun_out | head -30
python - <<'PY'
import json
for f in ("r2_bench_h2400_d6_b8192", "r2_bench_qm9"):
    d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    print(f, round(d["value"]), "graphs/s", round(d["ms_per_step"], 3), "ms", d["roofline"]["bound"], round(d["roofline"]["frac"], 3), d.get("step_tensor"), "e2e", d["e2e"] and round(d["e2e"]["value"]))
    print("   ", " ".join(f"{k}={v['ms_per_launch']:.3f}x{v['launches_per_step']:.0f}" for k, v in d["kernels"].items()))
PY
