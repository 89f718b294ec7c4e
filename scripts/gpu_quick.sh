#!/bin/bash
# Quick GPU loop: probe + TF32 parity tests + role timing + bench (no CPU baseline).
set -u
mkdir -p gpurun_out
timeout 300 python scripts/tc_probe.py > gpurun_out/tc_probe.log 2>&1; echo "probe rc=$?"; grep -v "^  " gpurun_out/tc_probe.log | tail -5
timeout 600 python -m pytest tests/test_gpu_tf32.py -q -m gpu -x 2>&1 | tail -3
if [ -f polymer_chemprop_b200/libwdmpnn_timing.so ]; then
  WDMPNN_LIB=/root/repo/polymer_chemprop_b200/libwdmpnn_timing.so timeout 300 python scripts/t2_timing.py > gpurun_out/t2_timing.log 2>&1; cat gpurun_out/t2_timing.log
fi
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline ${WD_BENCH_ARGS:-} > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_quick.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_quick.json"))
print(round(d["value"]), "graphs/s", round(d["ms_per_step"], 3), "ms; e2e", d.get("e2e") and round(d["e2e"]["value"]))
print("   ", " ".join(f"{k}={v['ms_per_launch']:.3f}x{v['launches_per_step']:.0f}" for k, v in d["kernels"].items()))
PY
for v in ${WD_VARIANTS:-}; do
  lib=/root/repo/polymer_chemprop_b200/libwdmpnn_$v.so
  [ -f "$lib" ] || continue
  WDMPNN_LIB=$lib timeout 300 python scripts/tc_probe.py 2>&1 | tail -1
  WDMPNN_LIB=$lib timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/bench_$v.json 2> gpurun_out/bench_$v.err
  python - <<PY
import json
d = json.load(open("gpurun_out/bench_$v.json"))
print("variant $v:", round(d["value"]), "graphs/s", round(d["ms_per_step"], 3), "ms")
print("   ", " ".join(f"{k}={v['ms_per_launch']:.3f}x{v['launches_per_step']:.0f}" for k, v in d["kernels"].items()))
PY
done
