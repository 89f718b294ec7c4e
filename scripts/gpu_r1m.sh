#!/bin/bash
# Session M: the round-end sequence as the driver runs it (tests, smoke, both bench arms) + the reworked clock sampler.
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/m_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/m_smoke.log | cut -c1-300
timeout 600 python bench.py --impl reference > gpurun_out/m_ref.json 2> gpurun_out/m_ref.err; echo "ref rc=$?"; cut -c1-400 gpurun_out/m_ref.json
timeout 600 python bench.py > gpurun_out/m_bench.json 2> gpurun_out/m_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/m_bench.err
python - <<PY
import json
d = json.load(open("gpurun_out/m_bench.json"))
print(round(d["value"]), "graphs/s", round(d["ms_per_step"], 4), "ms; e2e", round(d["e2e"]["value"]), "clocks", d["clocks"], "roofline", d["roofline"], "cpu", d["cpu_baseline"], "launches", d["gpu_launches"])
PY
