#!/bin/bash
# One gpurun call for the tensor-core path: contraction probe, TF32 parity tests, short benches.
set -u
mkdir -p gpurun_out
echo "== probe"; timeout 300 python scripts/tc_probe.py > gpurun_out/tc_probe.log 2>&1; echo "probe rc=$?"; tail -40 gpurun_out/tc_probe.log
echo "== pytest tf32"; timeout 900 python -m pytest tests/test_gpu_tf32.py -q -m gpu -x > gpurun_out/pytest_tf32.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_tf32.log
echo "== bench tf32"; timeout 600 python bench.py --steps 10 --warmup 3 --math tf32 --no-cpu-baseline > gpurun_out/bench_tf32.json 2> gpurun_out/bench_tf32.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_tf32.err; cat gpurun_out/bench_tf32.json
echo "== bench tf32 B=50"; timeout 300 python bench.py --steps 50 --warmup 5 --batch 50 --math tf32 --no-cpu-baseline > gpurun_out/bench_tf32_b50.json 2> gpurun_out/bench_tf32_b50.err; echo "rc=$?"; cat gpurun_out/bench_tf32_b50.json
