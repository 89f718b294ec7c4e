#!/bin/bash
# Round 2: GPU parity tests + A/B timing of the library in the tree.
set -u
mkdir -p gpurun_out
echo "== pytest -m gpu" ; timeout 1500 python -m pytest tests -q -m gpu -p no:cacheprovider ${PYTEST_ARGS:-} > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" ; tail -25 gpurun_out/pytest_gpu.log | cut -c1-250
echo "== A/B"; timeout 300 python scripts/fwd_ab.py 2>&1 | tail -2 | tee -a gpurun_out/fwd_ab.log
