#!/bin/bash
# One `ncu --set full` capture of the hot kernels (one launch each) of a bench step; raw CSV for profiles/.
set -u
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:k_tc_gemm4|k_tc_gemm3|k_tc_dw5|k_message_grad|k_aggregate' -c 14 \
  -o gpurun_out/full_r1h -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_full.log | cut -c1-200
ncu -i gpurun_out/full_r1h.ncu-rep --page raw --csv > gpurun_out/full_r1h_raw.csv 2>/dev/null; wc -l gpurun_out/full_r1h_raw.csv
