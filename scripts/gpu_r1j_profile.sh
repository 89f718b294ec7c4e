#!/bin/bash
# Session J: tests + smoke + benches + reference arm, then (each only after its own command exited 0 without ncu)
# the ncu launch list and one `ncu --set full` capture of the hot kernels of a bench step.
set -u
mkdir -p gpurun_out
T=${WD_TAG:-r1_k}
timeout 1200 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${T}_smoke.log
timeout 900 python bench.py --steps 30 --warmup 5 > gpurun_out/${T}_bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -2 gpurun_out/bench.err
timeout 300 python bench.py --steps 100 --warmup 10 --batch 50 --no-cpu-baseline > gpurun_out/${T}_bench_b50.json 2> gpurun_out/bench_b50.err; echo "bench b50 rc=$?"
timeout 300 python bench.py --steps 50 --warmup 10 --batch 512 --no-cpu-baseline > gpurun_out/${T}_bench_b512.json 2> gpurun_out/bench_b512.err; echo "bench b512 rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 300 python bench.py --steps 20 --warmup 5 --kind qm9 --no-cpu-baseline > gpurun_out/${T}_bench_qm9_b4096.json 2> gpurun_out/bench_qm9.err; echo "qm9 rc=$?"
timeout 600 python bench.py --steps 5 --warmup 3 --hidden 2400 --depth 6 --batch 8192 --no-cpu-baseline --no-e2e > gpurun_out/${T}_bench_h2400_d6_b8192.json 2> gpurun_out/bench_h2400.err; echo "h2400 rc=$?"
timeout 600 python bench.py --steps 5 --warmup 3 --math fp32 --no-cpu-baseline --no-e2e > gpurun_out/${T}_bench_fp32.json 2> gpurun_out/bench_fp32.err; echo "fp32 rc=$?"
python - <<PY
import json
for f in ("${T}_bench", "${T}_bench_b512", "${T}_bench_b50", "${T}_bench_qm9_b4096", "${T}_bench_h2400_d6_b8192", "${T}_bench_fp32"):
    try:
        d = json.load(open(f"gpurun_out/{f}.json"))
    except Exception as e:
        print(f, "unreadable", e); continue
    e, ef = d.get("e2e"), d.get("e2e_full_tensors")
    print(f, round(d["value"]), "graphs/s", round(d["ms_per_step"], 4), "ms; with events", round(d.get("ms_per_step_with_events", 0), 4), "; e2e", e and (round(e["value"]), round(e["ms_per_step"], 3), e["h2d_bytes_per_step"]),
          "; e2e full", ef and (round(ef["value"]), round(ef["ms_per_step"], 3)), "; cpu", d.get("cpu_baseline") and round(d["cpu_baseline"]["value"]), "; roofline", d["roofline"]["kernel"], round(d["roofline"]["frac"], 3))
    print("   ", " ".join(f"{k}={v['ms_per_launch']:.3f}x{v['launches_per_step']:.0f}" for k, v in d["kernels"].items()))
PY
echo "== ncu launch list"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_ncu_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu.log 2>&1; echo "ncu rc=$?"
echo "== ncu --set full"
timeout 1200 ncu --set full --clock-control none --import-source on \
  -k 'regex:k_tc_gemm4|k_tc_gemm3|k_tc_dw5|k_message_grad|k_aggregate|k_readout|k_reduce_partials|k_colsum' --launch-skip 60 -c 32 \
  -o /tmp/${T}_full -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -2 gpurun_out/ncu_full.log | cut -c1-200
ncu -i /tmp/${T}_full.ncu-rep --page raw --csv > gpurun_out/${T}_ncu_full_raw.csv 2>/dev/null; wc -l gpurun_out/${T}_ncu_full_raw.csv
python scripts/ncu_summary.py gpurun_out/${T}_ncu_full_raw.csv > gpurun_out/${T}_ncu_full_table.md; head -3 gpurun_out/${T}_ncu_full_table.md; ls -la /tmp/${T}_full.ncu-rep; du -sh gpurun_out
