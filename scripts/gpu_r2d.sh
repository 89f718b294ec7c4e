#!/bin/bash
# Round 2: store test, N=2 data-parallel bench (gradient sink, all-reduce share, store-fed e2e, train step)
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_store.py tests/test_gpu_optim.py -x -q -m gpu -p no:cacheprovider 2>&1 | tail -3
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "rc=$?"; tail -5 gpurun_out/bench_n2.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_n2.json"))
print(round(d["value"]), "graphs/s", round(d["ms_per_step"], 3), "ms; e2e", round(d["e2e"]["value"]), d["e2e"].get("store_graphs_per_rank"), "host upload", round(d["e2e_host_upload"]["value"]))
print("allreduce", d["allreduce"]); print("train_step", d["train_step"]); print("launches", d["gpu_launches"])
PY
