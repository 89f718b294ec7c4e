#!/bin/bash
# Round-2 scaling record (run under gpurun --gpus N): the default bench line at N GPUs, as the driver launches it.
set -u
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2_e_scale_n$N.json 2> gpurun_out/r2_e_scale_n$N.err
echo "rc=$?"; tail -3 gpurun_out/r2_e_scale_n$N.err | cut -c1-200
python - <<PY
import json
d = json.loads(open("gpurun_out/r2_e_scale_n$N.json").read().strip().splitlines()[-1])
print("N=$N", round(d["value"]), "graphs/s", round(d["ms_per_step"], 3), "ms; e2e", d["e2e"] and round(d["e2e"]["value"]), "allreduce", d.get("allreduce"), "train", d.get("train_step") and round(d["train_step"]["value"]))
PY
