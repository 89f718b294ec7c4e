#!/bin/bash
# N-GPU run: bench at N=1 and N=$1 (torchrun, NCCL), as the driver launches it.
set -u
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err; echo "n1 rc=$?"; tail -2 gpurun_out/scale_n1.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err; echo "n$N rc=$?"; tail -3 gpurun_out/scale_n$N.err
python - <<PY
import json
for n in (1, $N):
    try:
        d=json.loads([l for l in open(f'gpurun_out/scale_n{n}.json') if l.startswith('{')][-1])
        print(n, 'GPUs:', round(d['value']), 'graphs/s', round(d['ms_per_step'],3), 'ms/step; e2e', d['e2e'] and round(d['e2e']['value']), d['config']['parallelism'])
    except Exception as e: print(n, 'unreadable', e)
PY
