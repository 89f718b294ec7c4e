#!/bin/bash
# Scaling run on one box: bench at N = 1, 2, 4, ... up to $1 GPUs (torchrun, NCCL), as the driver launches it.
set -u
NMAX=${1:-8}
mkdir -p gpurun_out
nvidia-smi -L
for N in 1 2 4 8; do
  [ $N -le $NMAX ] || continue
  if [ $N -eq 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
  fi
  echo "n$N rc=$?"; tail -2 gpurun_out/scale_n$N.err
done
python - <<PY
import json
base = None
for n in (1, 2, 4, 8):
    try:
        d = json.loads([l for l in open(f'gpurun_out/scale_n{n}.json') if l.startswith('{')][-1])
        base = base or d['value']
        print(n, 'GPUs:', round(d['value']), 'graphs/s', round(d['ms_per_step'], 3), 'ms/step; e2e', d['e2e'] and round(d['e2e']['value']),
              d['config']['parallelism'], 'x%.2f' % (d['value'] / base))
    except Exception as e:
        print(n, 'unreadable', e)
PY
