#!/bin/bash
# Multi-GPU run (gpurun --gpus 8): the box's aggregate pinned-copy rates, then bench.py at N = 8, 4, 2 under torchrun with
# both e2e result sets.  Usage: tools/gpu_scale.sh <tag>
set -u
TAG=${1:-r2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/${TAG}_topo.txt 2>&1
nproc >> gpurun_out/${TAG}_topo.txt
python tools/h2d_aggregate.py > gpurun_out/${TAG}_h2d_aggregate.jsonl 2> gpurun_out/${TAG}_scale.err; cat gpurun_out/${TAG}_h2d_aggregate.jsonl
: > gpurun_out/${TAG}_scale.jsonl
for N in 8 4 2; do
  for OUT in kept full; do
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) \
      bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline --e2e-out $OUT 2>> gpurun_out/${TAG}_scale.err | tail -1 >> gpurun_out/${TAG}_scale.jsonl
  done
done
python - <<PY
import json
for l in open('gpurun_out/${TAG}_scale.jsonl'):
    try:
        d = json.loads(l)
        print(d['n_gpus'], d['e2e']['out'], 'value %.0f e2e %.0f' % (d['value'], d['e2e']['value']))
    except Exception as e:
        print('bad line', l[:200])
PY
