#!/bin/bash
# BASELINE config 4 on one multi-GPU box (gpurun --gpus 8): the 10 000-pair EuRoC sequence on 1, 2, 4, 8 GPUs.
mkdir -p gpurun_out
: > gpurun_out/${1:-r2}_sequence.jsonl
python tools/bench_sequence.py --pairs 10000 --check 600 2>/dev/null | tail -1 >> gpurun_out/${1:-r2}_sequence.jsonl
for N in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29540 + N)) \
    tools/bench_sequence.py --pairs 10000 --check 600 2>/dev/null | tail -1 >> gpurun_out/${1:-r2}_sequence.jsonl
done
python - <<PY
import json
for l in open('gpurun_out/${1:-r2}_sequence.jsonl'):
    d = json.loads(l); print(d['n_gpus'], 'pairs/s %.0f' % d['value'], 'seconds %.4f' % d['seconds'], d.get('sharded_equals_sequential'))
PY
