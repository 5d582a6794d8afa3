#!/bin/bash
# Multi-GPU run (gpurun --gpus 8): bench.py at N = 8, 4, 2 under torchrun (kept-only result set), the full-slab result set
# at N = 8, and the host-thread probe at N = 8 (1 and 2 threads per rank instead of 3: is the e2e leg host-bound?).
# With AGG=1 first the box's aggregate pinned-copy rates (tools/h2d_aggregate.py).  Usage: tools/gpu_scale.sh <tag>
set -u
TAG=${1:-r2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/${TAG}_topo.txt 2>&1
nproc >> gpurun_out/${TAG}_topo.txt
if [ "${AGG:-0}" = 1 ]; then
  python tools/h2d_aggregate.py > gpurun_out/${TAG}_h2d_aggregate.jsonl 2> gpurun_out/${TAG}_scale.err; cat gpurun_out/${TAG}_h2d_aggregate.jsonl
fi
: > gpurun_out/${TAG}_scale.jsonl
run() {  # N, extra args...
  local N=$1; shift
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) \
    bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline "$@" 2>> gpurun_out/${TAG}_scale.err | tail -1 >> gpurun_out/${TAG}_scale.jsonl
}
run 8 --e2e-out kept
run 4 --e2e-out kept
run 2 --e2e-out kept
run 8 --e2e-out full
[ "${PROBE:-0}" = 1 ] && run 8 --e2e-out kept --e2e-threads 1
[ "${PROBE:-0}" = 1 ] && run 8 --e2e-out kept --e2e-threads 2
python - <<PY
import json
for l in open('gpurun_out/${TAG}_scale.jsonl'):
    try:
        d = json.loads(l)
        print(d['n_gpus'], d['e2e']['out'], 'threads', d['e2e']['host_threads'], 'value %.0f e2e %.0f' % (d['value'], d['e2e']['value']))
    except Exception as e:
        print('bad line', l[:200])
PY
