#!/bin/bash
# 8-GPU e2e probes (gpurun --gpus 8): does the host-side pipeline shape move the e2e rate at 8 GPUs?
mkdir -p gpurun_out
: > gpurun_out/${1:-r2}_scale_probe.jsonl
run() {
  local tag=$1; shift
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus 8 --steps 6 --warmup 3 --no-cpu-baseline --e2e-out kept "$@" 2>> gpurun_out/r2_scale_probe.err | tail -1 | \
    python -c "import json,sys; d=json.loads(sys.stdin.read()); print(json.dumps({'probe': '$tag', 'value': d['value'], 'e2e': d['e2e']['value'], 'threads': d['e2e']['host_threads'], 'pairs_per_call': d['e2e']['pairs_per_call']}))" | tee -a gpurun_out/r2_scale_probe.jsonl
}
run "default (3 threads, chunk 64, 3 lanes)"
run "chunk 32, 4 lanes" --debug-set pipe_chunk=32 --debug-set pipe_lanes=4
run "256 pairs per call" --pairs-per-step 256 --pool 320
run "256 pairs per call, 2 threads" --pairs-per-step 256 --pool 320 --e2e-threads 2
