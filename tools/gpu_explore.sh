#!/bin/bash
# exploration run: stream-count sweep, per-kernel breakdown at batch 1 and at the 4K stress shape, triangulation line
mkdir -p gpurun_out
summ() { python - "$1" "$2" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[2], 'value=%.0f ms/step=%.3f' % (d['value'], d['ms_per_step']), 'e2e=%s' % (d['e2e'] and round(d['e2e']['value'])))
    print('   ', ' '.join('%s=%.2f' % kv for kv in d['roofline']['kernel_us_per_pair'].items()))
except Exception as e:
    print(sys.argv[2], 'FAILED', e)
PY
}
for s in ${STREAMS:-2}; do
  python bench.py --steps 10 --warmup 3 --no-cpu-baseline --skip-e2e --streams $s > gpurun_out/x_streams$s.json 2> gpurun_out/x_err.log; summ gpurun_out/x_streams$s.json "kitti streams=$s"
done
python bench.py --steps 50 --warmup 5 --no-cpu-baseline --skip-e2e --streams 1 --pairs-per-step 1 --pool 8 > gpurun_out/x_b1.json 2>> gpurun_out/x_err.log; summ gpurun_out/x_b1.json "kitti batch1"
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --skip-e2e --streams 1 --workload 4k --pairs-per-step 8 --pool 8 > gpurun_out/x_4k.json 2>> gpurun_out/x_err.log; summ gpurun_out/x_4k.json "4k batch8"
python bench.py --steps 20 --warmup 3 --workload 4k --pairs-per-step 1 --pool 4 --no-cpu-baseline --skip-e2e --streams 1 > gpurun_out/x_4k_b1.json 2>> gpurun_out/x_err.log; summ gpurun_out/x_4k_b1.json "4k batch1"
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --triangulate > gpurun_out/x_tri.json 2>> gpurun_out/x_err.log; summ gpurun_out/x_tri.json "kitti +triangulate"
tail -5 gpurun_out/x_err.log
