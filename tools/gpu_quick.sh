#!/bin/bash
# quick GPU loop: parity tests + short bench lines (value leg with 1 / 2 / 4 streams)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -15
for s in ${STREAMS:-2}; do
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --streams $s 2>&1 | tail -1 > gpurun_out/quick_bench_s$s.json
python - <<PY
import json
d=json.loads(open('gpurun_out/quick_bench_s$s.json').read())
print('streams=$s value=%.0f e2e=%.0f ms/step=%.3f h2d=%.1fGB/s (bound %.0f pairs/s)'%(d['value'],d['e2e']['value'],d['ms_per_step'],d['e2e']['h2d_link_gbs'],d['e2e']['h2d_link_bound']))
print(' '.join('%s=%.2f'%(k,v) for k,v in d['roofline']['kernel_us_per_pair'].items()))
PY
done
