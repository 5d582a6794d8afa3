#!/bin/bash
# A/B loop over an environment switch: tools/gpu_ab.sh VAR "v1 v2 ..." [extra bench args]
# Runs the GPU parity tests once per value when PYTEST=1, then a short device-only bench line per value.
VAR=$1; VALS=$2; shift 2
mkdir -p gpurun_out
for v in $VALS; do
  export $VAR=$v
  if [ "${PYTEST:-0}" = "1" ]; then python -m pytest tests -m gpu -x -q 2>&1 | tail -2; fi
  python bench.py --steps 10 --warmup 3 --no-cpu-baseline --skip-e2e "$@" 2>/dev/null | tail -1 > gpurun_out/ab_${VAR}_$v.json
  python - <<PY
import json
d=json.loads(open('gpurun_out/ab_${VAR}_$v.json').read())
print('$VAR=$v value=%.0f ms/step=%.3f'%(d['value'],d['ms_per_step']), ' '.join('%s=%.2f'%(k,x) for k,x in d['roofline']['kernel_us_per_pair'].items()))
PY
done
