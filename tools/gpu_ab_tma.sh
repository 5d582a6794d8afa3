#!/bin/bash
# A/B of the TMA-staged blur (experiment build) against the cp.async blur: parity tests, then per-kernel times.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-configs --skip-e2e 2>gpurun_out/ab_err.txt | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('product lib: value=%.0f' % d['value'], d['roofline']['kernel_us_per_pair'])"
export B2ORB_LIB=$PWD/pose-estimation_b200/libb2orb_variants.so
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "launch_time or stages or stereo_frontend" 2>&1 | tail -3
for v in 0 16 1 17; do
  python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-configs --skip-e2e --debug-set blur_band=$v 2>gpurun_out/ab_err.txt | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('blur_band=$v value=%.0f blur=%.3f us/pair' % (d['value'], d['roofline']['kernel_us_per_pair']['blur']))"
done
tools/ubench/pipes > gpurun_out/r2_pipes.txt 2>&1; tail -3 gpurun_out/r2_pipes.txt
