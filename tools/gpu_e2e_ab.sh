#!/bin/bash
# e2e leg A/B: pipeline chunk / lanes / host threads.  Usage: tools/gpu_e2e_ab.sh
for cfg in "64 3 2" "32 3 2" "128 3 2" "64 4 2" "64 3 1" "64 3 3" "32 4 3"; do
  set -- $cfg
  B2ORB_PIPE_CHUNK=$1 B2ORB_PIPE_LANES=$2 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --e2e-threads $3 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('chunk=$1 lanes=$2 threads=$3 value=%.0f e2e=%.0f link=%.1f GB/s'%(d['value'],d['e2e']['value'],d['e2e']['h2d_link_gbs']))"
done
