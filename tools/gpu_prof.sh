#!/bin/bash
# One gpurun call: plain run (must exit 0) then one ncu --set full capture of the kernels named in $1 (regex).
# Usage: tools/gpu_prof.sh <kernel-regex> <tag> [launch-skip] [launch-count]
set -u
K=${1:-fast_score_kernel}
TAG=${2:-prof}
SKIP=${3:-2}
CNT=${4:-2}
mkdir -p gpurun_out
P="python bench.py --steps 2 --warmup 1 --pairs-per-step 64 --pool 64 --skip-e2e --no-cpu-baseline --streams 1"
$P > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"$K" -s $SKIP -c $CNT -f -o gpurun_out/${TAG}_prof $P > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu full rc=$?"; tail -2 gpurun_out/${TAG}_ncu.log
