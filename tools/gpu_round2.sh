#!/bin/bash
# One gpurun call of round 2: GPU tests, smoke, the bench line + reference arm, the ncu launch list and one --set full
# capture of every kernel of one device-resident step (64 pairs, one stream).  Usage: tools/gpu_round2.sh <tag> [what]
#   what: all (default) | bench (skip tests) | prof (only the ncu captures)
set -u
TAG=${1:-r2}
WHAT=${2:-all}
mkdir -p gpurun_out
if [ "$WHAT" = all ]; then
  python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log
  python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${TAG}_smoke.log
fi
if [ "$WHAT" != prof ]; then
  python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/${TAG}_bench.err; cut -c1-1500 gpurun_out/${TAG}_bench.json
  python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2>> gpurun_out/${TAG}_bench.err; cut -c1-300 gpurun_out/${TAG}_bench_ref.json
fi
P="python bench.py --steps 2 --warmup 1 --pairs-per-step 64 --pool 64 --skip-e2e --no-cpu-baseline --no-configs --streams 1"
$P > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${TAG}_launches.csv $P > gpurun_out/${TAG}_ncu1.log 2>&1
echo "ncu list rc=$?"
# 16 of our launches per step (+ the one-off bind kernel): skip bind + two steps, capture one whole step
KR='regex:(resize|fast_score_win|fast_vote|octree|blur|describe|describe_tma|match_tc|stereo_filter|circular|bind)_kernel'
ncu --set full --clock-control none --import-source on -k "$KR" -s 33 -c 16 -f -o gpurun_out/${TAG}_prof $P > gpurun_out/${TAG}_ncu2.log 2>&1
echo "ncu full rc=$?"; tail -2 gpurun_out/${TAG}_ncu2.log
